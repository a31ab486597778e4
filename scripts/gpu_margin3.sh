#!/bin/bash
# adaptive lazy-refit margin: GPU suite, then the bench lines of the scenes it could help or hurt (no --refit-margin: the default)
O=gpurun_out; mkdir -p $O; TAG=${1:-r2za}
timeout 1500 python -m pytest tests -x -q -m gpu > $O/gpu_tests.log 2>&1; echo "pytest rc=$?" >> $O/gpu_tests.log; tail -4 $O/gpu_tests.log
for w in twist_beam_10m twist_beam_1m s01 drop_pile_32 drop_pile_4096; do
  timeout 900 python bench.py --no-cpu-baseline --workload $w > $O/${TAG}_adaptive_$w.json 2> $O/${TAG}_adaptive_$w.err; echo "rc=$?" >> $O/${TAG}_adaptive_$w.err
  python - <<PY
import json
try:
    d=json.load(open("$O/${TAG}_adaptive_$w.json")); r=d["roofline"]
    print("$w adaptive: ms/step", round(d["ms_per_step"],2), "G", round(d["value"]/1e9,3), "e2e", round(d["e2e"]["value"]/1e9,3), "contacts", r.get("contacts"))
except Exception as e: print("$w failed", e)
PY
done
