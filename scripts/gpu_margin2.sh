#!/bin/bash
O=gpurun_out; mkdir -p $O; TAG=${1:-r2z}
run() { # workload margin
  timeout 900 python bench.py --no-cpu-baseline --workload $1 --refit-margin $2 > $O/${TAG}_m_$1_$2.json 2> $O/${TAG}_m_$1_$2.err; echo "rc=$?" >> $O/${TAG}_m_$1_$2.err
  python - <<PY
import json
try:
    d=json.load(open("$O/${TAG}_m_$1_$2.json")); r=d["roofline"]
    print("$1 margin $2: ms/step", round(d["ms_per_step"],2), "contacts", r.get("contacts"))
except Exception as e: print("$1 $2 failed", e)
PY
}
for m in 0.38 0.52 0.6; do run twist_beam_10m $m; done
for w in s01 drop_pile_32 drop_pile_4096 twist_beam_1m; do for m in 0.2 0.45; do run $w $m; done; done
