#!/bin/bash
# default bench (DCD on) at several lazy-refit margins, then the other configs' bench lines
O=gpurun_out; mkdir -p $O; TAG=${1:-r2y}
for m in 0.2 0.3 0.45 0.7; do
  timeout 600 python bench.py --no-cpu-baseline --refit-margin $m > $O/${TAG}_margin_$m.json 2> $O/${TAG}_margin_$m.err; echo "rc=$?" >> $O/${TAG}_margin_$m.err
  python - <<PY
import json
d=json.load(open("$O/${TAG}_margin_$m.json")); r=d["roofline"]
print("margin $m: ms/step", round(d["ms_per_step"],2), "G", round(d["value"]/1e9,3), "phases", {k:round(v,2) for k,v in r["phases_ms_per_step"].items()})
PY
done
bash scripts/gpu_configs.sh $TAG
