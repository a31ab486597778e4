#!/bin/bash
# small scenes: graph of per-colour launches (mode 1) against multi-colour launches everywhere (mode 2)
O=gpurun_out; mkdir -p $O; TAG=${1:-r2ze}
for w in model_drop_teapot twist_beam_small drop_pile_32 twist_beam_1m_nodcd; do for m in 1 2; do
  timeout 600 python bench.py --no-cpu-baseline --workload $w --multi-colour-launch $m > $O/${TAG}_$w_m$m.json 2> $O/${TAG}_$w_m$m.err; echo "rc=$?" >> $O/${TAG}_$w_m$m.err
  python - <<PY
import json
try:
    d=json.load(open("$O/${TAG}_$w_m$m.json"))
    print("$w mode $m: ms/step", round(d["ms_per_step"],3), "launches", d["gpu_launches"])
except Exception as e: print("$w $m failed", e)
PY
done; done
