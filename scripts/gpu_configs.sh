#!/bin/bash
# bench lines of the other BASELINE configs (1, 2, 3, 5) with their cpu_baseline, one GPU; contact-rich states via pre-roll
O=gpurun_out; mkdir -p $O; TAG=${1:-r2}
for wl in model_drop_teapot s01 drop_pile_32 cloth_1m drop_pile_4096; do
  timeout 1500 python bench.py --workload $wl --steps 5 --warmup 3 > $O/${TAG}_bench_${wl}.json 2> $O/${TAG}_bench_${wl}.err; echo "rc=$?" >> $O/${TAG}_bench_${wl}.err
  echo "== $wl"; cut -c1-260 $O/${TAG}_bench_${wl}.json; python - <<PY
import json
try:
    d=json.load(open("$O/${TAG}_bench_${wl}.json"))
    r=d["roofline"]; print("  ms/step", round(d["ms_per_step"],2), "frac", round(r["frac"],3), "contacts", r["contacts"], "cpu", d["cpu_baseline"] and round(d["cpu_baseline"]["value"]/1e6,2))
except Exception as e: print("  (no line)", e)
PY
  tail -1 $O/${TAG}_bench_${wl}.err
done
