#!/bin/bash
# quick look at the default bench line (no CPU baseline) and the collision parity tests
O=gpurun_out; mkdir -p $O; TAG=${1:-quick}
timeout 600 python -m pytest tests/test_gpu_collision_parity.py tests/test_gpu_ref_meshes.py -x -q -m gpu > $O/${TAG}_col_tests.log 2>&1; echo "rc=$?" >> $O/${TAG}_col_tests.log; tail -3 $O/${TAG}_col_tests.log
timeout 600 python bench.py --no-cpu-baseline > $O/${TAG}_default.json 2> $O/${TAG}_default.err; echo "rc=$?"
python - <<PY
import json
d=json.load(open("$O/${TAG}_default.json")); r=d["roofline"]
print("default: ms/step", round(d["ms_per_step"],2), "G", round(d["value"]/1e9,3), "e2e", round(d["e2e"]["value"]/1e9,3), "launch us", round(r["avg_launch_us"],2), "share", round(r["kernel_share_of_step"],3), "dram_frac", round(r.get("dram_frac",0),3))
PY
