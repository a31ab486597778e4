#!/bin/bash
# tile sweep iteration: quick parity subset, default bench, one ncu capture of the sweep
O=gpurun_out; mkdir -p $O; TAG=${1:-r2c}
timeout 600 python -m pytest tests/test_gpu_solve_parity.py tests/test_gpu_ref_meshes.py tests/test_gpu_collision_parity.py -m gpu -x -q -k "not config2" > $O/${TAG}_tests.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_tests.log
tail -4 $O/${TAG}_tests.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "rc=$?" >> $O/${TAG}_bench.err
cat $O/${TAG}_bench.json; tail -2 $O/${TAG}_bench.err
timeout 800 ncu --set full --clock-control none --import-source on -k regex:k_sweep_tiles --launch-skip 40 -c 1 -o $O/${TAG}_sweep_full -f python bench.py --steps 1 --warmup 0 --no-cpu-baseline --profile-steps 1 > $O/${TAG}_ncu_full.log 2>&1; tail -2 $O/${TAG}_ncu_full.log
