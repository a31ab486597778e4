#!/bin/bash
# One GPU-box pass: parity tests, bench (1M and 10M-tet beams), ncu launch list and one full capture of the colour solve.
# Usage (from the repo root, on the GPU box via gpurun): bash scripts/gpu_check.sh [tag]
TAG=${1:-r1}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest.log
timeout 600 python bench.py --workload twist_beam_1m --steps 5 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_1m.json 2> $O/${TAG}_bench_1m.err; echo "rc=$?" >> $O/${TAG}_bench_1m.err
timeout 1200 python bench.py --steps 5 --warmup 3 > $O/${TAG}_bench_10m.json 2> $O/${TAG}_bench_10m.err; echo "rc=$?" >> $O/${TAG}_bench_10m.err
timeout 600 python bench.py --workload drop_pile_32 --steps 5 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_pile32.json 2> $O/${TAG}_bench_pile32.err; echo "rc=$?" >> $O/${TAG}_bench_pile32.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches_10m.csv \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline --profile-steps 1 > $O/${TAG}_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solve_colour --launch-skip 40 -c 4 -o $O/${TAG}_solve_full -f \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline --profile-steps 1 > $O/${TAG}_ncu_full.log 2>&1
ls -la $O
