#!/bin/bash
# one gpurun call: GPU tests, smoke, cloth bench, launch list + full capture of the cloth colour solve
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gpu_tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py --workload cloth_1m --steps 3 --warmup 3 > gpurun_out/bench_cloth1m.json 2> gpurun_out/bench_cloth1m.log; echo "bench rc=$?" >> gpurun_out/bench_cloth1m.log
if [ -s gpurun_out/bench_cloth1m.json ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 5000 --csv --log-file gpurun_out/launches_cloth250k.csv \
      python bench.py --workload cloth_250k --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solve_colour --launch-skip 2000 --launch-count 4 \
      -o gpurun_out/cloth_solve_full -f python bench.py --workload cloth_1m --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
fi
tail -5 gpurun_out/gpu_tests.log; tail -3 gpurun_out/smoke.log; cat gpurun_out/bench_cloth1m.json; tail -5 gpurun_out/bench_cloth1m.log
