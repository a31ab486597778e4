#!/bin/bash
# what the driver runs at round end: GPU tests, smoke, default bench (ours + reference arm)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gpu_tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.log; echo "bench rc=$?" >> gpurun_out/bench_default.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.log; echo "ref rc=$?" >> gpurun_out/bench_reference.log
tail -3 gpurun_out/gpu_tests.log; tail -2 gpurun_out/smoke.log; cat gpurun_out/bench_default.json; cat gpurun_out/bench_reference.json; tail -n 2 gpurun_out/bench_default.log; tail -n 2 gpurun_out/bench_reference.log
