#!/bin/bash
# the multi-colour launch: its bit-identity test first (short timeout: it spins on a grid barrier), then the whole GPU suite and the bench lines
mkdir -p gpurun_out; O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_solve_parity.py -m gpu -q -x -k "multi_colour" > $O/multi_test.log 2>&1; rc=$?; echo "rc=$rc" >> $O/multi_test.log; tail -15 $O/multi_test.log
if [ $rc -ne 0 ]; then exit 1; fi
timeout 1500 python -m pytest tests -m gpu -q > $O/gpu_tests.log 2>&1; echo "pytest rc=$?" >> $O/gpu_tests.log; tail -5 $O/gpu_tests.log
timeout 600 python bench.py --no-cpu-baseline > $O/ml_bench_default.json 2> $O/ml_bench_default.log; echo "rc=$?" >> $O/ml_bench_default.log; cat $O/ml_bench_default.json; tail -2 $O/ml_bench_default.log
timeout 600 python bench.py --no-cpu-baseline --workload twist_beam_10m_nodcd > $O/ml_bench_nodcd.json 2> $O/ml_bench_nodcd.log; echo "rc=$?" >> $O/ml_bench_nodcd.log; cat $O/ml_bench_nodcd.json; tail -2 $O/ml_bench_nodcd.log
