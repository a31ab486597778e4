#!/bin/bash
# round-2 validation: all GPU tests, smoke, the default bench (config 4 as SURVEY §8(d) writes it: DCD on) and its detection-off twin
O=gpurun_out; mkdir -p $O; TAG=${1:-r2g}
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_gpu_tests.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_gpu_tests.log
tail -5 $O/${TAG}_gpu_tests.log
timeout 300 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> $O/${TAG}_smoke.log; tail -3 $O/${TAG}_smoke.log
timeout 900 python bench.py > $O/${TAG}_bench_default.json 2> $O/${TAG}_bench_default.err; echo "rc=$?" >> $O/${TAG}_bench_default.err
cat $O/${TAG}_bench_default.json; tail -2 $O/${TAG}_bench_default.err
timeout 900 python bench.py --workload twist_beam_10m_nodcd --no-cpu-baseline > $O/${TAG}_bench_nodcd.json 2> $O/${TAG}_bench_nodcd.err; echo "rc=$?" >> $O/${TAG}_bench_nodcd.err
cat $O/${TAG}_bench_nodcd.json
