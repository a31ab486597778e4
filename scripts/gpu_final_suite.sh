#!/bin/bash
# final GPU check of the round: the new constraint tests first (not -x, so every failure shows), then the whole GPU suite
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_cloth_constraints.py tests/test_cloth_host.py -q -m gpu -p no:cacheprovider > gpurun_out/final_constraints.log 2>&1
echo "constraints rc=$?" >> gpurun_out/final_constraints.log
tail -25 gpurun_out/final_constraints.log
timeout 600 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/final_suite.log 2>&1
echo "suite rc=$?" >> gpurun_out/final_suite.log
tail -8 gpurun_out/final_suite.log
