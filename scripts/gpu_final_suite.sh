#!/bin/bash
# final GPU check of the round: the whole GPU suite in one pass (no -x: every failure shows), log kept under gpurun_out/
mkdir -p gpurun_out
timeout 300 python -m pytest tests -q -m gpu -rf -p no:cacheprovider > gpurun_out/final_suite.log 2>&1
echo "suite rc=$?" >> gpurun_out/final_suite.log
tail -15 gpurun_out/final_suite.log
