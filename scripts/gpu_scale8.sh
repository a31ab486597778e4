#!/bin/bash
# 8-GPU strong-scaling point of the default workload (10.2 M-tet twist beam), launched the way the driver does
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 8 --steps 5 --warmup 3 \
    > gpurun_out/scale_8gpu.json 2> gpurun_out/scale_8gpu.log; echo "rc=$?" >> gpurun_out/scale_8gpu.log
cat gpurun_out/scale_8gpu.json; tail -5 gpurun_out/scale_8gpu.log
