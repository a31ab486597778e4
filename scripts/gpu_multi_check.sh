#!/bin/bash
# N-GPU pass (gpurun --gpus N): bit-identity of the slab-partitioned beam and of the object-sharded pile against one GPU, then the scaling lines
N=${1:-2}; TAG=${2:-r2}; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR scripts/check_multigpu.py > $O/${TAG}_check_multigpu_${N}gpu.txt 2>&1; echo "rc=$?" >> $O/${TAG}_check_multigpu_${N}gpu.txt
grep -E "check_multigpu|rc=" $O/${TAG}_check_multigpu_${N}gpu.txt
timeout 600 $TR scripts/gpu_shard_check.py drop_pile_8 30 > $O/${TAG}_shard_check_pile8_${N}gpu.txt 2>&1; echo "rc=$?" >> $O/${TAG}_shard_check_pile8_${N}gpu.txt
grep -E "^rank|RESULT|rc=" $O/${TAG}_shard_check_pile8_${N}gpu.txt
timeout 900 $TR bench.py --gpus $N --workload twist_beam_10m_nodcd --no-cpu-baseline > $O/${TAG}_scale_${N}gpu_nodcd.json 2> $O/${TAG}_scale_${N}gpu_nodcd.err; echo "rc=$?" >> $O/${TAG}_scale_${N}gpu_nodcd.err
cat $O/${TAG}_scale_${N}gpu_nodcd.json | cut -c1-330; tail -2 $O/${TAG}_scale_${N}gpu_nodcd.err
timeout 900 $TR bench.py --gpus $N --no-cpu-baseline > $O/${TAG}_scale_${N}gpu.json 2> $O/${TAG}_scale_${N}gpu.err; echo "rc=$?" >> $O/${TAG}_scale_${N}gpu.err
cat $O/${TAG}_scale_${N}gpu.json | cut -c1-330; tail -2 $O/${TAG}_scale_${N}gpu.err
