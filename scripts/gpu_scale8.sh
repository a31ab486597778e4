#!/bin/bash
# 8-GPU pass: bit-identity of the slab-partitioned beam (torchrun path) and of the C++ host (--gpus 8), then the strong-scaling lines
O=gpurun_out; mkdir -p $O; TAG=${1:-r2}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --master-port 29511"
GAIA_CHECK_DIMS=96,8,8 timeout 600 $TR --nproc-per-node 8 scripts/check_multigpu.py > $O/${TAG}_check_multigpu_8gpu.txt 2>&1; echo "rc=$?" >> $O/${TAG}_check_multigpu_8gpu.txt
grep -E "check_multigpu|rc=" $O/${TAG}_check_multigpu_8gpu.txt
timeout 600 python scripts/check_cpp_multigpu.py 4 4 > $O/${TAG}_check_cpp_multigpu_4gpu.txt 2>&1; tail -2 $O/${TAG}_check_cpp_multigpu_4gpu.txt
for N in 2 4 8; do
  timeout 900 $TR --nproc-per-node $N bench.py --gpus $N --workload twist_beam_10m_nodcd --no-cpu-baseline > $O/${TAG}_scale_${N}gpu_nodcd.json 2> $O/${TAG}_scale_${N}gpu_nodcd.err; echo "rc=$?" >> $O/${TAG}_scale_${N}gpu_nodcd.err
  cut -c1-230 $O/${TAG}_scale_${N}gpu_nodcd.json
  timeout 900 $TR --nproc-per-node $N bench.py --gpus $N --no-cpu-baseline > $O/${TAG}_scale_${N}gpu.json 2> $O/${TAG}_scale_${N}gpu.err; echo "rc=$?" >> $O/${TAG}_scale_${N}gpu.err
  cut -c1-230 $O/${TAG}_scale_${N}gpu.json
done
timeout 900 $TR --nproc-per-node 8 bench.py --gpus 8 --workload drop_pile_4096 --no-cpu-baseline > $O/${TAG}_pile4096_8gpu.json 2> $O/${TAG}_pile4096_8gpu.err; echo "rc=$?" >> $O/${TAG}_pile4096_8gpu.err
cut -c1-230 $O/${TAG}_pile4096_8gpu.json
timeout 600 $TR --nproc-per-node 4 scripts/gpu_shard_check.py drop_pile_32 30 > $O/${TAG}_shard_check_pile32_4gpu.txt 2>&1; grep -E "^rank|RESULT" $O/${TAG}_shard_check_pile32_4gpu.txt | cut -c1-200
