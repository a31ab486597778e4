#!/bin/bash
# N-GPU: bit-identity of the partitioned beam, then the two scaling lines (DCD on = the default bench, and detection off)
N=${1:-2}; TAG=${2:-r2w}; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
F=$O/${TAG}_checks_${N}gpu.txt; : > $F
GAIA_CHECK_DCD=1 GAIA_CHECK_DIMS=120,12,12 timeout 600 $TR scripts/check_multigpu.py >> $F 2>&1; echo "rc=$?" >> $F
timeout 600 $TR scripts/check_multigpu.py >> $F 2>&1; echo "rc=$?" >> $F
grep -E "check_multigpu\]|rc=" $F | sed 's/.*layout=/layout=/'
timeout 900 $TR bench.py --gpus $N --no-cpu-baseline > $O/${TAG}_scale_${N}gpu.json 2> $O/${TAG}_scale_${N}gpu.err; echo "rc=$?" >> $O/${TAG}_scale_${N}gpu.err
cut -c1-200 $O/${TAG}_scale_${N}gpu.json; tail -1 $O/${TAG}_scale_${N}gpu.err
timeout 900 $TR bench.py --gpus $N --workload twist_beam_10m_nodcd --no-cpu-baseline > $O/${TAG}_scale_${N}gpu_nodcd.json 2> $O/${TAG}_scale_${N}gpu_nodcd.err; echo "rc=$?" >> $O/${TAG}_scale_${N}gpu_nodcd.err
cut -c1-200 $O/${TAG}_scale_${N}gpu_nodcd.json; tail -1 $O/${TAG}_scale_${N}gpu_nodcd.err
