#!/bin/bash
# N-GPU pass for cross-partition contacts (gpurun --gpus N): the folded beam (every contact joins the first and the last rank)
# and the twisting beam with detection on, both in the replicated layout, must equal the single-GPU run bit for bit;
# then the DCD-on scaling line in that layout.
N=${1:-2}; TAG=${2:-r2s}; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
F=$O/${TAG}_cross_contacts_${N}gpu.txt; : > $F
GAIA_CHECK_WORKLOAD=folded GAIA_CHECK_LAYOUT=replicated GAIA_CHECK_DIMS=16,2,2 GAIA_CHECK_FRAMES=8 timeout 600 $TR scripts/check_multigpu.py >> $F 2>&1; echo "rc=$?" >> $F
GAIA_CHECK_WORKLOAD=folded GAIA_CHECK_LAYOUT=replicated GAIA_CHECK_DIMS=96,8,8,0.05,0.03,0.6,2,10 GAIA_CHECK_FRAMES=8 timeout 600 $TR scripts/check_multigpu.py >> $F 2>&1; echo "rc=$?" >> $F
GAIA_CHECK_WORKLOAD=folded GAIA_CHECK_LAYOUT=ring GAIA_CHECK_DIMS=16,2,2 GAIA_CHECK_FRAMES=8 timeout 600 $TR scripts/check_multigpu.py >> $F 2>&1; echo "rc=$? (ring layout on the folded beam: expected to differ, it cannot see these contacts)" >> $F
GAIA_CHECK_DCD=1 GAIA_CHECK_LAYOUT=replicated GAIA_CHECK_DIMS=120,12,12 timeout 600 $TR scripts/check_multigpu.py >> $F 2>&1; echo "rc=$?" >> $F
grep -E "check_multigpu|rc=" $F
timeout 900 $TR bench.py --gpus $N --layout replicated --no-cpu-baseline > $O/${TAG}_scale_${N}gpu_replicated.json 2> $O/${TAG}_scale_${N}gpu_replicated.err; echo "rc=$?" >> $O/${TAG}_scale_${N}gpu_replicated.err
cut -c1-400 $O/${TAG}_scale_${N}gpu_replicated.json; tail -2 $O/${TAG}_scale_${N}gpu_replicated.err
