#!/bin/bash
# quick 2-GPU probe: which side loses bit-identity - the partitioned run or the single run, with / without the multi-colour launch
N=${1:-2}; TAG=${2:-probe}; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
F=$O/${TAG}_halo_probe_${N}gpu.txt; : > $F
for P in 0 1; do for S in 0 1; do
GAIA_CHECK_MULTI_PART=$P GAIA_CHECK_MULTI_SINGLE=$S timeout 300 $TR scripts/check_multigpu.py >> $F 2>&1; echo "rc=$?" >> $F
done; done
grep -E "check_multigpu|rc=" $F
