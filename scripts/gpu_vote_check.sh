#!/bin/bash
# N-GPU pass for the "fused exchange while nobody holds a contact" rule: bit-identity with detection on (ring and replicated
# layouts, contact-free and contact-rich), then the DCD-on scaling lines.
N=${1:-2}; TAG=${2:-r2t}; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
F=$O/${TAG}_vote_checks_${N}gpu.txt; : > $F
GAIA_CHECK_DCD=1 GAIA_CHECK_LAYOUT=ring GAIA_CHECK_DIMS=120,12,12 timeout 600 $TR scripts/check_multigpu.py >> $F 2>&1; echo "rc=$?" >> $F
GAIA_CHECK_DCD=1 GAIA_CHECK_LAYOUT=replicated GAIA_CHECK_DIMS=120,12,12 timeout 600 $TR scripts/check_multigpu.py >> $F 2>&1; echo "rc=$?" >> $F
GAIA_CHECK_WORKLOAD=folded GAIA_CHECK_LAYOUT=replicated GAIA_CHECK_DIMS=96,8,8,0.05,0.03,0.6,2,10 GAIA_CHECK_FRAMES=8 timeout 600 $TR scripts/check_multigpu.py >> $F 2>&1; echo "rc=$?" >> $F
timeout 600 $TR scripts/check_multigpu.py >> $F 2>&1; echo "rc=$?" >> $F
timeout 600 $TR scripts/gpu_shard_check.py drop_pile_8 30 > $O/${TAG}_shard_check_pile8_${N}gpu.txt 2>&1; echo "rc=$?" >> $O/${TAG}_shard_check_pile8_${N}gpu.txt
grep -E "check_multigpu|rc=" $F; grep -E "RESULT|rc=" $O/${TAG}_shard_check_pile8_${N}gpu.txt
for L in ring replicated; do
  timeout 900 $TR bench.py --gpus $N --layout $L --no-cpu-baseline > $O/${TAG}_scale_${N}gpu_${L}.json 2> $O/${TAG}_scale_${N}gpu_${L}.err; echo "rc=$?" >> $O/${TAG}_scale_${N}gpu_${L}.err
  cut -c1-330 $O/${TAG}_scale_${N}gpu_${L}.json; tail -2 $O/${TAG}_scale_${N}gpu_${L}.err
done
timeout 900 $TR bench.py --gpus $N --workload twist_beam_10m_nodcd --no-cpu-baseline > $O/${TAG}_scale_${N}gpu_nodcd.json 2> $O/${TAG}_scale_${N}gpu_nodcd.err; echo "rc=$?" >> $O/${TAG}_scale_${N}gpu_nodcd.err
cut -c1-330 $O/${TAG}_scale_${N}gpu_nodcd.json; tail -2 $O/${TAG}_scale_${N}gpu_nodcd.err
