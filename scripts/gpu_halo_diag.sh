#!/bin/bash
# where does the per-colour exchange cost go? (timing experiments: results of debug runs are wrong on purpose)
N=${1:-4}
for dbg in 0 1 2 3; do
  echo -n "N=$N HALO_DEBUG=$dbg: "
  GAIA_B200_HALO_DEBUG=$dbg timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+dbg)) \
      bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],3), 'ms/frame', round(d['value']/1e9,2), 'G')"
done
echo -n "N=$N unfused push: "
GAIA_B200_HALO_FUSED=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29610 \
      bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],3), 'ms/frame', round(d['value']/1e9,2), 'G')"
