"""torchrun --nproc-per-node N scripts/gpu_shard_check.py [workload] [frames]: the object-sharded multi-GPU run of a
many-object scene against the single-GPU run of the same scene (every rank also runs the whole scene on its own GPU):
owned objects must come out bit-identical."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import bench
from gaia_b200 import capi, multigpu

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
torch.cuda.set_device(lr)
name = sys.argv[1] if len(sys.argv) > 1 else "drop_pile_8"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 30
wl = bench.WORKLOADS[name]()
s = multigpu.build_object_sharded_scene(wl, rank, world, lr)
sh = s.sharding
ref = bench.build_scene(capi.Scene, wl, device=lr)
t0 = time.time()
for f in range(frames):
    s.run_frames(1)
s.synchronize()
t1 = time.time()
ref.run_frames(frames); ref.synchronize()
worst, nOwned = 0.0, 0
for k in sh.local[rank]:
    if sh.owner[k] != rank:
        continue
    d = np.abs(s.positions(sh.local_id(rank, k)) - ref.positions(int(k))).max()
    worst = max(worst, float(d)); nOwned += 1
st, sr = s.stats(), ref.stats()
print(f"rank {rank}: {nOwned} owned + {len(sh.local[rank]) - nOwned} ghost objects of {len(wl['meshes'])}; {frames} frames in {t1 - t0:.2f}s; "
      f"max |x_sharded - x_single| over owned objects = {worst:.3e}; contacts dcd/ccd sharded {st['numContactsDCD']}/{st['numContactsCCD']} "
      f"single {sr['numContactsDCD']}/{sr['numContactsCCD']}", flush=True)
t = torch.tensor([worst], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print("RESULT bit-identical" if t.item() == 0.0 else f"RESULT differs: {t.item():.3e}", flush=True)
dist.barrier(); dist.destroy_process_group()
