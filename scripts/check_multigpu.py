#!/usr/bin/env python
"""Run under torchrun on N GPUs: the partitioned scene (per-colour halo exchange over NVLink peer stores) must
reproduce the single-GPU run of the same workload bit-exactly.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 scripts/check_multigpu.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from gaia_b200 import capi, multigpu  # noqa: E402
import bench  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dims = [float(x) if "." in x else int(x) for x in os.environ.get("GAIA_CHECK_DIMS", "40,6,6").split(",")]
    what = os.environ.get("GAIA_CHECK_WORKLOAD", "twist")
    layout = os.environ.get("GAIA_CHECK_LAYOUT", "ring")          # ring: one-ring ghosts; replicated: whole mesh on every rank
    frames = int(os.environ.get("GAIA_CHECK_FRAMES", "3"))
    if what == "folded":          # self-contact between the two ends of one mesh = contacts between the first and last rank
        from gaia_b200 import workloads
        wl = workloads.workload_folded_beam(*dims)
    else:
        wl = bench.workload_twist_beam(*dims, collisions=os.environ.get("GAIA_CHECK_DCD", "0") == "1")
    build = multigpu.build_replicated_scene if layout == "replicated" else multigpu.build_partitioned_scene
    s = build(wl, rank, world, local)
    if os.environ.get("GAIA_CHECK_MULTI_PART", "1") == "0":
        s.set_multi_colour_launch(False)
    s.run_frames(frames)
    s.synchronize()
    nV = len(wl["meshes"][0][0])
    got = multigpu.gather_owned_positions(s, s.partition, nV)
    st = s.stats()
    solves = torch.tensor([st["vertexSolves"]], dtype=torch.float64, device="cuda")
    dist.all_reduce(solves)
    ok = True
    if rank == 0:
        single = bench.build_scene(capi.Scene, wl, device=local)
        if os.environ.get("GAIA_CHECK_MULTI_SINGLE", "1") == "0":
            single.set_multi_colour_launch(False)
        single.run_frames(frames)
        ref = single.positions(0)
        moved = float(np.abs(ref - wl["meshes"][0][0]).max())
        same = np.array_equal(got, ref)
        con = single.contacts()
        print(f"[check_multigpu] {wl['name']} | layout={layout} world={world} frames={frames} moved={moved:.4f} bit_exact={same} "
              f"maxdiff={float(np.abs(got - ref).max()):.3e} solves {int(solves.item())} vs {single.stats()['vertexSolves']} "
              f"contacts(single, last detection)={len(con)} multi(part,single)={os.environ.get('GAIA_CHECK_MULTI_PART', '1')},{os.environ.get('GAIA_CHECK_MULTI_SINGLE', '1')}", flush=True)
        ok = same and moved > 1e-3 and int(solves.item()) == single.stats()["vertexSolves"]
        single.close()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    s.close()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
    main()
