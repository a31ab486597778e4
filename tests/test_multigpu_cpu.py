"""CPU, world_size 2 (gloo): the multi-GPU host logic — slab partitioning, ghost layers, send/recv tables, per-colour
halo protocol — driven with the CPU oracle as the per-partition solver. A partitioned run must reproduce the
single-partition run BIT-EXACTLY: every owned vertex sees the same stencil, the same colour order and the same neighbour
positions."""
import os
import socket

import numpy as np
import pytest

from gaia_b200 import capi, meshio, multigpu


def _workload():
    nx, ny, nz = 10, 2, 2
    verts, tets = meshio.kuhn_beam(nx, ny, nz, h=0.2, origin=(0, 0.5, 0))
    colors = meshio.greedy_coloring(len(verts), tets)       # unequal colour sizes on purpose
    fixed = np.nonzero(verts[:, 0] < 1e-6)[0].astype(np.int32)
    phys = capi.default_physics_params(numSubsteps=2, iterations=4, worldBoundsMin=[-10, 0, -10], worldBoundsMax=[10, 20, 10])
    col = capi.default_collision_params(allowDCD=0, allowCCD=0)
    obj = capi.default_object_params(miu=5e4, lmbd=5e5, density=100, dampingHydrostatic=1e-5, dampingDeviatoric=1e-5)
    right = np.nonzero(verts[:, 0] > nx * 0.2 - 1e-6)[0].astype(np.int32)
    return verts, tets, colors, fixed, phys, col, obj, right


def test_partition_tables_are_consistent():
    verts, tets, colors, fixed, *_ = _workload()
    for world in (2, 3, 4):
        owner = multigpu.slab_owner(verts, world)
        assert set(owner.tolist()) == set(range(world))
        parts = [multigpu.Partition(verts, tets, owner, r, world) for r in range(world)]
        # every vertex owned exactly once; every tet held by each owner of one of its vertices
        cnt = np.zeros(len(verts), int)
        for p in parts:
            cnt[p.local_verts[p.owned]] += 1
        assert (cnt == 1).all()
        for r, p in enumerate(parts):
            # full stencil: all tets incident to an owned vertex are local
            inc = np.isin(tets, p.local_verts[p.owned]).any(1)
            assert np.array_equal(np.nonzero(inc)[0], p.local_tets)
            for q, ids in p.send.items():
                # what r sends to q is exactly what q expects from r, in the same global order
                assert np.array_equal(p.local_verts[ids], parts[q].local_verts[parts[q].recv[r]])
            assert set(p.recv) == {q for q in range(world) if q != r and (p.ghost_owner == q).any()}
            lc = p.restrict_colors(multigpu.global_color_order(colors))
            assert sum(len(c) for c in lc) == len(p.local_verts)
            assert meshio.check_coloring(len(p.local_verts), p.tets, lc)


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from oracle.binding import OracleScene
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    verts, tets, colors, fixed, phys, col, obj, right = _workload()
    owner = multigpu.slab_owner(verts, world)
    part = multigpu.Partition(verts, tets, owner, rank, world)
    o = OracleScene(phys, col, threads=1)
    multigpu.add_partition_mesh(o, part, obj, colors, fixed, rank, world)
    ids = part.restrict_ids(right)
    if len(ids):
        o.add_translator(0, ids, [0.3, 0.0, 0.1])
    o.finalize()
    groups = o.parallel_groups()
    group_of = np.zeros(len(part.local_verts), np.int64)
    for g, pairs in enumerate(groups):
        group_of[pairs[:, 1]] = g

    def exchange(g):
        """push my boundary vertices of colour g (all colours for g < 0) to their ghosts on the neighbours"""
        pos = o.positions(0)
        msg = {}
        for q, sid in part.send.items():
            sel = sid if g < 0 else sid[group_of[sid] == g]
            msg[q] = (part.local_verts[sel], pos[sel])
        everyone = [None] * world
        dist.all_gather_object(everyone, msg)
        for src, m in enumerate(everyone):
            if src == rank or rank not in m:
                continue
            gids, p = m[rank]
            pos[part.g2l[gids]] = p
        o.set_state(0, pos, None)

    for frame in range(3):
        for sub in range(phys.numSubsteps):
            o.phase_begin_substep()
            exchange(-1)
            o.phase_prepare()
            for it in range(phys.iterations):
                for g in range(len(groups)):
                    o.phase_solve_group(g, True)
                    exchange(g)
            o.phase_end_substep()
    pos = o.positions(0)
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), np.concatenate([part.local_verts[part.owned][:, None].astype(np.float64), pos[part.owned]], 1))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_partitioned_oracle_equals_single_bit_exact(tmp_path, world):
    import torch.multiprocessing as mp
    from oracle.binding import OracleScene
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    verts, tets, colors, fixed, phys, col, obj, right = _workload()
    o = OracleScene(phys, col, threads=1)
    o.add_tet_mesh(verts, tets, obj, colors, fixed)
    o.add_translator(0, right, [0.3, 0.0, 0.1])
    o.finalize()
    o.run_frames(3)
    ref = o.positions(0)
    got = np.zeros_like(ref)
    seen = np.zeros(len(ref), bool)
    for r in range(world):
        a = np.load(tmp_path / f"rank{r}.npy")
        ids = a[:, 0].astype(np.int64)
        got[ids] = a[:, 1:].astype(np.float32); seen[ids] = True
    assert seen.all()
    assert np.abs(ref - verts).max() > 1e-3          # something moved
    assert np.array_equal(got, ref)
