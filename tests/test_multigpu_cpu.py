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


# ------------------------------------------------------------------------------ cross-partition contacts (replicated layout)
def _folded_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from gaia_b200 import workloads
    from oracle.binding import OracleScene
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    wl = workloads.workload_folded_beam()
    verts, tets, obj, colors, fixed = wl["meshes"][0]
    owner = multigpu.slab_owner(wl["flat"], world)          # slabs along the UNBENT beam: the two arms' ends are far-apart ranks
    o = OracleScene(wl["phys"], wl["col"], threads=1)
    multigpu.add_replicated_mesh(o, verts, tets, obj, colors, fixed, owner, rank, world)
    for d in wl["deformers"]:
        o.add_translator(0, d[1], d[2], d[3])
    o.finalize()
    groups = o.parallel_groups()
    mine = owner == rank

    def exchange(g):
        """every rank's owned vertices of colour g (all for g < 0) to every other rank"""
        pos = o.positions(0)
        ids = np.nonzero(mine)[0] if g < 0 else groups[g][:, 1][mine[groups[g][:, 1]]]
        everyone = [None] * world
        dist.all_gather_object(everyone, (ids, pos[ids]))
        for src, (i, p) in enumerate(everyone):
            if src != rank:
                pos[i] = p
        o.set_state(0, pos, None)

    phys = wl["phys"]
    cross = 0
    for frame in range(6):
        for sub in range(phys.numSubsteps):
            o.phase_begin_substep()
            exchange(-1)
            o.phase_prepare()
            for it in range(phys.iterations):
                for g in range(len(groups)):
                    o.phase_solve_group(g, it >= phys.frictionStartIter)
                    exchange(g)
                ici = phys.intermediateCollisionIterations
                if ici > 0 and it % ici == ici - 1:
                    o.phase_intermediate_cd()
            o.phase_end_substep()
    pos = o.positions(0)
    np.save(os.path.join(out_dir, f"fold_rank{rank}.npy"), np.concatenate([np.nonzero(mine)[0][:, None].astype(np.float64), pos[mine]], 1))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_cross_partition_contacts_replicated_layout_bit_exact(tmp_path, world):
    """A beam folded onto itself, cut into slabs along its length: every contact joins the first and the last rank. Solve
    partitioned, whole mesh replicated as collision geometry, a colour's owned vertices pushed to every rank
    (multigpu.build_replicated_scene is the device statement) - identical to the single run, contacts and friction included."""
    import torch.multiprocessing as mp
    from gaia_b200 import workloads
    from oracle.binding import OracleScene
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    mp.spawn(_folded_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    wl = workloads.workload_folded_beam()
    o = workloads.build_scene(OracleScene, wl, threads=1)
    o.run_frames(6)
    con = o.contacts()
    assert len(con) > 10
    owner = multigpu.slab_owner(wl["flat"], world)
    ref = o.positions(0)
    got = np.zeros_like(ref)
    seen = np.zeros(len(ref), bool)
    for r in range(world):
        a = np.load(tmp_path / f"fold_rank{r}.npy")
        ids = a[:, 0].astype(np.int64)
        got[ids] = a[:, 1:].astype(np.float32); seen[ids] = True
    assert seen.all()
    assert np.array_equal(got, ref)
    # and the contacts really cross the cut: query vertex and hit face belong to different ranks, none of them neighbours
    faces = o.surface_faces(0)
    q_owner = owner[con["queryVertex"]]
    f_owner = owner[faces[con["hitFace"]]]
    crossing = (f_owner != q_owner[:, None]).all(1)
    assert crossing.sum() > 10
    assert (np.abs(f_owner[crossing] - q_owner[crossing, None]).min(1) == world - 1).any()


# ------------------------------------------------------------------------------------------------ many-object sharding
def _pile_workload():
    """Six blocks in three columns (two stacked per column, columns overlapping sideways): contacts inside a column and
    across columns, so that with 2-3 ranks there are contacts between objects of different owners."""
    meshes = []
    shapes = [(3, 2, 2, 0.25), (2, 2, 3, 0.22), (3, 3, 2, 0.2)]
    for col_ in range(3):
        y = 0.25
        for lvl in range(2):
            nx, ny, nz, h = shapes[(col_ + lvl) % 3]
            v, t = meshio.kuhn_beam(nx, ny, nz, h=h, origin=(0.0, 0.0, 0.0))
            colors = meshio.kuhn_coloring(nx, ny, nz) if (col_ + lvl) % 2 == 0 else meshio.greedy_coloring(len(v), t)
            i = 2 * col_ + lvl
            obj = capi.default_object_params(miu=1e5 * (1 + 0.3 * i), lmbd=1e6, density=50 + 5 * i, frictionDynamic=0.2 + 0.05 * i,
                                             translation=[0.62 * col_ + 0.03 * lvl, y, 0.05 * lvl], rotation=[0.0, 0.1 * i, 0.03 * i],
                                             initialVelocity=[0.1 * (1 - col_), -2.0 - lvl, 0.05 * i])
            meshes.append((v, t, obj, colors, np.zeros(0, np.int32)))
            y += ny * h - 0.06
    phys = capi.default_physics_params(numSubsteps=2, iterations=5, worldBoundsMin=[-10, 0, -10], worldBoundsMax=[10, 20, 10],
                                       collisionStiffness=1e5)
    col = capi.default_collision_params(allowDCD=1, allowCCD=1)
    return dict(meshes=meshes, phys=phys, col=col)


def test_global_groups_restate_the_merge():
    from oracle.binding import OracleScene
    wl = _pile_workload()
    o = OracleScene(wl["phys"], wl["col"], threads=1)
    for v, t, obj, colors, fixed in wl["meshes"]:
        o.add_tet_mesh(v, t, obj, colors, fixed)
    o.finalize()
    groups = o.parallel_groups()
    arranged = multigpu.global_groups([m[3] for m in wl["meshes"]])
    assert all(len(a) == len(groups) for a in arranged)
    for g, pairs in enumerate(groups):
        want = np.concatenate([np.stack([np.full(len(a[g]), k, np.int32), a[g]], 1) for k, a in enumerate(arranged)])
        assert np.array_equal(pairs, want)


def test_solver_lanes_rule_uses_the_global_group_sizes():
    """scene_solver_lanes: one answer per GLOBAL scene (8 when every merged parallel group is below the bound), whatever a
    rank holds of it - that is what keeps a sharded run on the same summation association as the single-GPU run."""
    from oracle.binding import OracleScene
    wl = _pile_workload()
    colors = [m[3] for m in wl["meshes"]]
    o = OracleScene(wl["phys"], wl["col"], threads=1)
    for v, t, obj, c, fixed in wl["meshes"]:
        o.add_tet_mesh(v, t, obj, c, fixed)
    o.finalize()
    largest = max(len(g) for g in o.parallel_groups())
    assert multigpu.scene_solver_lanes(colors) == 8                          # a small pile
    assert multigpu.scene_solver_lanes(colors, small=largest + 1) == 8       # the bound is on the largest merged group
    assert multigpu.scene_solver_lanes(colors, small=largest) == 4
    assert multigpu.scene_solver_lanes(colors[:1], small=largest) == 8       # a rank's share alone would answer differently


def test_object_sharding_tables():
    wl = _pile_workload()
    boxes = multigpu.object_boxes(wl["meshes"], meshio.place)
    for world in (2, 3):
        sh = multigpu.ObjectSharding(boxes, world, margin=0.05)
        assert sorted(np.concatenate([np.nonzero(sh.owner == r)[0] for r in range(world)]).tolist()) == list(range(6))
        for r in range(world):
            assert (sh.tier[r][sh.owner == r] == 0).all()
            for q, ks in sh.sends(r).items():
                assert ks == sh.recvs(q)[r]
        # an object touching an owned object is always local
        sh0 = multigpu.ObjectSharding(boxes, world, margin=0.05)
        assert all(len(l) >= 2 for l in sh0.local)


def _pile_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from oracle.binding import OracleScene
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    wl = _pile_workload()
    meshes = wl["meshes"]
    sh = multigpu.ObjectSharding(multigpu.object_boxes(meshes, meshio.place), world, margin=0.05)
    o = OracleScene(wl["phys"], wl["col"], threads=1)
    multigpu.add_sharded_meshes(o, meshes, sh, rank)
    o.finalize()
    groups = o.parallel_groups()
    sends = sh.sends(rank)

    def exchange(g):
        msg = {}
        for q, ks in sends.items():
            items = []
            for k in ks:
                lm = sh.local_id(rank, k)
                pos = o.positions(lm)
                sel = np.arange(len(pos)) if g < 0 else groups[g][groups[g][:, 0] == lm][:, 1]
                items.append((k, sel, pos[sel]))
            msg[q] = items
        everyone = [None] * world
        dist.all_gather_object(everyone, msg)
        for src, m in enumerate(everyone):
            if src == rank or rank not in m:
                continue
            for k, sel, p in m[rank]:
                lm = sh.local_id(rank, k)
                pos = o.positions(lm)
                pos[sel] = p
                o.set_state(lm, pos, None)

    for frame in range(4):
        for sub in range(wl["phys"].numSubsteps):
            o.phase_begin_substep()
            exchange(-1)
            o.phase_prepare()
            for it in range(wl["phys"].iterations):
                for g in range(len(groups)):
                    o.phase_solve_group(g, True)
                    exchange(g)
            o.phase_end_substep()
    out = {int(k): o.positions(sh.local_id(rank, k)) for k in sh.local[rank] if sh.owner[k] == rank}
    np.save(os.path.join(out_dir, f"pile_rank{rank}.npy"), np.array([out], dtype=object), allow_pickle=True)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_object_sharded_oracle_equals_single_bit_exact(tmp_path, world):
    import torch.multiprocessing as mp
    from oracle.binding import OracleScene
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    mp.spawn(_pile_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    wl = _pile_workload()
    o = OracleScene(wl["phys"], wl["col"], threads=1)
    for v, t, obj, colors, fixed in wl["meshes"]:
        o.add_tet_mesh(v, t, obj, colors, fixed)
    o.finalize()
    o.run_frames(4)
    assert len(o.contacts()) > 0                       # the scene really is in contact
    seen = set()
    for r in range(world):
        d = np.load(tmp_path / f"pile_rank{r}.npy", allow_pickle=True)[0]
        for k, p in d.items():
            assert np.array_equal(p, o.positions(k)), (r, k, np.abs(p - o.positions(k)).max())
            seen.add(k)
    assert seen == set(range(6))


def test_cpp_partitioner_matches_the_python_one():
    """gaia_b200_partition_* (csrc/partition.cpp, what the C++ multi-GPU host and multigpu.py use) against the numpy statement
    in multigpu.Partition: owners, local meshes, owned masks, send / recv lists and restricted colourings identical."""
    from gaia_b200 import capi, meshio, multigpu
    for (nx, ny, nz), world in (((24, 4, 4), 2), ((40, 3, 5), 4), ((9, 7, 3), 3)):
        v, t = meshio.kuhn_beam(nx, ny, nz, h=0.1)
        colors = meshio.kuhn_coloring(nx, ny, nz)
        owner_py = multigpu.slab_owner(v, world)
        owner_c = capi.partition_slab_owner(v, world)
        assert np.array_equal(owner_py, owner_c)
        for rank in range(world):
            a = multigpu.Partition(v, t, owner_py, rank, world)
            b = capi.SlabPartition(len(v), t, owner_c, rank, world)
            assert np.array_equal(a.local_verts, b.local_verts) and np.array_equal(a.local_tets, b.local_tets)
            assert np.array_equal(a.tets, b.tets) and np.array_equal(a.owned, b.owned)
            assert sorted(a.send) == sorted(b.send) and sorted(a.recv) == sorted(b.recv)
            for q in a.send:
                assert np.array_equal(a.send[q], b.send[q])
            for q in a.recv:
                assert np.array_equal(a.recv[q], b.recv[q])
            ca, cb = a.restrict_colors(colors), b.restrict_colors(colors)
            assert len(ca) == len(cb) and all(np.array_equal(x, y) for x, y in zip(ca, cb))
            fixed = np.arange(0, len(v), 7, dtype=np.int32)
            assert np.array_equal(a.restrict_ids(fixed), b.restrict_ids(fixed))
            b.close()


def test_tet_surface_without_a_device_matches_the_oracle():
    from gaia_b200 import capi, meshio
    from gaia_b200.workloads import ref_mesh
    from oracle.binding import OracleScene
    for v, t in (meshio.kuhn_beam(5, 3, 2, h=0.2), ref_mesh("teapot")[:2]):
        ids, faces = capi.tet_surface(len(v), t)
        o = OracleScene(capi.default_physics_params(), capi.default_collision_params(allowDCD=0, allowCCD=0), threads=1)
        o.add_tet_mesh(v, t, capi.default_object_params(), [np.arange(len(v), dtype=np.int32)])
        o.finalize()
        assert np.array_equal(ids, o.surface_vertex_ids(0)) and np.array_equal(faces, o.surface_faces(0))
