"""Multi-GPU scenes: one process per GPU, each holding one partition of the scene.

The reference is single-process; this layer is new. A mesh is cut into slabs along its longest axis:
a partition OWNS the vertices of its slab and additionally holds every tet incident to an owned vertex, i.e. a one-ring
GHOST layer of vertices owned by neighbours. The VBD vertex update is purely local, so owned vertices see their complete
stencil and no force reduction is needed; after every colour sweep the new positions of boundary vertices of that colour
are written into the neighbours' ghost slots (libgaia_b200's halo exchange: peer stores over NVLink + device-side epoch
flags). `torch.distributed` is used for set-up only (exchange of IPC handles and ghost slot tables).

Everything that decides WHO owns WHAT is pure numpy and deterministic from the global mesh, so every rank derives the same
partition without communication; tests/test_multigpu_cpu.py runs the same logic on two gloo ranks with the CPU oracle.
"""
import numpy as np


def slab_owner(verts, world, axis=None):
    """Owner rank of every vertex: equal-count slabs along the longest axis; vertices sharing a coordinate stay together."""
    verts = np.asarray(verts)
    if axis is None:
        axis = int(np.argmax(verts.max(0) - verts.min(0)))
    x = verts[:, axis]
    order = np.argsort(x, kind="stable")
    owner = np.empty(len(verts), np.int32)
    bounds = [0]
    for r in range(1, world):
        cut = (len(verts) * r) // world
        # move the cut to the next change of coordinate so that a coordinate layer is never split
        while cut < len(verts) and cut > 0 and x[order[cut]] == x[order[cut - 1]]:
            cut += 1
        bounds.append(cut)
    bounds.append(len(verts))
    for r in range(world):
        owner[order[bounds[r]:bounds[r + 1]]] = r
    return owner


class Partition:
    """What rank `rank` holds of one global mesh.
    local_verts  global ids of the local vertices (ascending): owned first? no - ascending global id, owned and ghosts mixed
    local_tets   global ids of the local tets (ascending) ; tets [nT,4] in LOCAL vertex ids
    owned        bool mask over local vertices
    send[q]      local vertex ids owned here that are ghosts on rank q (ascending global id)
    recv[q]      local vertex ids of ghosts owned by rank q (ascending global id) — same global sequence as q's send[rank]
    """

    def __init__(self, verts, tets, owner, rank, world):
        tets = np.asarray(tets)
        own_t = owner[tets]                                   # [nT,4] owner of each tet vertex
        mine = (own_t == rank).any(1)
        self.rank, self.world = rank, world
        self.local_tets = np.nonzero(mine)[0]
        lt = tets[self.local_tets]
        self.local_verts = np.unique(lt)
        g2l = -np.ones(len(verts), np.int64)
        g2l[self.local_verts] = np.arange(len(self.local_verts))
        self.g2l = g2l
        self.tets = g2l[lt].astype(np.int32)
        self.verts = np.asarray(verts, np.float32)[self.local_verts]
        self.owned = owner[self.local_verts] == rank
        self.ghost_owner = owner[self.local_verts]
        self.send, self.recv = {}, {}
        for q in range(world):
            if q == rank:
                continue
            # vertices I own that appear in a tet with a vertex owned by q  == ghosts of q owned by me
            tq = (own_t == q).any(1) & mine
            if tq.any():
                vs = np.unique(tets[tq])
                mine_v = vs[owner[vs] == rank]
                if len(mine_v):
                    self.send[q] = g2l[mine_v].astype(np.int32)
            ghosts_q = self.local_verts[(self.ghost_owner == q)]
            if len(ghosts_q):
                self.recv[q] = g2l[ghosts_q].astype(np.int32)

    def restrict_colors(self, colors):
        """Global colouring -> local colouring, colour order kept (every partition sweeps the same colour together)."""
        return [self.g2l[c[self.g2l[c] >= 0]].astype(np.int32) for c in (np.asarray(c) for c in colors)]

    def restrict_ids(self, ids):
        ids = np.asarray(ids, np.int64)
        l = self.g2l[ids]
        return l[l >= 0].astype(np.int32)


def global_color_order(colors):
    """The order the single-GPU path sweeps colours in: size-descending, ties keep file order (stable) —
    M/TetMesh/TetMeshFEM.cpp:122-133 with the documented stable rule."""
    idx = sorted(range(len(colors)), key=lambda i: -len(colors[i]))
    return [colors[i] for i in idx]


def add_partition_mesh(scene, part, obj, colors, fixed, rank, world):
    scene.set_partition_mode(rank, world)
    mesh = scene.add_tet_mesh(part.verts, part.tets, obj, part.restrict_colors(global_color_order(colors)), part.restrict_ids(fixed))
    scene.set_ghosts(mesh, np.nonzero(~part.owned)[0].astype(np.int32))
    return mesh


def build_partitioned_scene(wl, rank, world, local_rank):
    """bench.py / tests: the workload `wl` (single mesh + deformers) partitioned over `world` GPUs; returns the Scene of
    this rank with its halo exchange set up. Requires an initialised torch.distributed process group."""
    import torch.distributed as dist
    from . import capi
    assert len(wl["meshes"]) == 1, "slab partitioning handles one mesh per scene (many-object scenes shard by object)"
    verts, tets, obj, colors, fixed = wl["meshes"][0]
    owner = slab_owner(verts, world)
    part = Partition(verts, tets, owner, rank, world)
    s = capi.Scene(wl["phys"], wl["col"], device=local_rank)
    add_partition_mesh(s, part, obj, colors, fixed, rank, world)
    for d in wl["deformers"]:
        ids = part.restrict_ids(d[1])
        if not len(ids):
            continue
        if d[0] == "rotator":
            s.add_rotator(0, ids, d[2], d[3], d[4], d[5])
        else:
            s.add_translator(0, ids, d[2], d[3])
    s.finalize()
    s.partition = part
    # set-up exchange: my IPC handles + the device slots of my ghosts, per owner
    mine = {"handles": s.halo_ipc_handles(), "ghost_slots": {q: s.sorted_ids(0, ids) for q, ids in part.recv.items()}}
    everyone = [None] * world
    dist.all_gather_object(everyone, mine)
    for q, ids in sorted(part.send.items()):
        remote = everyone[q]["ghost_slots"][rank]
        assert len(remote) == len(ids), (rank, q, len(remote), len(ids))
        mv = np.stack([np.zeros(len(ids), np.int32), ids], 1)
        s.halo_add_peer(q, everyone[q]["handles"], mv, remote)
    s.halo_finalize()
    dist.barrier()
    return s


def gather_owned_positions(scene, part, n_global):
    """All ranks' owned positions assembled in global vertex order on every rank (tests / output)."""
    import torch
    import torch.distributed as dist
    pos = scene.positions(0)
    mine = (part.local_verts[part.owned], pos[part.owned])
    everyone = [None] * dist.get_world_size()
    dist.all_gather_object(everyone, mine)
    out = np.zeros((n_global, 3), np.float32)
    for ids, p in everyone:
        out[ids] = p
    return out
