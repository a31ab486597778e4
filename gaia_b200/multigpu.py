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
        """Global colouring -> local colouring, colour order kept (every partition sweeps the same colour together).
        Inside a colour the vertices this rank SENDS sit behind the first 2/5 of the interior vertices: when their tiles come up
        the neighbours have long finished the previous colour (no stall on the epoch flag, which a boundary-first order paid at
        the start of every colour), and the peer stores and the system-scope fence still overlap with the remaining interior
        tiles instead of trailing the launch (a vertex update does not depend on the order inside a colour)."""
        boundary = np.zeros(len(self.local_verts), bool)
        for ids in self.send.values():
            boundary[ids] = True
        out = []
        for c in (np.asarray(c) for c in colors):
            l = self.g2l[c[self.g2l[c] >= 0]]
            inner, edge = l[~boundary[l]], l[boundary[l]]
            k = (2 * len(inner)) // 5
            out.append(np.concatenate([inner[:k], edge, inner[k:]]).astype(np.int32))
        return out

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
    # partitioning is the library's (csrc/partition.cpp, the same calls the C++ host's --gpus path makes); the numpy
    # `Partition` above states the same rules independently and is what the CPU tests compare it with
    owner = capi.partition_slab_owner(verts, world)
    part = capi.SlabPartition(len(verts), tets, owner, rank, world)
    part.verts = np.asarray(verts, np.float32)[part.local_verts]
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
    s.halo_set_world([e["handles"] for e in everyone])      # all ranks: agreement on "no contacts anywhere" (fused exchange)
    s.halo_finalize()
    dist.barrier()
    return s


class ReplicatedPartition:
    """Ownership record of the replicated layout (every rank holds the whole mesh): what `gather_owned_positions` needs."""

    def __init__(self, owner, rank):
        self.local_verts = np.arange(len(owner), dtype=np.int32)
        self.owned = np.asarray(owner) == rank


def add_replicated_mesh(scene, verts, tets, obj, colors, fixed, owner, rank, world):
    """The whole mesh on this rank, every vertex another rank owns as a ghost: solved nowhere but on its owner, overwritten
    by the owner's pushes. Colour order = the global one, so every rank numbers the vertices identically."""
    scene.set_partition_mode(rank, world)
    mesh = scene.add_tet_mesh(verts, tets, obj, global_color_order(colors), fixed)
    scene.set_ghosts(mesh, np.nonzero(np.asarray(owner) != rank)[0].astype(np.int32))
    return mesh


def build_replicated_scene(wl, rank, world, local_rank):
    """Cross-partition contacts for ONE mesh cut into slabs (SURVEY.md §8(e) "ghost BVH regions", taken to its simplest exact
    form): the SOLVE is partitioned - a rank keeps records only for the vertices of its slab - but every rank holds the whole
    mesh as collision geometry, and after every colour each rank pushes ALL its vertices of that colour to ALL ranks (not
    just a one-ring boundary to its neighbours). Positions are therefore identical everywhere at every point where the
    single-GPU step reads them, every rank finds every contact by itself (detection is deterministic), applies the ones
    that touch its own vertices - own contacts and face-side relations alike - and nothing but positions is ever exchanged.
    Bit-identical to the single-GPU run, far ends of a folded beam touching across ranks included. Cost: detection is not
    partitioned and the exchange moves a colour's worth of positions per step (6 MB per rank and colour on the 10 M-tet beam)."""
    import torch.distributed as dist
    from . import capi
    assert len(wl["meshes"]) == 1
    verts, tets, obj, colors, fixed = wl["meshes"][0]
    owner = capi.partition_slab_owner(wl.get("flat", verts), world)      # "flat": coordinates to cut along, if not the mesh's own
    s = capi.Scene(wl["phys"], wl["col"], device=local_rank)
    add_replicated_mesh(s, verts, tets, obj, colors, fixed, owner, rank, world)
    for d in wl["deformers"]:
        if d[0] == "rotator":
            s.add_rotator(0, d[1], d[2], d[3], d[4], d[5])
        else:
            s.add_translator(0, d[1], d[2], d[3])
    s.finalize()
    s.partition = ReplicatedPartition(owner, rank)
    mine = np.nonzero(owner == rank)[0].astype(np.int32)
    slots = s.sorted_ids(0, mine)                       # the same numbering on every rank: my slot IS the remote slot
    everyone = [None] * world
    dist.all_gather_object(everyone, s.halo_ipc_handles())
    mv = np.stack([np.zeros(len(mine), np.int32), mine], 1)
    for q in range(world):
        if q != rank:
            s.halo_add_peer(q, everyone[q], mv, slots)
    s.halo_set_world(everyone)
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


# ------------------------------------------------------------------------------------------------ many-object scenes
def global_groups(color_lists):
    """The merge of per-mesh colours into global parallel groups (M/VBD/VBDPhysics.cpp:236-274, findSmallestParallelGroup
    :99-122) restated on sizes alone, for ALL meshes of the global scene: returns, per mesh, its colours arranged by global
    group id (empty where the mesh has no colour in that group). A partition hands these to add_tet_mesh in partition mode,
    where colour i goes to group i, so every partition sweeps the same global group at the same time."""
    n_groups = max(len(c) for c in color_lists)
    sizes = [0] * n_groups
    out = []
    for colors in color_lists:
        ordered = global_color_order(colors)
        free = list(range(len(ordered)))
        arranged = [np.zeros(0, np.int32) for _ in range(n_groups)]
        for c in ordered:
            pick = 0
            for i in range(1, len(free)):
                if sizes[free[i]] < sizes[free[pick]]:
                    pick = i
            g = free.pop(pick)
            arranged[g] = np.asarray(c, np.int32)
            sizes[g] += len(c)
        out.append(arranged)
    return out


def scene_solver_lanes(color_lists, small=65536):
    """Lanes per vertex of the colour sweep (Scene.set_solver_lanes) from the GLOBAL scene: 8 when every global parallel
    group is small, else 4. Computed from all meshes of the scene, never from a rank's share, so that a sharded run and the
    single-GPU run of the same scene sum in the same association and stay bit-identical."""
    n_groups = max(len(c) for c in color_lists)
    sizes = [0] * n_groups
    for arranged in global_groups(color_lists):
        for g, c in enumerate(arranged):
            sizes[g] += len(c)
    return 8 if max(sizes) < small else 4


def object_boxes(meshes, transform):
    """[n,2,3] world-space boxes of every object at rest (transform(verts, obj) applies the object's placement)."""
    boxes = np.empty((len(meshes), 2, 3), np.float64)
    for k, m in enumerate(meshes):
        p = transform(m[0], m[2])
        boxes[k, 0] = p.min(0); boxes[k, 1] = p.max(0)
    return boxes


class ObjectSharding:
    """Whole objects assigned to ranks by equal-count slabs of their centroid along `axis`; each rank additionally holds
    GHOST copies (every vertex a ghost: never solved there, overwritten by the owner's per-colour pushes) of
      tier 1: objects of other ranks that may touch an owned object, and
      tier 2: objects that may touch a tier-1 object - they only make sure a tier-1 vertex picks the same single contact
              here as on its owner (the reference keeps ONE contact per vertex, rules R1-R3),
    "may touch" = the boxes, grown by `margin`, overlap in the coordinates `axes` (the vertical is left out for a falling
    pile: objects keep their column). The assignment is static and deterministic from the rest boxes: every rank derives
    every rank's tables without communication."""

    def __init__(self, boxes, world, margin, axis=0, axes=(0, 2)):
        n = len(boxes)
        self.world = world
        cen = 0.5 * (boxes[:, 0, axis] + boxes[:, 1, axis])
        order = np.argsort(cen, kind="stable")
        self.owner = np.empty(n, np.int32)
        for r in range(world):
            self.owner[order[(n * r) // world:(n * (r + 1)) // world]] = r
        lo = boxes[:, 0][:, list(axes)] - margin
        hi = boxes[:, 1][:, list(axes)] + margin
        touch = np.ones((n, n), bool)
        for d in range(lo.shape[1]):
            touch &= (lo[:, None, d] <= hi[None, :, d]) & (lo[None, :, d] <= hi[:, None, d])
        self.local, self.tier = [], []
        for r in range(world):
            owned = self.owner == r
            t1 = touch[owned].any(0) & ~owned
            t2 = touch[t1].any(0) & ~owned & ~t1 if t1.any() else np.zeros(n, bool)
            tier = np.full(n, -1, np.int32)
            tier[owned] = 0; tier[t1] = 1; tier[t2] = 2
            self.tier.append(tier)
            self.local.append(np.nonzero(tier >= 0)[0])          # ascending global id: local mesh order = global order

    def local_id(self, rank, k):
        return int(np.searchsorted(self.local[rank], k))

    def sends(self, rank):
        """{peer: [global object ids owned by `rank` that `peer` ghosts, ascending]}"""
        out = {}
        for q in range(self.world):
            if q == rank:
                continue
            ks = [int(k) for k in self.local[q] if self.owner[k] == rank]
            if ks:
                out[q] = ks
        return out

    def recvs(self, rank):
        """{peer: [global object ids owned by `peer` that `rank` ghosts, ascending]}"""
        out = {}
        for k in self.local[rank]:
            q = int(self.owner[k])
            if q != rank:
                out.setdefault(q, []).append(int(k))
        return out


def add_sharded_meshes(scene, meshes, sharding, rank):
    """Adds the local objects (owned + ghosts) of `rank` in global order; returns the list of global ids added."""
    arranged = global_groups([m[3] for m in meshes])
    scene.set_partition_mode(rank, sharding.world)
    for k in sharding.local[rank]:
        v, t, obj, _, fixed = meshes[k]
        mesh = scene.add_tet_mesh(v, t, obj, arranged[k], fixed)
        if sharding.owner[k] != rank:
            scene.set_ghosts(mesh, np.arange(len(v), dtype=np.int32))
    return sharding.local[rank]


def build_object_sharded_scene(wl, rank, world, local_rank, margin=None):
    """bench.py / tests: the many-object workload `wl` sharded by object over `world` GPUs.
    Requires an initialised torch.distributed process group (used at set-up only)."""
    import torch.distributed as dist
    from . import capi, meshio
    meshes = wl["meshes"]
    boxes = object_boxes(meshes, meshio.place)
    if margin is None:
        margin = 0.25 * float(np.median(boxes[:, 1] - boxes[:, 0]))
    sh = ObjectSharding(boxes, world, margin)
    s = capi.Scene(wl["phys"], wl["col"], device=local_rank)
    add_sharded_meshes(s, meshes, sh, rank)
    s.finalize()
    s.set_solver_lanes(scene_solver_lanes([m[3] for m in meshes]))      # from the global scene: same on every rank and on one GPU
    s.sharding = sh
    recv = sh.recvs(rank)
    mine = {"handles": s.halo_ipc_handles(),
            "ghost_slots": {q: np.concatenate([s.sorted_ids(sh.local_id(rank, k), np.arange(len(meshes[k][0]), dtype=np.int32)) for k in ks])
                            for q, ks in recv.items()}}
    everyone = [None] * world
    dist.all_gather_object(everyone, mine)
    for q, ks in sorted(sh.sends(rank).items()):
        mv = np.concatenate([np.stack([np.full(len(meshes[k][0]), sh.local_id(rank, k), np.int32), np.arange(len(meshes[k][0]), dtype=np.int32)], 1)
                             for k in ks])
        remote = everyone[q]["ghost_slots"][rank]
        assert len(remote) == len(mv), (rank, q, len(remote), len(mv))
        s.halo_add_peer(q, everyone[q]["handles"], mv, remote)
    s.halo_finalize()
    dist.barrier()
    return s
