"""Seeded inputs shared by tests/golden/make_golden.py (which feeds them to the reference's compiled code) and
tests/test_oracle_vs_ref.py (which replays them through the CPU oracle)."""
import numpy as np

from gaia_b200 import capi, meshio


def single_tet_cases(n=16, seed=12, dampH=2e-4, dampD=3e-4):
    """n one-tet meshes (every vertex has exactly one incident tet), for the damped Neo-Hookean comparison: the
    reference's GPU-variant damping branch scales the ACCUMULATED Hessian, so only one-incident-tet vertices have a
    closed-form relation to the CPU semantics."""
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        verts = (np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], np.float32) * 0.3 + rng.normal(0, 0.04, (4, 3))).astype(np.float32)
        tets = np.array([[0, 1, 2, 3]], np.int32)
        out.append(neohookean_case(seed + 100 + i, dampH, dampD, mesh=(verts, tets, [np.array([k], np.int32) for k in range(4)])))
    return out


def neohookean_case(seed, dampH, dampD, mesh=None):
    """A small Kuhn beam, randomly deformed (5 % of the cell size) with a random previous position; everything the
    reference's per-vertex Neo-Hookean routine reads, produced by the oracle's own setup (TetMeshFEM::initialize)."""
    from oracle.binding import OracleScene
    rng = np.random.default_rng(seed)
    if mesh is None:
        verts, tets = meshio.kuhn_beam(3, 2, 2, h=0.3, origin=(0, 1, 0))
        colors = meshio.kuhn_coloring(3, 2, 2)
    else:
        verts, tets, colors = mesh
    obj = capi.default_object_params(miu=1e5, lmbd=1e6, density=50, dampingHydrostatic=dampH, dampingDeviatoric=dampD,
                                     rotation=[0.2, 0.1, -0.3], scale=[1.1, 0.9, 1.0])
    phys = capi.default_physics_params(numSubsteps=4)
    o = OracleScene(phys, capi.default_collision_params(allowDCD=0, allowCCD=0), threads=1)
    o.add_tet_mesh(verts, tets, obj, colors)
    o.finalize()
    rest = o.positions(0)
    pos = (rest + rng.normal(0, 0.015, rest.shape)).astype(np.float32)
    prev = (pos + rng.normal(0, 0.004, rest.shape)).astype(np.float32)
    infos, nbrTets, nbrOrder = o.vertex_tet_csr(0)
    return dict(pos=np.ascontiguousarray(pos), prev=np.ascontiguousarray(prev), tets=np.ascontiguousarray(tets, np.int32),
                DmInv=o.dminv(0), restVol=o.rest_volumes(0), infos=infos, nbrTets=nbrTets, nbrOrder=nbrOrder,
                miu=1e5, lmbd=1e6, dampH=dampH, dampD=dampD, dt=float(np.float32(phys.timeStep) / np.float32(4)),
                verts=verts, colors=colors, obj=obj, phys=phys, scene=o)


def solve3x3_cases(n=64, seed=5):
    rng = np.random.default_rng(seed)
    Ms = np.zeros((n, 9), np.float32); Bs = rng.normal(0, 10, (n, 3)).astype(np.float32)
    for i in range(n):
        A = rng.normal(0, 1, (3, 3))
        S = A @ A.T * 10 ** rng.uniform(-2, 6)
        if i % 8 == 0:                       # rank deficient -> the failure rule
            S = np.outer(A[:, 0], A[:, 0]) * 1e3
        if i % 16 == 3:
            S = np.zeros((3, 3))
        Ms[i] = S.astype(np.float32).reshape(-1, order="F")
    return Ms, Bs


def ccd_cases(n=200, seed=9):
    """Random triangle/point trajectories: about half of them cross."""
    rng = np.random.default_rng(seed)
    x0 = rng.uniform(-1, 1, (n, 3, 3)).astype(np.float32)
    x1 = (x0 + rng.normal(0, 0.3, (n, 3, 3))).astype(np.float32)
    c0 = x0.mean(1); nrm = np.cross(x0[:, 1] - x0[:, 0], x0[:, 2] - x0[:, 0])
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    w = rng.dirichlet([1, 1, 1], n)
    inside = np.einsum("nk,nkd->nd", w, x0)
    start = np.where(rng.random((n, 1)) < 0.7, inside, c0 + rng.normal(0, 1.5, (n, 3)))
    p0 = (start + nrm * rng.uniform(0.05, 0.6, (n, 1))).astype(np.float32)
    p1 = (start - nrm * rng.uniform(-0.2, 0.8, (n, 1)) + rng.normal(0, 0.05, (n, 3))).astype(np.float32)
    return (np.ascontiguousarray(x0.reshape(n, 9)), np.ascontiguousarray(x1.reshape(n, 9)), np.ascontiguousarray(p0),
            np.ascontiguousarray(p1))


def cubic_cases(n=200, seed=3):
    rng = np.random.default_rng(seed)
    coefs = np.zeros((n, 4), np.float32)
    for i in range(n):
        k = i % 4
        if k == 0:      # three roots, some inside [0,1]
            r = rng.uniform(-0.5, 1.5, 3)
            c = np.poly(r)[::-1] * rng.uniform(0.5, 4)
        elif k == 1:    # one real root
            r = rng.uniform(-0.5, 1.5)
            c = np.polymul([1, -r], [1, rng.uniform(-1, 1), rng.uniform(1, 3)])[::-1]
        elif k == 2:
            c = rng.normal(0, 1, 4)
        else:           # nearly quadratic
            c = np.append(rng.normal(0, 1, 3), rng.normal(0, 1e-3))
        coefs[i] = c.astype(np.float32)
    return coefs


def contact_cases(n=60, seed=21):
    rng = np.random.default_rng(seed)
    tri = rng.uniform(-1, 1, (n, 3, 3)).astype(np.float32)
    triPrev = (tri + rng.normal(0, 0.01, tri.shape)).astype(np.float32)
    nrm = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]); nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    w = rng.dirichlet([1, 1, 1], n)
    w[n // 2:] = rng.uniform(-0.4, 1.2, (n - n // 2, 3))     # second half: closest point on an edge or a vertex
    foot = np.einsum("nk,nkd->nd", w, tri)
    side = np.where(np.arange(n) % 5 == 4, 1.0, -1.0)[:, None]   # every 5th case lies outside (no force)
    p = (foot + side * nrm * rng.uniform(0.01, 0.2, (n, 1))).astype(np.float32)
    pPrev = (p + rng.normal(0, 0.02, p.shape)).astype(np.float32)
    return dict(p=np.ascontiguousarray(p), pPrev=np.ascontiguousarray(pPrev), tri=np.ascontiguousarray(tri.reshape(n, 9)),
                triPrev=np.ascontiguousarray(triPrev.reshape(n, 9)), k=1e5, mu0=0.2, mu1=0.4, eps0=0.01, eps1=0.03, dt=1.0 / 240)


def boundary_cases(n=48, seed=33):
    rng = np.random.default_rng(seed)
    mn = np.array([-2.0, 0.0, -3.0], np.float32); mx = np.array([2.0, 5.0, 3.0], np.float32)
    p = rng.uniform(mn + 0.1, mx - 0.1, (n, 3)).astype(np.float32)
    for i in range(n):
        if i % 6 == 5:
            continue                         # inside: no force
        d = i % 3
        p[i, d] = (mn[d] - rng.uniform(1e-3, 0.2)) if (i // 3) % 2 == 0 else (mx[d] + rng.uniform(1e-3, 0.2))
    scale = np.where(np.arange(n)[:, None] % 4 == 0, 1e-4, 0.03)   # displacements below and above epsU = epsV*dt
    pPrev = (p + rng.normal(0, 1, p.shape) * scale).astype(np.float32)
    return dict(p=np.ascontiguousarray(p), pPrev=np.ascontiguousarray(pPrev), mn=mn, mx=mx, kB=1e5, muB=0.3, epsVB=1.0, dt=1.0 / 240)


def bowl_cases(n=160, seed=35):
    """points around the bowl ground (the inside of a sphere): inside (no force), outside below / above the centre (above only
    counts with the cap), just outside, far outside; displacements below and above epsU = epsV * dt"""
    rng = np.random.default_rng(seed)
    center = np.array([0.3, 1.0, -0.2], np.float32); radius = 2.0
    d = rng.normal(0, 1, (n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = np.where(np.arange(n) % 4 == 0, rng.uniform(0.5, 1.99, n), np.where(np.arange(n) % 4 == 1, rng.uniform(2.0001, 2.01, n), rng.uniform(2.01, 2.6, n)))
    p = (center + d * r[:, None]).astype(np.float32)
    scale = np.where(np.arange(n)[:, None] % 3 == 0, 1e-4, 0.03)
    pPrev = (p + rng.normal(0, 1, p.shape) * scale).astype(np.float32)
    return dict(p=np.ascontiguousarray(p), pPrev=np.ascontiguousarray(pPrev), center=center, radius=radius, kB=1e5, muB=0.3, epsVB=1.0, dt=1.0 / 240)


def point_in_tet_cases(n=400, seed=41):
    """tets of mixed size (some nearly flat: |6V| around the reference's 1e-5 rejection bound) with query points inside,
    outside, on faces / edges / vertices and just off them"""
    rng = np.random.RandomState(seed)
    tets = rng.uniform(-1, 1, (n, 4, 3)).astype(np.float32)
    tets[::7] *= 0.02                                         # small tets: volume near CMP_EPSILON
    tets[3::11, 3] = tets[3::11, 0] + 0.3 * (tets[3::11, 1] - tets[3::11, 0]) + 0.3 * (tets[3::11, 2] - tets[3::11, 0]) + \
        rng.uniform(-2e-5, 2e-5, (len(tets[3::11]), 3)).astype(np.float32)     # nearly coplanar
    w = rng.dirichlet(np.ones(4), n).astype(np.float32)
    kind = rng.randint(0, 6, n)
    w[kind == 1] = rng.uniform(-0.3, 1.3, (int((kind == 1).sum()), 4)).astype(np.float32)      # mostly outside
    w[kind == 2, 0] = 0.0                                                                       # on a face
    w[kind == 3, :2] = 0.0                                                                      # on an edge
    w[kind == 4] = np.eye(4, dtype=np.float32)[rng.randint(0, 4, int((kind == 4).sum()))]       # at a vertex
    w /= np.where(np.abs(w.sum(1, keepdims=True)) > 1e-6, w.sum(1, keepdims=True), 1.0)
    p = np.einsum("nk,nkd->nd", w, tets).astype(np.float32)
    p[kind == 5] += rng.uniform(-1e-6, 1e-6, (int((kind == 5).sum()), 3)).astype(np.float32)
    return np.ascontiguousarray(p), np.ascontiguousarray(tets)


def closest_point_triangle_cases(n=400, seed=43):
    """random triangles (some slivers) and query points over all seven Voronoi regions (vertex / edge / interior)"""
    rng = np.random.RandomState(seed)
    tri = rng.uniform(-1, 1, (n, 3, 3)).astype(np.float32)
    tri[::9, 2] = tri[::9, 0] + 0.5 * (tri[::9, 1] - tri[::9, 0]) + rng.uniform(-1e-3, 1e-3, (len(tri[::9]), 3)).astype(np.float32)
    w = rng.uniform(-0.8, 1.8, (n, 3)).astype(np.float32)
    w /= w.sum(1, keepdims=True) + 1e-3
    nrm = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    p = np.einsum("nk,nkd->nd", w, tri) + rng.uniform(-0.5, 0.5, (n, 1)) * nrm
    p[::13] = tri[::13, 1]                                    # exactly at a vertex
    return np.ascontiguousarray(p, np.float32), np.ascontiguousarray(tri)


def segment_cases(n=400, seed=47):
    """segment pairs: general position, parallel, sharing an end point, degenerate (zero length)"""
    rng = np.random.RandomState(seed)
    s = rng.uniform(-1, 1, (n, 4, 3)).astype(np.float32)      # p0, p1, q0, q1
    s[::5, 3] = s[::5, 2] + (s[::5, 1] - s[::5, 0])           # parallel
    s[1::17, 2] = s[1::17, 0]                                 # shared end point
    s[2::19, 1] = s[2::19, 0]                                 # zero-length first segment
    s[4::23, 3] = s[4::23, 2]                                 # zero-length second segment
    return np.ascontiguousarray(s)


def coloring_graphs():
    """name -> (numNodes, unique undirected edges [nE,2] with a < b): seeded random graphs and the vertex graphs of two
    reference meshes (tests/golden/ref_root), as M/GraphColoring/TetMeshVertexGraph.h builds them"""
    import os
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    from gaia_b200 import meshio
    from gaia_b200.workloads import ref_mesh
    g = {}
    rng = np.random.RandomState(77)
    for n, deg in ((60, 4), (300, 7), (1000, 12)):
        e = rng.randint(0, n, (n * deg // 2, 2))
        e = e[e[:, 0] != e[:, 1]]
        e = np.unique(np.sort(e, axis=1), axis=0)
        g[f"rand{n}"] = (n, e.astype(np.int32))
    for m in ("teapot", "beam_1k"):
        v, t, _ = ref_mesh(m)
        offs, dst = meshio.vertex_adjacency(len(v), t)
        src = np.repeat(np.arange(len(v)), np.diff(offs))
        keep = src < dst
        g[m] = (len(v), np.stack([src[keep], dst[keep]], 1).astype(np.int32))
    return g


def cloth_repulsion_cases(n=300, seed=51):
    """(dis, thickness, contactRadius, contactStiffness) rows for the three contact energy types: distances on both sides of
    tau = r/2 and of the CMP_EPSILON cut, and beyond the radius"""
    rng = np.random.RandomState(seed)
    r = rng.uniform(0.05, 0.5, n).astype(np.float32)
    th = (rng.uniform(0, 0.02, n) * (rng.rand(n) < 0.7)).astype(np.float32)
    dis = (th + r * rng.uniform(-0.05, 1.3, n)).astype(np.float32)
    dis[::11] = th[::11] + np.float32(1e-5) * rng.uniform(0.5, 1.5, len(dis[::11])).astype(np.float32)
    dis[1::13] = th[1::13] + r[1::13] * 0.5
    k = rng.uniform(1e3, 1e6, n).astype(np.float32)
    return dis, th, r, k


def cloth_material_case(n=7, seed=61):
    """A small square cloth, bent and stretched at random (with a random previous position, so that both damping terms act):
    everything the reference's per-vertex cloth material routine reads. Mesh set-up (DmInv, rest areas, edge infos) is the
    cloth oracle's; the vertex -> face and vertex -> hinge tables are rebuilt here in ascending order."""
    from gaia_b200 import cloth
    from oracle.binding_cloth import OracleClothScene
    rng = np.random.default_rng(seed)
    verts, faces = cloth.grid_cloth(n, size=1.0, origin=(0.0, 1.0, 0.0))
    phys = cloth.default_cloth_physics_params(numSubsteps=4, iterations=1, handleCollision=0)
    obj = cloth.default_cloth_object_params(miu=1.2e4, lambda_=2.1e4, bendingStiffness=6.0, dampingStVK=1.5e-3, dampingBending=2.5e-3, density=0.3)
    o = OracleClothScene(phys, threads=1)
    o.add_tri_mesh(verts, faces, obj, cloth.grid_cloth_coloring(n))
    o.finalize()
    rest = o.positions(0)
    h = 1.0 / (n - 1)
    pos = (rest + rng.normal(0, 0.08 * h, rest.shape)).astype(np.float32)
    prev = (pos + rng.normal(0, 0.02 * h, rest.shape)).astype(np.float32)
    o.set_state(0, pos, None); o.set_prev_positions(0, prev)
    faces = np.ascontiguousarray(faces, np.int32)
    edges = o.edges(0)
    nV = len(rest)
    vf = [[] for _ in range(nV)]
    for f, tri in enumerate(faces):
        for k in range(3):
            vf[tri[k]].append((f, k))
    vb = [[] for _ in range(nV)]
    for e, E in enumerate(edges):
        if E[5] < 0:
            continue                                    # boundary edge: no hinge
        for k in range(4):
            vb[E[k]].append((e, k))

    def csr(lists):
        off = np.zeros(nV + 1, np.int32); a = []; b = []
        for v, l in enumerate(lists):
            off[v + 1] = off[v] + len(l); a += [x[0] for x in l]; b += [x[1] for x in l]
        return off, np.asarray(a, np.int32), np.asarray(b, np.int32)
    dminv, area = o.dminv(0)
    return dict(scene=o, rest=np.ascontiguousarray(rest), pos=pos, prev=prev, faces=faces, edges=np.ascontiguousarray(edges, np.int32),
                dminv=np.ascontiguousarray(dminv), area=np.ascontiguousarray(area), vf=csr(vf), vb=csr(vb),
                miu=1.2e4, lmbd=2.1e4, bend=6.0, dampS=1.5e-3, dampB=2.5e-3, dt=float(np.float32(phys.timeStep) / np.float32(4)),
                thres=float(phys.degenerateTriangleThres))


def cloth_contact_case(energy_type=1, seed=71):
    """Two small cloths a few centimetres apart (different friction parameters), both wrinkled at random with random previous
    positions, inside a world box that cuts through them: vertex-face pairs in both directions, edge-edge pairs, and vertices
    outside the box on several axes - the inputs of the reference's per-contact and boundary routines."""
    from gaia_b200 import cloth
    from oracle.binding_cloth import OracleClothScene
    rng = np.random.default_rng(seed)
    phys = cloth.default_cloth_physics_params(numSubsteps=4, iterations=1, handleCollision=1, contactRadius=0.12, thickness=0.01,
                                              contactStiffness=2e4, contactEnergyTypes=energy_type, boundaryContactStiffness=3e4,
                                              boundaryFrictionDynamic=0.4, boundaryFrictionEpsValue=0.05,
                                              worldBoundsMin=[0.2, 0.005, -10.0], worldBoundsMax=[0.9, 10.0, 0.7])
    o = OracleClothScene(phys, threads=1)
    meshes = []
    for n, size, org, mu, eps in ((5, 1.0, (0.0, 0.0, 0.0), 0.3, 0.05), (4, 0.8, (0.1, 0.06, 0.12), 0.5, 0.2)):
        v, f = cloth.grid_cloth(n, size=size, origin=org)
        obj = cloth.default_cloth_object_params(frictionDynamic=mu, frictionEpsV=eps, bendingStiffness=1.0)
        o.add_tri_mesh(v, f, obj, cloth.grid_cloth_coloring(n))
        meshes.append([np.ascontiguousarray(f, np.int32), mu, eps])
    o.finalize()
    out = []
    for i, (f, mu, eps) in enumerate(meshes):
        rest = o.positions(i)
        pos = (rest + rng.normal(0, 0.02, rest.shape)).astype(np.float32)
        prev = (pos + rng.normal(0, 0.01, rest.shape)).astype(np.float32)
        o.set_state(i, pos, None); o.set_prev_positions(i, prev)
        out.append(dict(pos=np.ascontiguousarray(pos), prev=np.ascontiguousarray(prev), faces=f, edges=np.ascontiguousarray(o.edges(i), np.int32), mu=mu, eps=eps))
    vf = []           # (vertex mesh, vertex, face mesh, face): the face with the nearest centroid, both directions
    for a, b in ((0, 1), (1, 0)):
        cen = out[b]["pos"][out[b]["faces"]].mean(1)
        for v in range(len(out[a]["pos"])):
            vf.append((a, v, b, int(np.argmin(((cen - out[a]["pos"][v]) ** 2).sum(1)))))
    ee = []           # (mesh 1, edge, mesh 2, edge): nearest midpoints
    mid = [m["pos"][m["edges"][:, :2]].mean(1) for m in out]
    for e1 in range(len(out[0]["edges"])):
        ee.append((0, e1, 1, int(np.argmin(((mid[1] - mid[0][e1]) ** 2).sum(1)))))
    dt = float(np.float32(phys.timeStep) / np.float32(4))
    params8 = np.array([2e4, 0.12, 0.01, dt, 3e4, 0.4, 0.05, float(energy_type)], np.float32)
    return dict(scene=o, meshes=out, vf=vf, ee=ee, params8=params8, bmin=np.array([0.2, 0.005, -10.0], np.float32), bmax=np.array([0.9, 10.0, 0.7], np.float32))


def cloth_constraint_case(seed=91):
    """Sewing and vertex-face attachments: two simulated 5 x 5 cloths and a collider sheet, all wrinkled at
    random. Seams inside cloth 0 with ratios that include both ends (0 and 1) and a vertex that sits in three seams;
    attachments of every kind the reference's set-up loop distinguishes: to a fixed point, to a face of the other simulated
    cloth (one barycentric coordinate exactly 0: that face vertex gets no reaction entry), to a face of the collider, and to
    a face of the vertex's own cloth; with and without per-constraint stiffness (None = the struct's default 1e4)."""
    from gaia_b200 import cloth
    rng = np.random.default_rng(seed)
    meshes = []
    for n, org in ((5, (0.0, 0.5, 0.0)), (5, (0.05, 0.56, 0.02)), (4, (0.0, 0.4, 0.0))):
        v, f = cloth.grid_cloth(n, 1.0, origin=org)
        meshes.append(((v + rng.normal(0, 0.02, v.shape)).astype(np.float32), np.ascontiguousarray(f, np.int32), n))
    seams = np.array([[0, 6, 7], [12, 1, 2], [12, 20, 21], [24, 12, 13], [3, 18, 19], [9, 15, 16]], np.int32)
    ratios = np.array([0.3, 0.0, 1.0, 0.65, 0.5, 0.125], np.float32)
    # (mesh, vertex), (target mesh or -1, three face vertices), barycentrics or position
    v_infos = np.array([[0, 4], [0, 10], [1, 7], [0, 22], [1, 0], [1, 7]], np.int32)
    f_infos = np.array([[-1, 0, 0, 0], [1, 5, 6, 10], [2, 0, 1, 4], [0, 8, 13, 14], [0, 1, 2, 6], [-1, 0, 0, 0]], np.int32)
    barys = np.array([[0.1, 0.9, -0.2], [0.25, 0.0, 0.75], [0.2, 0.3, 0.5], [0.6, 0.3, 0.1], [1.0 / 3, 1.0 / 3, 1.0 / 3], [0.3, 0.55, 0.1]], np.float32)
    stiff = np.array([1e4, 3e3, 5e4, 8e3, 2e4, 1.5e3], np.float32)
    return dict(meshes=meshes, n_sim=2, seams=seams, ratios=ratios, sew_k=7.5e3, v_infos=v_infos, f_infos=f_infos, barys=barys, stiff=stiff)


def deformer_cases():
    """A Kuhn beam whose two end faces are driven: (kind, selected vertices, centre, axis or speed, angular velocity, end time). The
    rotator's axis is not a unit vector and does not pass through the centre on purpose (the reference normalises nothing); end
    times fall inside the run."""
    verts, tets = meshio.kuhn_beam(4, 2, 2, h=0.25, origin=(0.1, 0.5, -0.2))
    left = np.nonzero(verts[:, 0] < 0.11)[0].astype(np.int32); right = np.nonzero(verts[:, 0] > 1.09)[0].astype(np.int32)
    return dict(verts=verts, tets=tets, colors=meshio.kuhn_coloring(4, 2, 2),
                cases=[(0, left, [0.0, 0.7, 0.1], [1.0, 0.05, -0.02], 2.5, -1.0), (0, right, [1.1, 0.75, 0.0], [0.9, 0.0, 0.1], -4.0, 0.05),
                       (1, right, [0.0, 0.0, 0.0], [0.3, -0.1, 0.05], 0.0, -1.0), (1, left, [0.0, 0.0, 0.0], [-0.2, 0.4, 0.0], 0.0, 0.06)])


def run_oracle_deformer(case, c, frames=6):
    """the oracle with one deformer: (curTime after each frame, positions of the selected vertices after each frame)"""
    from oracle.binding import OracleScene
    kind, sel, centre, a, w, end = case
    phys = capi.default_physics_params(numSubsteps=3, iterations=2, worldBoundsMin=[-10, -10, -10], worldBoundsMax=[10, 10, 10])
    o = OracleScene(phys, capi.default_collision_params(allowDCD=0, allowCCD=0), threads=1)
    o.add_tet_mesh(c["verts"], c["tets"], capi.default_object_params(miu=1e4, lmbd=1e5, density=100), c["colors"])
    if kind == 0:
        o.add_rotator(0, sel, centre, a, w, end)
    else:
        o.add_translator(0, sel, a, end)
    o.finalize()
    times, pos = [], []
    for _ in range(frames):
        o.run_frames(1)
        times.append(o.cur_time); pos.append(o.positions(0)[sel].copy())
    return np.array(times, np.float64), np.array(pos, np.float32), o


def cloth_deformer_cases():
    """a 6 x 6 cloth whose two edges are driven: (kind, selected vertices, centre, axis or speed, angular velocity, start, end)"""
    from gaia_b200 import cloth
    v, f = cloth.grid_cloth(6, 1.0, origin=(0.0, 1.0, 0.0))
    left = np.arange(0, 6, dtype=np.int32); right = np.arange(30, 36, dtype=np.int32)
    return dict(verts=v, faces=f, cases=[(0, left, [0.0, 1.0, 0.1], [0.0, 0.05, 1.0], 3.0, 0.0, -1.0), (0, right, [0.2, 1.1, 0.0], [0.1, 0.0, 0.9], -5.0, 0.0, 0.04),
                                         (1, right, [0, 0, 0], [0.3, -0.2, 0.05], 0.0, 0.0, -1.0), (1, left, [0, 0, 0], [-0.2, 0.4, 0.0], 0.0, 0.02, 0.06)])


def run_cloth_oracle_deformer(case, c, frames=5, substeps=4):
    """the cloth oracle with one deformer: (curTime of every substep, dt, positions of the selected vertices after each frame)"""
    from gaia_b200 import cloth
    from oracle.binding_cloth import OracleClothScene
    kind, sel, centre, a, w, t0, t1 = case
    phys = cloth.default_cloth_physics_params(numSubsteps=substeps, iterations=2, handleCollision=0)
    o = OracleClothScene(phys, threads=1)
    o.add_tri_mesh(c["verts"], c["faces"], cloth.default_cloth_object_params(density=0.2, miu=1e3), cloth.grid_cloth_coloring(6))
    if kind == 0:
        o.add_rotator(0, sel, centre, a, w, t1)
    else:
        o.add_translator(0, sel, a, t0, t1)
    o.finalize()
    dt = np.float32(phys.timeStep) / np.float32(substeps)
    t = np.float32(0.0)
    times, pos = [], []
    for _ in range(frames):
        o.run_frames(1)
        for _ in range(substeps):
            t = np.float32(t + dt); times.append(t)
        assert abs(o.cur_time - float(t)) < 1e-9
        pos.append(o.positions(0)[sel].copy())
    return np.array(times, np.float32), float(dt), np.array(pos, np.float32)


def collider_sequence_cases(seed=151):
    """A collider with four key meshes at frames 1, 3, 4, 8 and the calls one step makes - every (frame 0..9, substep 0..2, iteration
    0..3) in order - for three settings: per-iteration interpolation over the first 2 iterations, over the first iteration only,
    and once per frame"""
    from gaia_b200 import cloth
    rng = np.random.default_rng(seed)
    v, f = cloth.grid_cloth(4, 0.8, origin=(0.0, 0.3, 0.0))
    keys = np.stack([v + rng.normal(0, 0.05, v.shape).astype(np.float32) + np.float32([0, 0.1 * k, 0]) for k in range(4)]).astype(np.float32)
    calls = np.array([(fr, sub, it) for fr in range(10) for sub in range(3) for it in range(4)], np.int32)
    return dict(verts=v, faces=f, keys=np.ascontiguousarray(keys), key_frames=np.array([1, 3, 4, 8], np.int32), calls=calls, substeps=3, iters=4,
                settings=[(1, 2), (1, 1), (0, 1)])


def run_cloth_oracle_collider(c, interp_substep, interp_iter):
    """the cloth oracle's update of the collider at every call: (positions [nCalls, nV, 3], updated flags)"""
    from gaia_b200 import cloth
    from oracle.binding_cloth import OracleClothScene
    phys = cloth.default_cloth_physics_params(numSubsteps=c["substeps"], iterations=c["iters"], handleCollision=0)
    o = OracleClothScene(phys, threads=1)
    sv, sf = cloth.grid_cloth(3, 0.5, origin=(0.0, 1.0, 0.0))
    o.add_tri_mesh(sv, sf, cloth.default_cloth_object_params(), cloth.grid_cloth_coloring(3))
    cid = o.add_tri_mesh(c["keys"][0], c["faces"], cloth.default_cloth_object_params(), simulated=False)
    o.set_collider_sequence(cid, c["keys"], c["key_frames"], bool(interp_substep), interp_iter)
    o.finalize()
    fn = o._lib.oracle_cloth_probe_update_colliders
    pos, upd = [], []
    for fr, sub, it in c["calls"]:
        upd.append(int(fn(o.h, int(fr), int(sub), int(it))))
        pos.append(o.positions(cid).copy())
    return np.array(pos, np.float32), np.array(upd, np.int32)


def group_merge_case(seed=161):
    """Five meshes with different numbers of colour classes of different sizes (Kuhn colourings, greedy colourings, and colourings
    with some classes merged) for the merge into global parallel groups: (meshes [(verts, tets, colours)], the colours as the oracle
    holds them = sorted by size, large first, ties in file order)"""
    rng = np.random.default_rng(seed)
    meshes = []
    for i, dims in enumerate(((3, 2, 2), (2, 2, 1), (4, 1, 1), (1, 1, 1), (2, 3, 2))):
        verts, tets = meshio.kuhn_beam(*dims, h=0.2, origin=(0.5 * i, 0.1, 0.0))
        colors = meshio.kuhn_coloring(*dims) if i % 2 == 0 else meshio.greedy_coloring(len(verts), tets)
        colors = [np.asarray(c, np.int32) for c in colors if len(c)]
        if i == 4:      # fewer, larger classes would not be a proper colouring; split instead: one class in two
            k = int(np.argmax([len(c) for c in colors])); c = colors.pop(k); colors += [c[: len(c) // 3], c[len(c) // 3:]]
        rng.shuffle(colors)
        meshes.append((verts, tets, colors))
    return meshes


def cloth_group_merge_case(seed=163):
    """four cloths with 9, 6, 9 and 4 colour classes of uneven sizes (grid colourings with classes merged or emptied, in shuffled
    order): the cloth framework takes the classes in file order and does not retire a group once a mesh has used it"""
    from gaia_b200 import cloth
    rng = np.random.default_rng(seed)
    meshes = []
    for i, n in enumerate((7, 5, 6, 4)):
        v, f = cloth.grid_cloth(n, 1.0, origin=(0.0, 1.0 + 0.2 * i, 0.0))
        cols = [np.asarray(c, np.int32) for c in cloth.grid_cloth_coloring(n)]
        if i == 1:
            cols = [np.concatenate([cols[0], cols[4]]), np.concatenate([cols[1], cols[5]]), cols[2], cols[3], np.concatenate([cols[6], cols[7]]), cols[8]]   # (not proper: only the merge is under test)
        if i == 3:
            cols = [np.concatenate(cols[0:3]), np.concatenate(cols[3:5]), np.concatenate(cols[5:8]), cols[8]]
        order = rng.permutation(len(cols))
        meshes.append((v, f, [cols[k] for k in order]))
    return meshes


def inertia_cases(n=200, seed=133):
    """masses, inertia targets and positions for the inertia term, with force / Hessian accumulators that already hold something"""
    rng = np.random.default_rng(seed)
    return dict(mass=(10.0 ** rng.uniform(-4, 1, n)).astype(np.float32), inertia=rng.uniform(-3, 3, (n, 3)).astype(np.float32),
                x=rng.uniform(-3, 3, (n, 3)).astype(np.float32), dtsq=float(np.float32(240.0) * np.float32(240.0)),
                f0=(rng.normal(0, 100, (n, 3))).astype(np.float32), h0=(rng.normal(0, 1e4, (n, 9))).astype(np.float32))


def vertex_step_tail_cases(n=240, seed=131):
    """(total force, total Hessian, position) triples for the tail of the vertex step: well-conditioned SPD Hessians of the size the
    step meets (mass / dt^2 + stiffness), forces just below and above the threshold |f|^2 = 1e-10, Hessians that are singular or
    nearly singular (the 3x3 solve fails: gradient step), tiny, indefinite"""
    rng = np.random.default_rng(seed)
    f = (rng.normal(0, 1, (n, 3)) * 10.0 ** rng.uniform(-2, 4, (n, 1))).astype(np.float32)
    H = np.zeros((n, 3, 3), np.float32)
    for i in range(n):
        A = rng.normal(0, 1, (3, 3)); S = A @ A.T + 0.1 * np.eye(3)
        kind = i % 8
        if kind == 5:
            S = np.outer(A[:, 0], A[:, 0]) + 1e-5 * np.eye(3)                 # rank 1 plus a whiff (an exactly singular Hessian whose cofactors vanish
            # makes the reference divide 0 by 0 and exit: "descentDirection has NaN")
        elif kind == 6:
            S = S - 2.0 * np.linalg.eigvalsh(S)[1] * np.eye(3)                # indefinite
        elif kind == 7:
            S = (np.outer(A[:, 0], A[:, 0]) + np.outer(A[:, 1], A[:, 1]) + 1e-7 * np.eye(3)) if i % 16 == 7 else S * 1e-12      # rank 2 plus a whiff; tiny
        H[i] = (S * 10.0 ** rng.uniform(2, 6)).astype(np.float32)
    f[3::8] = (rng.normal(0, 1, (len(f[3::8]), 3)) * rng.choice([2e-6, 5e-6, 8e-6, 2e-5], (len(f[3::8]), 1))).astype(np.float32)      # around the threshold 1e-5
    x = rng.uniform(-2, 2, (n, 3)).astype(np.float32)
    return dict(f=np.ascontiguousarray(f), h=np.ascontiguousarray(H.transpose(0, 2, 1).reshape(n, 9)), x=np.ascontiguousarray(x), step=0.9, step_gd=2e-6)


def step_schedule_case(dcd=1, ccd=1):
    """Three Kuhn blocks for the SCHEDULE of the step: block 0 with fixed vertices, block 1 dropped onto it (contacts from the first
    detection on), block 2 appearing in frame 1 (frameToAppear); 2 substeps x 6 iterations, friction from iteration 2 on,
    intermediate detection every 3 iterations, tree rebuilds every 2 / 1 / 2 frames (tets / surface / CCD). Single-threaded oracle
    (the trace is ordered). dcd / ccd: which detectors run (both; DCD alone; CCD alone)."""
    from oracle.binding import OracleScene
    phys = capi.default_physics_params(numSubsteps=2, iterations=6, frictionStartIter=2, intermediateCollisionIterations=3, collisionStiffness=1e5,
                                       worldBoundsMin=[-10, 0, -10], worldBoundsMax=[10, 20, 10], dcdTetMeshSceneBVHRebuildSteps=2,
                                       dcdSurfaceSceneBVHRebuildSteps=1, ccdBVHRebuildSteps=2)
    o = OracleScene(phys, capi.default_collision_params(allowDCD=dcd, allowCCD=ccd), threads=1)
    fixed = [np.array([0, 3, 7, 20], np.int32), np.zeros(0, np.int32), np.array([1], np.int32)]
    shapes = [((3, 2, 2, 0.25), (0.0, 0.3, 0.0), 0), ((2, 2, 2, 0.22), (0.1, 0.72, 0.05), 0), ((2, 1, 2, 0.2), (0.3, 1.4, 0.1), 1)]
    sizes = []
    for i, ((nx, ny, nz, h), org, appear) in enumerate(shapes):
        verts, tets = meshio.kuhn_beam(nx, ny, nz, h=h, origin=org)
        colors = meshio.kuhn_coloring(nx, ny, nz) if i != 1 else meshio.greedy_coloring(len(verts), tets)
        obj = capi.default_object_params(miu=1e5, lmbd=1e6, density=50 + 10 * i, frictionDynamic=0.2, initialVelocity=[0.0, -4.0 * i, 0.0], frameToAppear=appear)
        o.add_tet_mesh(verts, tets, obj, colors, fixed[i])
        sizes.append(len(verts))
    o.finalize()
    fixed_off = np.zeros(len(sizes) + 1, np.int32); fixed_off[1:] = np.cumsum(sizes)
    mask = np.zeros(int(fixed_off[-1]), np.uint8)
    for i, fx in enumerate(fixed):
        mask[fixed_off[i] + fx] = 1
    groups = o.parallel_groups()
    group_off = np.zeros(len(groups) + 1, np.int32); group_off[1:] = np.cumsum([2 * len(g) for g in groups])
    pairs = np.ascontiguousarray(np.concatenate([g.reshape(-1) for g in groups]), np.int32)
    svids = [o.surface_vertex_ids(m) for m in range(3)]
    sv_off = np.zeros(4, np.int32); sv_off[1:] = np.cumsum([len(x) for x in svids])

    def params(frame):      # ref_step_trace's params10
        return np.array([2, 6, 2, 3, frame, dcd, ccd, 2, 1, 2], np.int32)
    return dict(scene=o, params=params, fixed_off=fixed_off, mask=mask, group_off=group_off, pairs=pairs, n_groups=len(groups), sv_off=sv_off,
                svids=np.ascontiguousarray(np.concatenate(svids), np.int32), n_verts=int(fixed_off[-1]),
                active=[np.array([1, 1, 0], np.int32), np.array([1, 1, 1], np.int32), np.array([1, 1, 1], np.int32)])


def prepared_collision_cases():
    """Scene states whose detection results feed prepareCollisionDataCPU: the three-block schedule scene after one and two frames,
    and a four-block stack (config 2 in miniature) after 3 and 6 frames - DCD and CCD contacts, on the order of a hundred of them"""
    from conftest import make_contact_scene
    from oracle.binding import OracleScene
    o = step_schedule_case()["scene"]
    for tag in ("blocks3_f1", "blocks3_f2"):
        o.run_frames(1)
        yield tag, o                      # (a generator: the caller looks at the scene before it moves on)
    o = make_contact_scene(OracleScene, n_objects=4, threads=1)
    for tag in ("stack4_f3", "stack4_f6"):
        o.run_frames(3)
        yield tag, o


def prepared_collision_inputs(o):
    """what ref_prepare_collision_data takes, from the oracle's current state"""
    nM = o.num_meshes
    sizes = [o.mesh_sizes(m) for m in range(nM)]
    v_off = np.zeros(nM + 1, np.int32); sv_off = np.zeros(nM + 1, np.int32); f_off = np.zeros(nM + 1, np.int32)
    v_off[1:] = np.cumsum([s[0] for s in sizes]); sv_off[1:] = np.cumsum([s[2] for s in sizes]); f_off[1:] = np.cumsum([s[3] for s in sizes])
    active, rel, res = o.prepared_raw()
    cat = lambda xs, dt: np.ascontiguousarray(np.concatenate(xs), dt)
    groups = o.parallel_groups()
    return dict(nM=nM, v_off=v_off, sv_off=sv_off, f_off=f_off, vertex_group=cat([r[3] for r in rel], np.int32),
                svids=cat([o.surface_vertex_ids(m) for m in range(nM)], np.int32), faces=cat([o.surface_faces(m).reshape(-1) for m in range(nM)], np.int32),
                has=cat([r[0] for r in res], np.uint8), ccd=cat([r[1] for r in res], np.uint8), path=cat([r[2] for r in res], np.uint8),
                hit_mesh=cat([r[3] for r in res], np.int32), hit_face=cat([r[4] for r in res], np.int32),
                group_sizes=np.array([len(g) for g in groups], np.int32), active=active, rel=rel)


def cloth_bounds_case(seed=95):
    """What the VF / EE queries leave behind on a 7 x 7 cloth, made up at random: per vertex / face / edge the smallest distance met
    (the query radius where nothing was met, some below the thickness) and the number of contacts recorded (mostly none), and a
    mask that already has some vertices raised: the inputs of the reference's conservative-bounds loop."""
    from gaia_b200 import cloth
    from oracle.binding_cloth import OracleClothScene
    rng = np.random.default_rng(seed)
    R, thick, relax = 0.05, 0.004, 0.42
    v, f = cloth.grid_cloth(7, 1.0)
    o = OracleClothScene(cloth.default_cloth_physics_params(maxQueryDis=R, thickness=thick, conservativeStepRelaxation=relax, contactRadius=0.03), threads=1)
    o.add_tri_mesh(v, f, cloth.default_cloth_object_params(), cloth.grid_cloth_coloring(7))
    o.finalize()
    nV, nF, nE = o.mesh_sizes(0)

    def leftovers(n):
        d = np.where(rng.random(n) < 0.8, R, rng.uniform(0.0, R, n)).astype(np.float32)
        d[rng.random(n) < 0.015] = np.float32(0.5 * thick)                    # closer than the thickness: the bound clamps at 0
        c = np.where((d < 0.03) & (rng.random(n) < 0.6), rng.integers(1, 4, n), 0).astype(np.int32)
        return d, c
    vf_min, vf_cnt = leftovers(nV); face_min, face_cnt = leftovers(nF); ee_min, ee_cnt = leftovers(nE)
    mask = (rng.random(nV) < 0.15).astype(np.uint8)
    return dict(scene=o, sizes=(nV, nF, nE), R=R, thick=thick, relax=relax, vf_min=vf_min, vf_cnt=vf_cnt, face_min=face_min, face_cnt=face_cnt,
                ee_min=ee_min, ee_cnt=ee_cnt, mask=mask)


def build_cloth_constraint_scene(cls, case, phys=None, with_stiffness=True, **kw):
    """the scene of cloth_constraint_case on either implementation (colourings that keep coupled vertices apart are the
    caller's business: the probes do not step)"""
    from gaia_b200 import cloth
    s = cls(phys if phys is not None else cloth.default_cloth_physics_params(handleCollision=0), **kw)
    for i, (v, f, n) in enumerate(case["meshes"]):
        s.add_tri_mesh(v, f, cloth.default_cloth_object_params(), cloth.grid_cloth_coloring(n) if i < case["n_sim"] else (), simulated=i < case["n_sim"])
    s.set_sewing(0, case["seams"], case["ratios"], case["sew_k"])
    for i in range(len(case["v_infos"])):
        s.add_attachment(case["v_infos"][i, 0], case["v_infos"][i, 1], case["f_infos"][i, 0], case["f_infos"][i, 1:], case["barys"][i],
                         float(case["stiff"][i]) if with_stiffness else 1e4)
    s.finalize()
    return s


def traverse_cases(seed=81, n_rays=400):
    """Rays for TetMeshFEM::tetrahedralTraverseTo on two meshes - the reference's teapot (irregular valence, thin parts) and a
    Kuhn beam -, both deformed at random: from a point on a surface face towards the centroid of a random tet (the way DCD
    uses it: start = the tet behind that face, target = the tet holding the query point), with and without a distance limit,
    plus rays aimed away from their target (stopped by the boundary, the empty stack or the distance check)."""
    from gaia_b200.workloads import ref_mesh
    from oracle.binding import OracleScene
    rng = np.random.default_rng(seed)
    out = []
    for name in ("teapot", "kuhn"):
        if name == "teapot":
            verts, tets, colors = ref_mesh("teapot")
        else:
            verts, tets = meshio.kuhn_beam(6, 3, 3, h=0.25)
            colors = meshio.kuhn_coloring(6, 3, 3)
        o = OracleScene(capi.default_physics_params(), capi.default_collision_params(allowDCD=0, allowCCD=0), threads=1)
        o.add_tet_mesh(verts, tets, capi.default_object_params(), colors)
        o.finalize()
        rest = o.positions(0)
        scale = float(np.linalg.norm(rest.max(0) - rest.min(0)))
        pos = (rest + rng.normal(0, 0.002 * scale, rest.shape)).astype(np.float32)
        o.set_state(0, pos, None)
        t, nb, xs = o.tet_topology(0)
        faces = o.surface_faces(0)
        # the tet behind a surface face and the face's id in it (= the index of the tet vertex the face does not hold)
        key = {}
        for ti in range(len(t)):
            for i in range(4):
                if nb[ti, i] < 0:
                    key[tuple(sorted(int(x) for x in np.delete(t[ti], i)))] = (ti, i)
        rays = []
        for k in range(n_rays):
            f = int(rng.integers(len(faces)))
            st, sf = key[tuple(sorted(int(x) for x in faces[f]))]
            w = rng.dirichlet([2.0, 2.0, 2.0]).astype(np.float32)
            origin = (w[:, None] * pos[faces[f]]).sum(0).astype(np.float32)
            target = int(rng.integers(len(t)))
            cen = pos[t[target]].mean(0)
            d = cen - origin
            if k % 5 == 4:
                d = d + rng.normal(0, 0.6 * np.linalg.norm(d), 3)            # off target
            dist = float(np.linalg.norm(d))
            d = (d / max(dist, 1e-20)).astype(np.float32)
            max_dis = 0.0 if k % 3 == 0 else dist * float(rng.uniform(0.5, 1.6))
            rays.append((origin, d, np.float32(max_dis), st, sf, target))
        out.append(dict(name=name, scene=o, pos=np.ascontiguousarray(pos), tets=np.ascontiguousarray(t), nbr=np.ascontiguousarray(nb),
                        xor=np.ascontiguousarray(xs), rays=rays, eps=1e-10))
    return out


def feasible_region_cases(seed=91, n_pts=1500):
    """Query points around the surfaces of the reference's teapot and a Kuhn beam (both deformed at random): near face
    interiors, edges and corners, inside and outside, close and far - with the surface face they are tested against. The
    closest-point type that selects the branch is computed by the caller with closestPointTriangle."""
    rng = np.random.default_rng(seed)
    out = []
    for c in traverse_cases(seed=seed + 1, n_rays=1):
        o, pos = c["scene"], c["pos"]
        faces = np.ascontiguousarray(o.surface_faces(0), np.int32)
        f3, t2s, off, nb = o.surface_topology(0)
        scale = float(np.linalg.norm(pos.max(0) - pos.min(0)))
        pts = []
        for k in range(n_pts):
            f = int(rng.integers(len(faces)))
            a, b, cc = pos[faces[f]]
            mode = k % 3
            w = rng.dirichlet([1.0, 1.0, 1.0]) if mode == 0 else (np.array([*rng.dirichlet([1.0, 1.0]), 0.0])[rng.permutation(3)] if mode == 1 else np.eye(3)[rng.integers(3)])
            q = w[0] * a + w[1] * b + w[2] * cc
            n = np.cross(b - a, cc - a); n /= max(np.linalg.norm(n), 1e-20)
            edge = float(np.linalg.norm(b - a))
            q = q + n * rng.normal(0, 0.7 * edge) + rng.normal(0, 0.35 * edge, 3)
            pts.append((f, q.astype(np.float32)))
        out.append(dict(name=c["name"], scene=o, pos=pos, faces=faces, f3=np.ascontiguousarray(f3), t2s=np.ascontiguousarray(t2s), off=np.ascontiguousarray(off),
                        nb=np.ascontiguousarray(nb), pts=pts, eps=1e-2))
    return out


def cloth_feasible_region_cases(seed=95, n_pts=1500):
    """Query points around a wrinkled cloth (open boundary: boundary edges and corners included), with the face they are
    tested against; the closest-point type comes from closestPointTriangle."""
    rng = np.random.default_rng(seed)
    c = cloth_material_case(n=9, seed=seed + 1)
    o, pos = c["scene"], c["pos"]
    faces = c["faces"]
    f3, off, nb = o.topology(0)
    pts = []
    for k in range(n_pts):
        f = int(rng.integers(len(faces)))
        a, b, cc = pos[faces[f]]
        mode = k % 3
        w = rng.dirichlet([1.0, 1.0, 1.0]) if mode == 0 else (np.array([*rng.dirichlet([1.0, 1.0]), 0.0])[rng.permutation(3)] if mode == 1 else np.eye(3)[rng.integers(3)])
        q = w[0] * a + w[1] * b + w[2] * cc
        n = np.cross(b - a, cc - a); n /= max(np.linalg.norm(n), 1e-20)
        edge = float(np.linalg.norm(b - a))
        q = q + n * rng.normal(0, 0.5 * edge) + rng.normal(0, 0.3 * edge, 3)
        pts.append((f, q.astype(np.float32)))
    return dict(scene=o, pos=pos, faces=faces, f3=np.ascontiguousarray(f3), edges=c["edges"], off=np.ascontiguousarray(off), nb=np.ascontiguousarray(nb),
                pts=pts, eps=1e-3)


def step_sequence_cases(seed=101, n_steps=4):
    """Initial step + velocity update over four substeps for every initial-guess type the path supports, with velocity damping,
    a velocity cap inside / outside a no-gravity zone and fixed points; the positions "after the iterations" of each substep
    are given (the initial guess moved at random), so that the velocity history - which types 2, 4 and 6 read - evolves."""
    import ctypes
    powf = ctypes.CDLL("libm.so.6").powf
    powf.restype = ctypes.c_float; powf.argtypes = [ctypes.c_float, ctypes.c_float]
    rng = np.random.default_rng(seed)
    verts, tets = meshio.kuhn_beam(3, 2, 2, h=0.3, origin=(0, 1.0, 0))
    colors = meshio.kuhn_coloring(3, 2, 2)
    fixed = np.array([0, 5, 11], np.int32)
    cases = []
    for init_type, kw in ((0, {}), (1, dict(exponentialVelDamping=0.5, constantVelDamping=0.3)), (2, dict(maxVelocityMagnitude=1.5, hasNoGravZone=1, noGravZoneThreshold=1.3)),
                          (3, dict(constantVelDamping=4.0)), (4, dict(initRatio_g=0.6, exponentialVelDamping=0.8)), (5, dict(constantVelDamping=0.5)),
                          (6, dict(maxVelocityMagnitude=2.0, hasNoGravZone=1, noGravZoneThreshold=0.5))):
        cases.append((init_type, kw))

    def make(init_type, kw):
        from oracle.binding import OracleScene
        phys = capi.default_physics_params(numSubsteps=3, iterations=1, gravity=[0.4, -9.0, 0.2], worldBoundsMin=[-50, -50, -50], worldBoundsMax=[50, 50, 50])
        obj = capi.default_object_params(initializationType=init_type, density=40, **kw)
        o = OracleScene(phys, capi.default_collision_params(allowDCD=0, allowCCD=0), threads=1)
        o.add_tet_mesh(verts, tets, obj, colors, fixed)
        o.finalize()
        dt = np.float32(phys.timeStep) / np.float32(3)
        ed = float(powf(float(np.float32(kw.get("exponentialVelDamping", 1.0))), float(dt)))
        cd = float(np.float32(kw.get("constantVelDamping", 0.0)) * dt)
        params12 = np.array([dt, 0.4, -9.0, 0.2, 1.0, init_type, kw.get("initRatio_g", obj.initRatio_g), kw.get("maxVelocityMagnitude", obj.maxVelocityMagnitude),
                             kw.get("hasNoGravZone", 0), kw.get("noGravZoneThreshold", 0.0), ed, cd], np.float32)
        return o, params12
    n = len(verts)
    pos0 = (verts + rng.normal(0, 0.01, verts.shape)).astype(np.float32)
    vel0 = rng.normal(0, 1.0, verts.shape).astype(np.float32)
    noise = rng.normal(0, 0.01, (n_steps, n, 3)).astype(np.float32)
    return dict(cases=cases, make=make, pos0=pos0, vel0=vel0, noise=noise, fixed=fixed, n_steps=n_steps)


def run_oracle_step_sequence(o, c, forces=None):
    """the oracle through the same sequence: (positions after the initial step, inertia, velocities after the update, solved) per substep;
    forces (a list) collects the elastic forces at the positions each substep starts from - what initial-guess type 5 reads"""
    o.set_state(0, c["pos0"], c["vel0"])
    init, inert, vel, solved = [], [], [], []
    for s in range(c["n_steps"]):
        if forces is not None:
            forces.append(o.neohookean_forces(0))
        o.phase_begin_substep()
        init.append(o.positions(0)); inert.append(o.inertia(0))
        sol = (init[-1] + c["noise"][s]).astype(np.float32)
        sol[c["fixed"]] = init[-1][c["fixed"]]
        o.set_state(0, sol, None)
        o.phase_end_substep()
        vel.append(o.velocities(0)); solved.append(sol)
    return np.array(init), np.array(inert), np.array(vel), np.array(solved)


def cloth_step_sequence_cases(seed=111, n_steps=4):
    """Cloth initial step + velocity update over four substeps for every initial-guess type (0 inertia, 1 adaptive, 2 previous
    position, 3 inertia clipped to the conservative bound), with damping, the velocity cap / no-gravity zone and fixed points,
    plus candidate positions for the conservative-bound truncation."""
    from gaia_b200 import cloth
    rng = np.random.default_rng(seed)
    n = 5
    verts, faces = cloth.grid_cloth(n, size=1.0, origin=(0.0, 1.0, 0.0))
    fixed = np.array([0, 4], np.int32)
    cases = [(0, {}), (1, dict(exponentialVelDamping=0.9, constantVelDamping=0.05)), (2, dict(maxVelocityMagnitude=1.0, hasNoGravZone=1, noGravZoneThreshold=-0.3)),
             (3, dict(constantVelDamping=0.2))]

    def make(init_type, kw):
        from oracle.binding_cloth import OracleClothScene
        phys = cloth.default_cloth_physics_params(numSubsteps=3, iterations=1, handleCollision=1, gravity=[0.3, -9.0, 0.1])
        obj = cloth.default_cloth_object_params(initializationType=init_type, **kw)
        o = OracleClothScene(phys, threads=1)
        o.add_tri_mesh(verts, faces, obj, cloth.grid_cloth_coloring(n), fixed=fixed)
        o.finalize()
        dt = np.float32(phys.timeStep) / np.float32(3)
        params12 = np.array([dt, 0.3, -9.0, 0.1, 1.0, 1.0, init_type, kw.get("maxVelocityMagnitude", obj.maxVelocityMagnitude), kw.get("hasNoGravZone", 0),
                             kw.get("noGravZoneThreshold", 0.0), kw.get("exponentialVelDamping", 1.0), kw.get("constantVelDamping", 0.0)], np.float32)
        return o, params12
    nV = len(verts)
    pos0 = (verts + rng.normal(0, 0.01, verts.shape)).astype(np.float32)
    vel0 = rng.normal(0, 0.8, verts.shape).astype(np.float32)
    noise = rng.normal(0, 0.01, (n_steps, nV, 3)).astype(np.float32)
    bounds = rng.uniform(0.002, 0.08, nV).astype(np.float32)          # some smaller than a substep's inertia step, some larger
    trunc_verts = rng.integers(0, nV, 200).astype(np.int32)
    trunc_pos = (pos0[trunc_verts] + rng.normal(0, 0.04, (200, 3))).astype(np.float32)
    return dict(cases=cases, make=make, pos0=pos0, vel0=vel0, noise=noise, fixed=fixed, n_steps=n_steps, bounds=bounds, trunc_verts=trunc_verts, trunc_pos=trunc_pos)


def run_cloth_oracle_step_sequence(o, c):
    o.set_state(0, c["pos0"], c["vel0"])
    o.set_bounds(0, c["bounds"], c["pos0"])
    init, inert, vel, solved = [], [], [], []
    for s in range(c["n_steps"]):
        o.probe_initial_step()
        init.append(o.positions(0)); inert.append(o.inertia(0))
        sol = (init[-1] + c["noise"][s]).astype(np.float32)
        sol[c["fixed"]] = init[-1][c["fixed"]]
        o.set_state(0, sol, None)
        o.probe_update_velocity()
        vel.append(o.velocities(0)); solved.append(sol)
    return np.array(init), np.array(inert), np.array(vel), np.array(solved)


def boundary_cases_cpu(n=120, seed=35):
    """World-box cases for the CPU rule (penalty + friction per violated axis): none, one, two and all three axes violated, on
    either side, with displacements below and above epsU."""
    rng = np.random.default_rng(seed)
    mn = np.array([-2.0, 0.0, -3.0], np.float32); mx = np.array([2.0, 5.0, 3.0], np.float32)
    p = rng.uniform(mn + 0.1, mx - 0.1, (n, 3)).astype(np.float32)
    for i in range(n):
        for d in range(3):
            if (i >> d) & 1:
                p[i, d] = (mn[d] - rng.uniform(1e-3, 0.2)) if rng.integers(2) else (mx[d] + rng.uniform(1e-3, 0.2))
    scale = np.where(np.arange(n)[:, None] % 4 == 0, 1e-4, 0.03)
    pPrev = (p + rng.normal(0, 1, p.shape) * scale).astype(np.float32)
    return dict(p=np.ascontiguousarray(p), pPrev=np.ascontiguousarray(pPrev), mn=mn, mx=mx, kB=1e5, muB=0.3, epsVB=1.0, dt=1.0 / 240)
