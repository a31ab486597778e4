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
