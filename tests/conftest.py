import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Native pieces are built in-tree by __graft_entry__.build(); build on demand when missing."""
    need = [os.path.join(ROOT, "gaia_b200", "csrc", "libgaia_b200.so"), os.path.join(ROOT, "oracle", "liboracle.so")]
    if not all(os.path.exists(p) for p in need):
        import __graft_entry__
        __graft_entry__.build()


def has_gpu():
    try:
        import ctypes
        cuda = ctypes.CDLL("libcuda.so.1")
        n = ctypes.c_int(0)
        return cuda.cuInit(0) == 0 and cuda.cuDeviceGetCount(ctypes.byref(n)) == 0 and n.value > 0
    except OSError:
        return False


def make_beam_scene(cls, nx=6, ny=3, nz=3, h=0.25, y0=0.05, substeps=2, iterations=5, v0=(0.3, -5, 0.2), obj_kw=None, phys_kw=None,
                    col_kw=None, coloring="kuhn", **cls_kw):
    """A Kuhn-subdivided beam above the ground plane y=0 (config 1 shape: one tet mesh, ground contact only).
    The default initial velocity has tangential components on purpose: the reference's friction Hessian is discontinuous
    at zero tangential displacement (`if (uNorm > 0)`, M/VBD/VBDPhysics.cpp:2567), so a purely vertical drop makes the
    first impact depend on rounding noise in ANY two implementations."""
    from gaia_b200 import capi, meshio
    verts, tets = meshio.kuhn_beam(nx, ny, nz, h=h, origin=(0, y0, 0))
    colors = meshio.kuhn_coloring(nx, ny, nz) if coloring == "kuhn" else meshio.greedy_coloring(len(verts), tets)
    pk = dict(numSubsteps=substeps, iterations=iterations, worldBoundsMin=[-10, 0, -10], worldBoundsMax=[10, 20, 10])
    pk.update(phys_kw or {})
    ck = dict(allowDCD=0, allowCCD=0)
    ck.update(col_kw or {})
    ok = dict(miu=1e5, lmbd=1e6, density=50, initialVelocity=list(v0))
    ok.update(obj_kw or {})
    s = cls(capi.default_physics_params(**pk), capi.default_collision_params(**ck), **cls_kw)
    s.add_tet_mesh(verts, tets, capi.default_object_params(**ok), colors)
    s.finalize()
    return s


def rel_err(a, b):
    return float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / max(np.abs(b).max(), 1e-30))


def per_vertex_rel(a, b, scale=None):
    """north_star's "per-vertex positions within 1e-5 relative": max over vertices of |a_v - b_v| / max(|b_v|, L), L = the
    diagonal of b's bounding box unless given (a vertex near the origin is measured against the size of the scene)."""
    a = np.asarray(a, np.float64).reshape(-1, 3); b = np.asarray(b, np.float64).reshape(-1, 3)
    if scale is None:
        scale = float(np.linalg.norm(b.max(0) - b.min(0)))
    den = np.maximum(np.linalg.norm(b, axis=1), max(scale, 1e-30))
    return float((np.linalg.norm(a - b, axis=1) / den).max())


def make_contact_scene(cls, n_objects=2, ccd=1, dcd=1, substeps=2, iterations=6, phys_kw=None, col_kw=None, gap=-0.08, v_rel=(0.0, -6.0, 0.0),
                       **cls_kw):
    """Config-2 shape in miniature: Kuhn-grid blocks stacked with a small overlap (gap < 0) above the ground plane, each
    with its own material and colouring, the upper ones moving down. Returns the finalized scene."""
    from gaia_b200 import capi, meshio
    pk = dict(numSubsteps=substeps, iterations=iterations, worldBoundsMin=[-10, 0, -10], worldBoundsMax=[10, 20, 10], collisionStiffness=1e5)
    pk.update(phys_kw or {})
    ck = dict(allowDCD=dcd, allowCCD=ccd)
    ck.update(col_kw or {})
    s = cls(capi.default_physics_params(**pk), capi.default_collision_params(**ck), **cls_kw)
    y = 0.3
    shapes = [(4, 2, 3, 0.25), (3, 2, 2, 0.22), (2, 3, 2, 0.2), (3, 3, 3, 0.15)]
    for i in range(n_objects):
        nx, ny, nz, h = shapes[i % len(shapes)]
        ox = 0.11 * i - 0.05 * (i % 2); oz = 0.07 * (i % 3)
        verts, tets = meshio.kuhn_beam(nx, ny, nz, h=h, origin=(ox, y, oz))
        colors = meshio.kuhn_coloring(nx, ny, nz) if i % 2 == 0 else meshio.greedy_coloring(len(verts), tets)
        obj = capi.default_object_params(miu=1e5 * (1 + 0.5 * i), lmbd=1e6, density=50 + 10 * i, frictionDynamic=0.2 + 0.1 * i,
                                         initialVelocity=[v_rel[0] * i, v_rel[1] * i, v_rel[2] * i],
                                         rotation=[0.0, 0.15 * i, 0.05 * i] if i else [0, 0, 0])
        s.add_tet_mesh(verts, tets, obj, colors)
        y += ny * h + gap
    s.finalize()
    return s


def contacts_table(c):
    """structured contact array -> plain int array for exact comparison"""
    return np.stack([c[n] for n in c.dtype.names], 1) if len(c) else np.zeros((0, 8), np.int32)
