"""Backtracking line search on the vertex merit energy (the reference's APPLY_LOCAL_LINE_SEARCH, M/VBD/VBDPhysics.cpp:25,
2467-2474 - compiled out there; backTracingLineSearchVBD M/VBD/VBD_NeoHookean.h:115-153; evaluateVertexMeritEnergy
M/VBD/VBD_NeoHookean.cpp:90-124). The elastic part of the oracle's merit energy is pinned to the reference's own compiled per-tet
energy (tests/test_oracle_vs_ref.py::test_vertex_merit_energy_elastic_part); the search loop itself restates CPU members that
need Eigen + MeshFrame and are not buildable here, so these tests check it against an independent float64 statement: a colour sweep
with the line search on never raises the total merit energy. With the stable Neo-Hookean energy the merit energy of ONE vertex is
exactly quadratic in its position (|F|^2 is quadratic, det F affine), so a full Newton step is never rejected except by
rounding at convergence - which is why the reference can compile the search out; the tests therefore drive it with
stepSize = 2.5 (first trial overshoots the minimiser and is rejected, alpha = 1.25 is accepted; without the search the same
sweeps diverge). The GPU tests compare the CUDA variant (csrc/sweep.cu: k_sweep_linesearch, through the C ABI) with the oracle on the same inputs."""
import numpy as np
import pytest

from gaia_b200 import capi, meshio


def _scene(cls, noise=0.0, seed=3, iterations=4, ls=True, dims=(6, 3, 3), h=0.25, step=1.0, **kw):
    nx, ny, nz = dims
    verts, tets = meshio.kuhn_beam(nx, ny, nz, h=h, origin=(0, 2.0, 0))
    colors = meshio.kuhn_coloring(nx, ny, nz)
    phys = capi.default_physics_params(numSubsteps=1, iterations=iterations, worldBoundsMin=[-50, -50, -50], worldBoundsMax=[50, 50, 50],
                                       stepSize=step)
    col = capi.default_collision_params(allowDCD=0, allowCCD=0)
    obj = capi.default_object_params(miu=1e5, lmbd=1e6, density=50, initialVelocity=[0.3, -1.0, 0.2])
    s = cls(phys, col, **kw)
    s.add_tet_mesh(verts, tets, obj, colors)
    s.finalize()
    if noise:
        rng = np.random.default_rng(seed)
        s.set_state(0, (verts + rng.uniform(-noise * h, noise * h, verts.shape)).astype(np.float32), None)
    if ls:
        s.set_line_search(True, tau=0.5, tau_gd=0.1, max_iters=16)
    return s, verts, tets


def _total_merit_energy(o, tets, pos, inertia, mass, dt, miu=1e5, lmbd=1e6):
    """float64, written from the formula: sum_v 1/2 m |x - y|^2 / dt^2 + sum_t A (mu/2 (|F|^2 - 3) + lmbd/2 (det F - a)^2)"""
    Dm = o.dminv(0).astype(np.float64).reshape(-1, 3, 3).transpose(0, 2, 1)      # column-major storage -> [t, r, c]
    A = o.rest_volumes(0).astype(np.float64)
    p = pos.astype(np.float64)[tets]
    Ds = np.stack([p[:, 1] - p[:, 0], p[:, 2] - p[:, 0], p[:, 3] - p[:, 0]], 2)  # columns
    F = Ds @ Dm
    a = 1 + miu / lmbd
    el = A * (miu * 0.5 * ((F * F).sum((1, 2)) - 3) + lmbd * 0.5 * (np.linalg.det(F) - a) ** 2)
    d = pos.astype(np.float64) - inertia.astype(np.float64)
    return float((0.5 * mass.astype(np.float64) * (d * d).sum(1)).sum() / dt ** 2 + el.sum())


def test_oracle_line_search_never_raises_the_merit_energy_and_shrinks_steps():
    from oracle.binding import OracleScene
    o, verts, tets = _scene(OracleScene, noise=0.35, step=2.5, threads=1)
    dt = 1.0 / 60.0
    o.phase_begin_substep(); o.phase_prepare()
    y, m = o.inertia(0), o.vertex_masses(0)
    E = [_total_merit_energy(o, tets, o.positions(0), y, m, dt)]
    for it in range(4):
        for g in range(len(o.parallel_groups())):
            o.phase_solve_group(g, True)
            E.append(_total_merit_energy(o, tets, o.positions(0), y, m, dt))
    E = np.array(E)
    assert o.line_search_shrinks() > 0                                  # the search really rejected full steps somewhere
    assert (np.diff(E) <= 1e-6 * np.abs(E[:-1])).all(), np.diff(E) / E[:-1]
    assert E[-1] < 0.2 * E[0]
    # the same sweeps WITHOUT the search: steps of 2.5 x Newton overshoot and the energy grows (that is what the search is for)
    p, *_ = _scene(OracleScene, noise=0.35, ls=False, step=2.5, threads=1)
    p.phase_begin_substep(); p.phase_prepare()
    Ep = [_total_merit_energy(p, tets, p.positions(0), y, m, dt)]
    for it in range(4):
        for g in range(len(p.parallel_groups())):
            p.phase_solve_group(g, True)
        Ep.append(_total_merit_energy(p, tets, p.positions(0), y, m, dt))
    assert p.line_search_shrinks() == 0 and not (np.diff(Ep) <= 0).all()


def test_oracle_line_search_with_accepted_full_steps_equals_the_plain_step():
    """alpha * dx + x with alpha = stepSize accepted at the first trial is the plain update: a gentle scene far from
    convergence (one iteration) comes out identical wherever no trial was rejected."""
    from oracle.binding import OracleScene
    a, *_ = _scene(OracleScene, ls=True, iterations=1, threads=1)
    b, *_ = _scene(OracleScene, ls=False, iterations=1, threads=1)
    a.run_iterations(1); b.run_iterations(1)
    if a.line_search_shrinks() == 0:
        assert np.array_equal(a.positions(0), b.positions(0))
    else:
        assert np.abs(a.positions(0) - b.positions(0)).max() < 1e-3


@pytest.mark.gpu
@pytest.mark.parametrize("noise,step", [(0.0, 1.0), (0.35, 2.5)])
def test_gpu_line_search_matches_the_oracle(noise, step):
    """One and three iterations from the same state. Tolerance: 1e-5 relative to the scene size per vertex wherever both took
    the same number of trials; the accept / reject decisions compare float energies whose last bits differ between the two
    (the GPU sums force and Hessian in closed form), so a vertex may take one more or one less halving: at most 2 % of the
    vertices may differ by more, and the totals of rejected trials agree within 5 %."""
    from oracle.binding import OracleScene
    from conftest import per_vertex_rel
    for iters in (1, 3):
        g, verts, tets = _scene(capi.Scene, noise=noise, step=step)
        o, *_ = _scene(OracleScene, noise=noise, step=step, threads=2)
        g.run_iterations(iters); o.run_iterations(iters)
        pg, po = g.positions(0), o.positions(0)
        scale = float(np.linalg.norm(po.max(0) - po.min(0)))
        err = np.abs(pg.astype(np.float64) - po.astype(np.float64)).max(1) / scale
        assert (err > 1e-5).mean() <= 0.02, (noise, iters, float((err > 1e-5).mean()), float(err.max()))
        sg, so = g.line_search_shrinks(), o.line_search_shrinks()
        if noise:
            assert so > 0 and abs(sg - so) <= max(3, 0.05 * so), (sg, so)
        assert np.abs(po - verts).max() > 1e-4


@pytest.mark.gpu
def test_gpu_line_search_lowers_the_merit_energy_with_every_iteration():
    from oracle.binding import OracleScene
    o, verts, tets = _scene(OracleScene, noise=0.35, ls=False, step=2.5, threads=1)       # only for Dm^-1 / volumes / masses
    m, dt = o.vertex_masses(0), 1.0 / 60.0
    E = []
    for iters in (0, 1, 2, 4):
        g, *_ = _scene(capi.Scene, noise=0.35, step=2.5)
        g.run_iterations(iters)
        E.append(_total_merit_energy(o, tets, g.positions(0), g.inertia(0), m, dt))
    assert all(b <= a * (1 + 1e-6) for a, b in zip(E, E[1:])), E
    assert E[-1] < 0.2 * E[0]
    # switching the option off again restores the tile sweep (same result as a scene that never had it on)
    a, *_ = _scene(capi.Scene, noise=0.1)
    a.set_line_search(False)
    b, *_ = _scene(capi.Scene, noise=0.1, ls=False)
    a.run_iterations(2); b.run_iterations(2)
    assert np.array_equal(a.positions(0), b.positions(0))


@pytest.mark.gpu
def test_gpu_line_search_with_contacts_and_the_ground():
    """The contact variant of the kernel (contact + ground terms enter force and Hessian, NOT the merit energy, as in the
    reference - so the search rejects many steps here): stacked blocks in contact above the ground, one iteration from the
    same detected state. Same tolerance rule as above (decisions on float energies), and identical contact sets."""
    from oracle.binding import OracleScene
    from conftest import make_contact_scene, contacts_table
    g = make_contact_scene(capi.Scene, n_objects=3)
    o = make_contact_scene(OracleScene, n_objects=3, threads=2)
    for s in (g, o):
        s.set_line_search(True, tau=0.5, tau_gd=0.1, max_iters=16)
        s.run_iterations(1)
    assert len(o.contacts()) > 10 and np.array_equal(contacts_table(g.contacts()), contacts_table(o.contacts()))
    bad = tot = 0
    for m in range(3):
        pg, po = g.positions(m).astype(np.float64), o.positions(m).astype(np.float64)
        scale = float(np.linalg.norm(po.max(0) - po.min(0)))
        err = np.abs(pg - po).max(1) / scale
        bad += int((err > 1e-5).sum()); tot += len(err)
    assert bad <= 0.02 * tot, (bad, tot)
    sg, so = g.line_search_shrinks(), o.line_search_shrinks()
    assert so > 0 and abs(sg - so) <= max(3, 0.05 * so), (sg, so)
