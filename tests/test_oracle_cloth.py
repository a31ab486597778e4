"""CPU: the cloth oracle (CPU restatement of M/VBDCloth + ClothContactDetector). The reference cloth code cannot be built
here (Eigen/oneTBB/Embree absent) and ships no golden vectors, so the restatement is pinned to closed-form properties:
topology against a numpy recomputation, forces/Hessians against finite differences of the energies, contact sets against
brute force, and the physical invariants of a drape (no tunnelling, bounded steps)."""
import numpy as np
import pytest

from gaia_b200 import cloth
from oracle.binding_cloth import OracleClothScene, closest_points_segments


def make_scene(n=8, phys_kw=None, obj_kw=None, fixed=(), threads=2, height=1.0, sphere=None, brute=False, wrinkle=0.0, seed=0):
    v, f = cloth.grid_cloth(n, 1.0, origin=(0, height, 0))
    if wrinkle:
        rng = np.random.RandomState(seed)
        v = v + rng.uniform(-wrinkle, wrinkle, v.shape).astype(np.float32)
    cols = cloth.grid_cloth_coloring(n)
    pk = dict(numSubsteps=4, iterations=6, contactRadius=0.03, maxQueryDis=0.05, contactStiffness=1e4)
    pk.update(phys_kw or {})
    ok = dict(bendingStiffness=5.0, density=0.2, miu=1e3, dampingStVK=1e-6, dampingBending=1e-6)
    ok["lambda"] = 1e3
    ok.update(obj_kw or {})
    o = OracleClothScene(cloth.default_cloth_physics_params(**pk), threads=threads)
    o.set_brute_force(brute)
    o.add_tri_mesh(v, f, cloth.default_cloth_object_params(**ok), cols, fixed=fixed)
    if sphere is not None:
        sv, sf = cloth.uv_sphere(*sphere)
        o.add_tri_mesh(sv, sf, cloth.default_cloth_object_params(), simulated=False)
    o.finalize()
    return o, v, f


def test_topology_matches_numpy_recomputation():
    o, v, f = make_scene(7)
    nV, nF, nE = o.mesh_sizes(0)
    assert nV - nE + nF == 1                       # disc
    E = o.edges(0)
    und = {tuple(sorted(e[:2])) for e in E}
    assert len(und) == nE
    for e in E:
        eV1, eV2, n12, n21, f1, f2 = e
        # fId1 is oriented eV1 -> eV2 and eV12Next completes it; fId2 the other way round
        tri = list(f[f1]); i = tri.index(eV1)
        assert tri[(i + 1) % 3] == eV2 and tri[(i + 2) % 3] == n12
        if f2 >= 0:
            tri = list(f[f2]); i = tri.index(eV2)
            assert tri[(i + 1) % 3] == eV1 and tri[(i + 2) % 3] == n21
        else:
            assert n21 == -1
    # creation order: first edge of the first face is (v2 -> v0)
    assert tuple(E[0][:2]) == (f[0][2], f[0][0])
    masses = o.vertex_masses(0)
    assert np.isclose(masses.sum(), 0.2 * 1.0, rtol=1e-5)     # density * area


def test_parallel_groups_are_a_valid_colouring_sorted_by_size():
    n = 10
    o, v, f = make_scene(n)
    groups = o.parallel_groups()
    sizes = [len(g) for g in groups]
    assert sum(sizes) == n * n
    assert cloth.check_cloth_coloring(n * n, f, [g[:, 1] for g in groups])
    assert cloth.check_cloth_coloring(n * n, f, cloth.greedy_cloth_coloring(n * n, f))


def _fd_force(o, mesh, x, vid, eps):
    f = np.zeros(3)
    for d in range(3):
        xp = x.copy(); xp[vid, d] += eps
        xm = x.copy(); xm[vid, d] -= eps
        o.set_state(mesh, xp); ep = o.elastic_energy(mesh)
        o.set_state(mesh, xm); em = o.elastic_energy(mesh)
        f[d] = -(ep - em) / (2 * eps)
    o.set_state(mesh, x)
    return f


def test_material_force_is_minus_energy_gradient_and_hessian_its_jacobian():
    # no damping: force = -dE/dx exactly (StVK membrane + quadratic bending)
    o, v, f = make_scene(6, obj_kw=dict(dampingStVK=0.0, dampingBending=0.0, bendingStiffness=20.0), wrinkle=0.03)
    x = o.positions(0).astype(np.float32)
    o.set_prev_positions(0, x)
    for vid in (0, 7, 14, 21, 35):
        force, H = o.material_force_hessian(0, vid)
        fd = _fd_force(o, 0, x, vid, 2e-3)
        assert np.allclose(force, fd, rtol=3e-2, atol=3e-2 * np.abs(fd).max()), (vid, force, fd)
        # Hessian: finite difference of the analytic force
        Hfd = np.zeros((3, 3))
        for d in range(3):
            xp = x.copy(); xp[vid, d] += 1e-3
            xm = x.copy(); xm[vid, d] -= 1e-3
            o.set_state(0, xp); fp, _ = o.material_force_hessian(0, vid)
            o.set_state(0, xm); fm, _ = o.material_force_hessian(0, vid)
            Hfd[:, d] = -(fp.astype(np.float64) - fm) / 2e-3
        o.set_state(0, x)
        assert np.allclose(H, H.T, rtol=1e-4, atol=1e-3 * np.abs(H).max())
        assert np.allclose(H, Hfd, rtol=3e-2, atol=3e-2 * np.abs(H).max()), (vid, H, Hfd)


def test_rayleigh_damping_terms():
    # damping adds gamma/dt * H_material * (x - x_prev) to the force and scales the Hessian by (1 + gamma/dt)
    o0, v, f = make_scene(6, obj_kw=dict(dampingStVK=0.0, dampingBending=0.0), wrinkle=0.02)
    o1, _, _ = make_scene(6, obj_kw=dict(dampingStVK=2e-3, dampingBending=2e-3), wrinkle=0.02)
    x = o0.positions(0)
    prev = x - 0.01
    for o in (o0, o1):
        o.set_prev_positions(0, prev)
    dt = float(o0.phys.timeStep / o0.phys.numSubsteps)
    g = 2e-3 / dt
    for vid in (8, 15):
        f0, H0 = o0.material_force_hessian(0, vid)
        f1, H1 = o1.material_force_hessian(0, vid)
        assert np.allclose(H1, H0 * (1 + g), rtol=1e-4, atol=1e-4 * np.abs(H0).max())
        assert np.allclose(f1, f0 - g * (H0 @ (x[vid] - prev[vid])), rtol=1e-3, atol=1e-3 * np.abs(f0).max())


def test_contact_repulsion_follows_the_reference_two_stage_energy():
    """computeContactRepulsiveForce, contactEnergyTypes = 1 (M/VBDCloth/VBDClothPhysics.cpp:1548-1562): quadratic above
    tau = r/2, k2/d below with k2 = tau^2 k / 2. (With this k2 the force halves across tau: the reference's comment says
    "C2 continuous" but the code it ships is not; the restatement follows the code.)"""
    o, _, _ = make_scene(4, phys_kw=dict(contactRadius=0.1, thickness=0.01, contactStiffness=1e4))
    r, t, k = 0.1, 0.01, 1e4
    tau = r / 2
    dE, d2E = o.contact_repulsion(0.08 + t)
    assert np.isclose(dE, -k * (r - 0.08), rtol=1e-5) and np.isclose(d2E, k)
    k2 = 0.5 * tau * tau * k
    for d in (0.01, 0.03, 0.0499):
        dE, d2E = o.contact_repulsion(d + t)
        assert np.isclose(dE, -k2 / d, rtol=1e-5) and np.isclose(d2E, k2 / d ** 2, rtol=1e-5)
    # below CMP_EPSILON the quadratic branch is used again
    dE, d2E = o.contact_repulsion(t + 5e-6)
    assert np.isclose(dE, -k * (r - 5e-6), rtol=1e-4) and np.isclose(d2E, k)
    # always repulsive inside the radius
    assert all(o.contact_repulsion(x)[0] < 0 for x in np.linspace(t + 1e-3, t + r - 1e-4, 40))


def test_segment_segment_closest_points():
    rng = np.random.RandomState(3)
    for _ in range(200):
        p0, p1, q0, q1 = rng.uniform(-1, 1, (4, 3)).astype(np.float32)
        c1, c2, s, t = closest_points_segments(p0, p1, q0, q1)
        assert 0 <= s <= 1 and 0 <= t <= 1
        ss, tt = np.meshgrid(np.linspace(0, 1, 41), np.linspace(0, 1, 41), indexing="ij")
        pts1 = p0[None, None] + ss[..., None] * (p1 - p0)
        pts2 = q0[None, None] + tt[..., None] * (q1 - q0)
        brute = np.linalg.norm(pts1 - pts2, axis=-1).min()
        assert np.linalg.norm(c1 - c2) <= brute + 1e-5


def _two_layer_scene(brute, n=9, gap=0.02):
    phys = cloth.default_cloth_physics_params(numSubsteps=2, iterations=4, contactRadius=0.03, maxQueryDis=0.045)
    o = OracleClothScene(phys, threads=2)
    o.set_brute_force(brute)
    obj = cloth.default_cloth_object_params(bendingStiffness=1.0, density=0.2, miu=1e3)
    rng = np.random.RandomState(5)
    for layer in range(2):
        v, f = cloth.grid_cloth(n, 1.0, origin=(0.013 * layer, 1.0 + gap * layer, 0.007 * layer))
        v = v + rng.uniform(-0.004, 0.004, v.shape).astype(np.float32)
        o.add_tri_mesh(v, f, obj, cloth.grid_cloth_coloring(n))
    sv, sf = cloth.uv_sphere(0.2, 8, 12, center=(0, 0.78, 0))
    o.add_tri_mesh(sv, sf, cloth.default_cloth_object_params(), simulated=False)
    o.finalize()
    return o


def test_contact_sets_grid_equals_brute_force_and_are_canonical():
    a, b = _two_layer_scene(False), _two_layer_scene(True)
    a.detect_contacts(); b.detect_contacts()
    vfa, vfb = a.vf_contacts(), b.vf_contacts()
    eea, eeb = a.ee_contacts(), b.ee_contacts()
    assert len(vfa) > 50 and len(eea) > 50
    assert np.array_equal(vfa, vfb) and np.array_equal(eea, eeb)
    for m in range(2):
        assert np.array_equal(a.conservative_bounds(m), b.conservative_bounds(m))
    # sorted, no vertex holds the same closest primitive twice, no contact with an incident face
    key = vfa[:, :4]
    assert (np.lexsort(key.T[::-1]) == np.arange(len(key))).all()
    per_vertex = {}
    for vm, v, fm, f, pt, pid in vfa:
        assert (pt, pid) not in per_vertex.setdefault((vm, v), set())
        per_vertex[(vm, v)].add((pt, pid))
    # EE between simulated meshes is symmetric
    sim = eea[eea[:, 2] < 2]
    fwd = {tuple(r) for r in sim}
    assert all((m2, e2, m1, e1) in fwd for m1, e1, m2, e2 in sim)
    # collider vertices never query collider faces; collider edges are never query edges
    assert not ((vfa[:, 0] == 2) & (vfa[:, 2] == 2)).any()
    assert (eea[:, 0] < 2).all()
    # conservative bound = relaxation * (min distance - thickness), below the query distance
    assert (a.conservative_bounds(0) <= 0.45 * 0.045 + 1e-7).all() and (a.conservative_bounds(0) >= 0).all()


def test_cloth_drapes_over_sphere_without_tunnelling():
    R, c = 0.25, np.array([0, 0.6, 0], np.float32)
    o, v, f = make_scene(15, phys_kw=dict(numSubsteps=5, iterations=8, contactRadius=0.02, maxQueryDis=0.03, contactStiffness=1e5,
                                          gravity=[0, -10, 0]),
                         obj_kw=dict(density=0.2, miu=2e3, bendingStiffness=0.5), height=0.9, sphere=(R, 12, 20, c))
    x0 = o.positions(0)
    o.run_frames(30)
    x = o.positions(0)
    assert np.isfinite(x).all()
    d = np.linalg.norm(x - c, axis=1)
    assert d.min() > R * np.cos(np.pi / 20) - 0.02          # faceted sphere: stays outside (up to the facet sagitta)
    assert x[:, 1].mean() < x0[:, 1].mean() - 0.05           # it fell
    centre = np.argmin(np.linalg.norm(x0[:, [0, 2]], axis=1))
    assert x[centre, 1] > c[1] + R * 0.9                      # the middle rests on top of the sphere
    st = o.stats()
    assert st["numberCollisionDetection"] >= 30 * 5 and st["numContactsVF"] > 0


def test_accelerator_and_initialization_types_run():
    for it in (0, 1, 2, 3):
        o, v, f = make_scene(6, phys_kw=dict(applyAcceleration=1, acceleratorRho=0.9), obj_kw=dict(initializationType=it), fixed=[0, 5])
        o.run_frames(4)
        x = o.positions(0)
        assert np.isfinite(x).all() and np.allclose(x[[0, 5]], v[[0, 5]])
        assert x[:, 1].min() < 1.0 - 1e-3


def test_seam_filtered_pairs_are_skipped_before_they_are_measured():
    """ObjectParams::filteredPairsPath (TriMeshFEM::generate_sewing_filters, M/TriMesh/TriMesh.h:596-622; the ...WithSewing callbacks,
    M/CollisionDetector/ClothContactDetector.cpp:290-320, 815-852 - Embree-bound, so a restatement): on a sheet folded onto
    itself every 7th detected self contact is declared a filtered pair. Those pairs are gone (either way round for edges), the
    lists of vertices / edges that are in no filtered pair are untouched, a filtered vertex's conservative bound can only grow,
    the grid and brute force agree, and - as in the reference, which attaches the filtering callbacks only to meshes with seams -
    the lists do nothing without seams."""
    from cloth_filter_scene import build, pick_pairs
    base = build(OracleClothScene, dict(threads=2))
    base.detect_contacts()
    vf0, ee0 = base.vf_contacts(), base.ee_contacts()
    assert len(vf0) > 100 and len(ee0) > 100
    vf_pairs, ee_pairs = pick_pairs(vf0, ee0)
    o = build(OracleClothScene, dict(threads=2), vf_pairs=vf_pairs, ee_pairs=ee_pairs)
    o.detect_contacts()
    vf1, ee1 = o.vf_contacts(), o.ee_contacts()
    have_vf = {(int(r[1]), int(r[3])) for r in vf1}
    assert not any((int(v), int(f)) in have_vf for v, f in vf_pairs)
    have_ee = {(int(r[1]), int(r[3])) for r in ee1}
    assert not any((int(a), int(b)) in have_ee or (int(b), int(a)) in have_ee for a, b in ee_pairs)
    assert len(vf1) < len(vf0) and len(ee1) < len(ee0)
    touched_v = set(vf_pairs[:, 0].tolist()); touched_e = set(ee_pairs.ravel().tolist())
    keep = lambda rows, touched: np.array([r for r in rows if int(r[1]) not in touched])
    assert np.array_equal(keep(vf0, touched_v), keep(vf1, touched_v)) and np.array_equal(keep(ee0, touched_e), keep(ee1, touched_e))
    b0, b1 = base.conservative_bounds(0), o.conservative_bounds(0)
    assert (b1 >= b0).all() and (b1 > b0).any()
    bf = build(OracleClothScene, dict(threads=2), vf_pairs=vf_pairs, ee_pairs=ee_pairs)
    bf.set_brute_force(True); bf.detect_contacts()
    assert np.array_equal(bf.vf_contacts(), vf1) and np.array_equal(bf.ee_contacts(), ee1) and np.array_equal(bf.conservative_bounds(0), b1)
    ns = build(OracleClothScene, dict(threads=2), seams=False, vf_pairs=vf_pairs, ee_pairs=ee_pairs)
    ns.detect_contacts()
    assert np.array_equal(ns.vf_contacts(), vf0) and np.array_equal(ns.ee_contacts(), ee0)
