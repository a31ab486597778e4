"""Pin the CPU oracle to the reference's own arithmetic.

tests/golden/ref_vectors.npz holds outputs of the REFERENCE's compiled code (oracle/_ref/libgaia_ref.so, built from
/root/reference; generator: tests/golden/make_golden.py) on the seeded inputs of tests/golden_inputs.py. The oracle is
replayed on the same inputs. Tolerances: the 3x3 solve, the cubic solver and the CCD test are the same sequence of
float operations -> bit-exact; force/Hessian assemblies sum in a different association -> 2e-5 relative to the
largest entry (the reference's GPU-variant code adds the mu*|m|^2*A damping term twice when damping > 0 — SURVEY.md §8(a)
divergences — so the damped case is compared with that term added back)."""
import ctypes as C
import os

import numpy as np
import pytest

from golden_inputs import (boundary_cases_cpu, cloth_step_sequence_cases, run_cloth_oracle_step_sequence, step_sequence_cases, run_oracle_step_sequence, cloth_feasible_region_cases, feasible_region_cases, traverse_cases, cloth_contact_case, cloth_material_case, neohookean_case, single_tet_cases, solve3x3_cases, ccd_cases, cubic_cases, contact_cases, boundary_cases,
                           point_in_tet_cases, closest_point_triangle_cases, segment_cases, cloth_repulsion_cases)
from oracle.binding import load_oracle, load_ref, ref_available

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_vectors.npz"))
FP = C.POINTER(C.c_float)


def fp(a):
    return a.ctypes.data_as(FP)


def relmax(a, b):
    return float(np.abs(a.astype(np.float64) - b.astype(np.float64)).max() / max(np.abs(b).max(), 1e-30))


def test_neohookean_force_hessian_plain():
    c = neohookean_case(seed=11, dampH=0.0, dampD=0.0)
    o = c["scene"]
    o.set_state(0, c["pos"], None); o.set_prev_positions(0, c["prev"])
    F = np.zeros((len(c["pos"]), 3), np.float32); H = np.zeros((len(c["pos"]), 9), np.float32)
    for v in range(len(c["pos"])):
        F[v], H[v] = o.vertex_material_force_hessian(0, v)
    assert relmax(F, GOLD["nh_plain_force"]) < 2e-5
    assert relmax(H, GOLD["nh_plain_hessian"]) < 2e-5


def test_neohookean_force_hessian_damped():
    """CPU semantics (the oracle) vs the reference's GPU-variant code on one-tet meshes. The GPU variant
    (VBDPhysicsCompute.cu:2147-2155, 2196-2221) (i) carries mu*|m|^2*A inside the Hessian it scales by dampingHydrostatic
    and (ii) adds mu*|m|^2*A to the diagonal a second time; both extras are closed-form for a vertex with a single
    incident tet, so: H_gpu = H_cpu + tmp*(1 + gamma_h/dt) I,  f_gpu = f_cpu - tmp*gamma_h/dt (x - x_prev)."""
    for i, c in enumerate(single_tet_cases()):
        o = c["scene"]
        o.set_state(0, c["pos"], None); o.set_prev_positions(0, c["prev"])
        B = c["DmInv"][0].astype(np.float64).reshape(3, 3, order="F")
        rows = [-(B[0] + B[1] + B[2]), B[0], B[1], B[2]]
        for v in range(4):
            F, H = o.vertex_material_force_hessian(0, v)
            tmp = float(rows[v] @ rows[v]) * c["miu"] * float(c["restVol"][0])
            disp = c["pos"][v].astype(np.float64) - c["prev"][v].astype(np.float64)
            Hx = H.astype(np.float64); Hx[[0, 4, 8]] += tmp * (1 + c["dampH"] / c["dt"])
            Fx = F.astype(np.float64) - tmp * c["dampH"] / c["dt"] * disp
            assert relmax(Hx, GOLD[f"nh_damped{i}_hessian"][v]) < 3e-5
            assert relmax(Fx, GOLD[f"nh_damped{i}_force"][v]) < 3e-5
            # and the damping terms are not negligible in this fixture
            assert tmp * c["dampD"] / c["dt"] > 1e-3 * np.abs(H).max()


def test_vertex_merit_energy_elastic_part():
    """The line search's merit energy (oracle: vertex_merit_energy_parts, restating evaluateVertexMeritEnergy,
    M/VBD/VBD_NeoHookean.cpp:90-124) against the reference's own compiled per-tet energy evaluateNeoHookeanEnergy
    (VBDPhysicsCompute.cu:1775-1808 = A (mu Phi_D + lmbd Phi_H) minus its rest value 0.5 A lmbd (1 - a)^2): the elastic part of
    a vertex is the sum over its incident tets. 1e-5 relative (different float association; the rest value is added back)."""
    c = neohookean_case(seed=11, dampH=0.0, dampD=0.0)
    o = c["scene"]
    o.set_state(0, c["pos"], None)
    a = 1.0 + c["miu"] / c["lmbd"]
    rest = 0.5 * c["restVol"].astype(np.float64) * c["lmbd"] * (1.0 - a) ** 2
    per_tet = GOLD["nh_tet_energy"].astype(np.float64) + rest
    assert per_tet.min() > 0 and (np.abs(GOLD["nh_tet_energy"]) > 1e-3 * rest).any()          # deformed: not just the rest value
    infos, nbr, _ = c["infos"], c["nbrTets"], c["nbrOrder"]
    for v in range(len(c["pos"])):
        want = per_tet[nbr[infos[2 * v]: infos[2 * v] + infos[2 * v + 1]]].sum()
        inertia, elastic = o.vertex_merit_energy(0, v, c["pos"][v])
        assert abs(elastic - want) <= 1e-5 * want, (v, elastic, want)
        # moving the vertex changes the energy the way the per-vertex force says (first order), so the two are consistent
    v = 5
    F, _ = o.vertex_material_force_hessian(0, v)
    d = np.array([1e-3, -2e-3, 1.5e-3], np.float32)
    e0 = o.vertex_merit_energy(0, v, c["pos"][v])[1]; e1 = o.vertex_merit_energy(0, v, c["pos"][v] + d)[1]; e2 = o.vertex_merit_energy(0, v, c["pos"][v] - d)[1]
    fd = float(F.astype(np.float64) @ d)
    assert abs((e1 - e2) / 2 + fd) < 2e-2 * abs(fd) + 1e-6 * abs(e0)


def test_solve3x3_bit_exact():
    lib = load_oracle()
    Ms, Bs = solve3x3_cases()
    for i in range(len(Ms)):
        x = np.zeros(3, np.float32)
        ok = lib.oracle_solve3x3(fp(Ms[i]), fp(Bs[i]), fp(x))
        assert ok == GOLD["solve_ok"][i]
        assert np.array_equal(x, GOLD["solve_x"][i], equal_nan=True)
    assert 0 < GOLD["solve_ok"].sum() < len(Ms)       # both branches covered


def test_cubic_roots_bit_exact():
    lib = load_oracle()
    coefs = cubic_cases()
    for i in range(len(coefs)):
        r = np.zeros(3, np.float32)
        n = lib.oracle_cubic_roots(fp(coefs[i]), 0.0, 1.0, fp(r))
        assert n == GOLD["cubic_n"][i]
        assert np.array_equal(r[:n], GOLD["cubic_roots"][i][:n])
    assert set(GOLD["cubic_n"].tolist()) >= {0, 1, 2}


def test_ccd_tri_point_bit_exact():
    lib = load_oracle()
    x0, x1, p0, p1 = ccd_cases()
    for i in range(len(x0)):
        t = C.c_float(-1); b = np.zeros(3, np.float32)
        hit = lib.oracle_ccd_tri_point(fp(x0[i]), fp(x1[i]), fp(p0[i]), fp(p1[i]), C.byref(t), fp(b))
        assert hit == GOLD["ccd_hit"][i]
        if hit:
            assert t.value == GOLD["ccd_t"][i]
            assert np.array_equal(b, GOLD["ccd_bary"][i])
    assert 20 < GOLD["ccd_hit"].sum() < len(x0) - 20


def test_contact_force_hessian():
    lib = load_oracle()
    cc = contact_cases()
    nz = 0
    for i in range(len(cc["p"])):
        for order in range(4):
            f = np.zeros(3, np.float32); h = np.zeros(9, np.float32); d = C.c_float()
            lib.oracle_contact_force_hessian(fp(cc["p"][i]), fp(cc["pPrev"][i]), fp(cc["tri"][i]), fp(cc["triPrev"][i]), order, cc["k"],
                                             cc["mu0"], cc["mu1"], cc["eps0"], cc["eps1"], cc["dt"], fp(f), fp(h), C.byref(d))
            gf, gh = GOLD["contact_force"][i, order], GOLD["contact_hessian"][i, order]
            assert (np.abs(gf).max() > 0) == (np.abs(f).max() > 0)
            if np.abs(gf).max() > 0:
                nz += 1
                assert relmax(f, gf) < 2e-5 and relmax(h, gh) < 2e-5
    assert nz > 100


def test_contact_force_hessian_cpu_path():
    """The same probe against the reference's own CPU contact path - VBDPhysics::updateCollisionInfo and
    ::accumlateCollisionForceAndHessianPerCollision with computeVertexFriction, closestPointTriangle and computeFaceNormal
    (VBDPhysics.cpp:2205-2311, 3005-3047, 2563-2592; oracle/ref_shim/refcontactcpu.cpp). Penetration depth, force and Hessian of all
    four vertex orders; which cases carry a force must agree exactly, the depth bit for bit, force / Hessian to 1e-6 of the largest entry."""
    lib = load_oracle()
    cc = contact_cases()
    G = GOLD["contact_cpu"]
    got = np.zeros_like(G)
    for i in range(len(cc["p"])):
        for order in range(4):
            f = np.zeros(3, np.float32); h = np.zeros(9, np.float32); d = C.c_float()
            lib.oracle_contact_force_hessian(fp(cc["p"][i]), fp(cc["pPrev"][i]), fp(cc["tri"][i]), fp(cc["triPrev"][i]), order, cc["k"],
                                             cc["mu0"], cc["mu1"], cc["eps0"], cc["eps1"], cc["dt"], fp(f), fp(h), C.byref(d))
            got[i, order] = np.concatenate([[d.value], f, h])
    assert np.array_equal(np.abs(got[..., 1:4]).max(-1) > 0, np.abs(G[..., 1:4]).max(-1) > 0)
    assert (np.abs(G[..., 1:4]).max(-1) > 0).sum() > 100 and (G[:, 3, 0] <= 0).sum() >= 10
    print("contact cpu:", relmax(got[..., 0], G[..., 0]), relmax(got[..., 1:4], G[..., 1:4]), relmax(got[..., 4:], G[..., 4:]))
    assert np.array_equal(got[..., 0], G[..., 0])                     # penetration depth: bit for bit
    for i in range(len(cc["p"])):
        for order in range(4):
            if np.abs(G[i, order, 1:4]).max() > 0:
                assert relmax(got[i, order, 1:4], G[i, order, 1:4]) < 1e-6 and relmax(got[i, order, 4:], G[i, order, 4:]) < 1e-6


def test_boundary_force_hessian():
    lib = load_oracle()
    bc = boundary_cases()
    nz = 0
    for i in range(len(bc["p"])):
        f = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
        lib.oracle_boundary_force_hessian(fp(bc["p"][i]), fp(bc["pPrev"][i]), fp(bc["mn"]), fp(bc["mx"]), bc["kB"], bc["muB"], bc["epsVB"],
                                          bc["dt"], 1, fp(f), fp(h))
        gf, gh = GOLD["boundary_force"][i], GOLD["boundary_hessian"][i]
        if np.abs(gf).max() > 0:
            nz += 1
            assert relmax(f, gf) < 1e-6 and relmax(h, gh) < 1e-6
        else:
            assert np.abs(f).max() == 0
    assert nz >= 30


def test_boundary_force_hessian_cpu_semantics():
    """The rule the product implements - penalty AND friction for every violated axis - against the reference's own CPU function
    VBDPhysics::accumlateBoundaryForceAndHessian with computeVertexFriction (VBDPhysics.cpp:2821-2954, 2563-2592; oracle/ref_shim/
    refboundarycpu.cpp): 120 points with 0-3 violated axes, friction off and on. 1e-6 relative to the largest entry."""
    lib = load_oracle()
    bc = boundary_cases_cpu()
    G = GOLD["boundary_cpu"]
    multi = 0
    for fr in (0, 1):
        got = np.zeros_like(G[fr])
        for i in range(len(bc["p"])):
            f = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
            lib.oracle_boundary_force_hessian(fp(bc["p"][i]), fp(bc["pPrev"][i]), fp(bc["mn"]), fp(bc["mx"]), bc["kB"], bc["muB"], bc["epsVB"],
                                              bc["dt"], fr, fp(f), fp(h))
            got[i] = np.concatenate([f, h])
            multi += int((np.abs(G[fr, i, :3]) > 0).sum() >= 2)
        assert relmax(got[:, :3], G[fr][:, :3]) < 1e-6 and relmax(got[:, 3:], G[fr][:, 3:]) < 1e-6
        assert np.array_equal(got == 0, G[fr] == 0)
    assert multi > 60 and not np.array_equal(G[0], G[1])


def test_point_in_tet_bit_exact():
    """oracle point_in_tet == GAIA::pointInTet (CollisionGeometry.h:56-88) compiled from the reference: same accept / reject
    decision on every case, including nearly flat tets around the 1e-5 volume bound and points on faces, edges and vertices"""
    lib = load_oracle()
    p, tets = point_in_tet_cases()
    got = np.array([lib.oracle_point_in_tet(fp(p[i]), fp(tets[i, 0]), fp(tets[i, 1]), fp(tets[i, 2]), fp(tets[i, 3])) for i in range(len(p))], np.int32)
    assert np.array_equal(got, GOLD["pit_inside"])
    assert 40 < int(GOLD["pit_inside"].sum()) < len(p) - 40          # both answers are well represented


def test_closest_point_triangle_bit_exact():
    """oracle closest_point_triangle == GAIA::closestPointTriangle (CollisionGeometry.h:90-157): region type, closest point
    and barycentrics bit for bit over all seven regions"""
    lib = load_oracle()
    p, tri = closest_point_triangle_cases()
    for i in range(len(p)):
        c3 = np.zeros(3, np.float32); b3 = np.zeros(3, np.float32)
        k = lib.oracle_closest_point_triangle(fp(p[i]), fp(tri[i, 0]), fp(tri[i, 1]), fp(tri[i, 2]), fp(c3), fp(b3))
        assert k == GOLD["cpt_type"][i], i
        assert np.array_equal(c3, GOLD["cpt_point"][i]) and np.array_equal(b3, GOLD["cpt_bary"][i]), i
    assert len(set(GOLD["cpt_type"].tolist())) == 7


def test_closest_points_between_segments_bit_exact():
    """oracle closest_points_segments (cloth edge-edge contacts) == GAIA::get_closest_points_between_segments
    (CollisionGeometry.h:160-267), incl. parallel, touching and zero-length segments"""
    from oracle.binding_cloth import _lib as cloth_lib
    lib = cloth_lib()
    seg = segment_cases()
    for i in range(len(seg)):
        o8 = np.zeros(8, np.float32)
        lib.oracle_cloth_closest_points_segments(fp(seg[i, 0]), fp(seg[i, 1]), fp(seg[i, 2]), fp(seg[i, 3]), fp(o8))
        assert np.array_equal(o8, GOLD["seg_out"][i], equal_nan=True), (i, o8, GOLD["seg_out"][i])


def test_cloth_contact_repulsion_bit_exact():
    """oracle cloth_contact_repulsion == VBDClothSimulationFramework::computeContactRepulsiveForce (VBDClothPhysics.cpp:1533-1579)
    compiled from the reference: dE/dd and d2E/dd2 for the quadratic and the two-stage quadratic / logarithmic energies bit for
    bit; the pure logarithmic type (2) calls log(): identical to 2 ulp (libm vs the build's vectorised log)."""
    from gaia_b200.cloth import default_cloth_physics_params
    from oracle.binding_cloth import OracleClothScene
    dis, th, r, k = cloth_repulsion_cases()
    for et in range(3):
        for i in range(0, len(dis)):
            o = OracleClothScene(default_cloth_physics_params(thickness=float(th[i]), contactRadius=float(r[i]), contactStiffness=float(k[i]),
                                                              contactEnergyTypes=et), threads=1)
            a, b = o.contact_repulsion(float(dis[i]))
            o.close()
            g = GOLD["cloth_repulsion"][et, i]
            if et < 2:
                assert np.float32(a) == g[0] and np.float32(b) == g[1], (et, i, a, b, g)
            else:
                assert np.allclose([a, b], g, rtol=3e-7, atol=0, equal_nan=True), (et, i, a, b, g)


@pytest.mark.skipif(not ref_available(), reason="oracle/_ref/libgaia_ref.so is only built where /root/reference exists")
def test_cloth_material_force_hessian_and_bending_forms():
    """The cloth oracle's membrane (StVK) + bending force / Hessian per vertex, with both damping terms, and its quadratic
    bending forms, against the reference's own lines compiled into oracle/_ref (VBDTriMeshStVk::accumlateMaterialForceAndHessian
    = ::accumlateStVKForceAndHessian + ::accumlateBendingForceAndHessian, VBDTriMeshStVK.cpp:715-982; ::intializeEdgeQuadraticForms,
    :49-88; assembleMembraneHessian, VBDBaseTriMesh.h:149-160 - compiled against a stand-in <Eigen/Dense>, oracle/ref_shim/refcloth.cpp).
    Forms: bit-exact. Force / Hessian: 2e-6 relative to the largest entry (sums over 6 faces and up to 12 hinges in another order)."""
    c = cloth_material_case()
    o = c["scene"]
    assert np.array_equal(o.bending_forms(0).reshape(-1, 16), GOLD["cloth_Q"]) and np.abs(GOLD["cloth_Q"]).max() > 10
    n = len(c["pos"])
    F = np.zeros((n, 3), np.float32); H = np.zeros((n, 9), np.float32)
    for v in range(n):
        f, h = o.material_force_hessian(0, v)
        F[v] = f; H[v] = h.T.reshape(9)
    assert relmax(F, GOLD["cloth_material_force"]) < 2e-6
    assert relmax(H, GOLD["cloth_material_hessian"]) < 2e-6
    # both parts and both damping terms are live in this fixture
    assert np.abs(GOLD["cloth_bending_force"]).max() > 0.05 * np.abs(GOLD["cloth_stvk_force"]).max()
    assert np.abs(GOLD["cloth_stvk_force"] + GOLD["cloth_bending_force"] - GOLD["cloth_material_force"]).max() < 1e-5 * np.abs(GOLD["cloth_material_force"]).max()


@pytest.mark.parametrize("energy_type", [0, 1, 2])
def test_cloth_contact_force_hessian(energy_type):
    """One vertex-face / edge-edge contact at a time, all four vertex orders, friction on, against the reference's own lines
    compiled into oracle/_ref (VBDClothSimulationFramework::accumulateVFContactForceAndHessian, ::accumulateEEContactForceAndHessian
    with ::computeFriction and ::computeContactRepulsiveForce, VBDClothPhysics.cpp:1248-1399, 1500-1579; oracle/ref_shim/
    refclothcontact.cpp; closest points from the reference's own closestPointTriangle / get_closest_points_between_segments).
    Distances bit-exact; force / Hessian 1e-6 relative to the largest entry; NaNs (log energy inside the thickness) in the same
    places."""
    c = cloth_contact_case(energy_type)
    o = c["scene"]
    for key, pairs, probe in (("cloth_vf", c["vf"], o.probe_vf), ("cloth_ee", c["ee"], o.probe_ee)):
        G = GOLD[f"{key}_e{energy_type}"]
        got = np.zeros_like(G)
        for i, pr in enumerate(pairs):
            for order in range(4):
                d, f, h = probe(*pr, order)
                got[i, order] = np.concatenate([[d], f, h.T.reshape(9)])
        assert np.array_equal(got[:, :, 0], G[:, :, 0])
        assert np.array_equal(np.isnan(got), np.isnan(G))
        ok = ~np.isnan(G)
        active = (np.abs(np.nan_to_num(G[:, :, 1:4])).max((1, 2)) > 0).sum()
        assert active >= 25, active                                         # most pairs are inside the contact radius
        for sl in (slice(1, 4), slice(4, 13)):
            a, b = np.where(ok, got, 0)[:, :, sl], np.where(ok, G, 0)[:, :, sl]
            assert relmax(a, b) < 1e-6


def test_cloth_boundary_force_hessian():
    """::accumlateBoundaryForceAndHessian (VBDClothPhysics.cpp:1401-1463): vertices outside the world box on up to two axes."""
    c = cloth_contact_case(1)
    o = c["scene"]
    G = GOLD["cloth_boundary"]
    got = []
    for mi, m in enumerate(c["meshes"]):
        for v in range(len(m["pos"])):
            f, h = o.probe_boundary(mi, v)
            got.append(np.concatenate([f, h.T.reshape(9)]))
    got = np.asarray(got, np.float32)
    assert (np.abs(G[:, :3]).max(1) > 0).sum() > 20
    assert relmax(got[:, :3], G[:, :3]) < 1e-6 and relmax(got[:, 3:], G[:, 3:]) < 1e-6


def test_tetrahedral_traversal_integer_exact():
    """The oracle's tetrahedral_traverse_to against the reference's own TetMeshFEM::tetrahedralTraverseTo (TetMeshFEM.cpp:503-742 with
    exitFaceSelection / projectTo2DCoordinates / originBarycentric2DTriangle from TetMeshFEM.h, MeshFrame2's CPArrayStatic / CircularArray
    and CuMatrix::buildOrthonormalBasis, compiled into oracle/_ref: oracle/ref_shim/reftraverse.cu) on 800 rays through the
    reference's teapot and a Kuhn beam: same answer (target reached or not) and the same number of tets traversed, ray by ray."""
    for c in traverse_cases():
        G = GOLD["traverse_" + c["name"]]
        o = c["scene"]
        assert G[:, 0].sum() > 100 and (G[:, 2] == 3).sum() > 50 and (G[:, 2] == 2).sum() > 20 and G[:, 1].max() > 20
        assert (G[:, 2] != 1).all()                   # no static-stack overflow among the fixtures (the oracle's stack is unbounded)
        for i, (org, d, md, st, sf, tg) in enumerate(c["rays"]):
            ok, n = o.tet_traverse(0, org, d, md, st, sf, tg, c["eps"])
            assert ok == bool(G[i, 0]) and n == G[i, 1], (c["name"], i, ok, n, G[i])


def test_feasible_region_verdicts_exact():
    """The oracle's check_feasible_region against the reference's own DiscreteCollisionDetector::checkFeasibleRegion with
    checkEdgeFeasibleRegion / checkVertexFeasibleRegion (DiscreteCollisionDetector.cpp:664-700, 709-815, compiled with the
    vendored Embree math headers: oracle/ref_shim/reffeasible.cpp) on 3000 points around the reference's teapot and a Kuhn beam:
    every closest-point type, both verdicts, identical point by point."""
    for c in feasible_region_cases():
        G = GOLD["feasible_" + c["name"]]
        o = c["scene"]
        for t in range(7):
            assert (G[:, 0] == t).sum() > 50 and 0 < G[G[:, 0] == t, 1].sum()
        assert (G[G[:, 0] != 6, 1] == 0).sum() > 500                      # many rejections outside the interior type
        for i, (f, q) in enumerate(c["pts"]):
            assert o.check_feasible_region(0, q, f, int(G[i, 0]), c["eps"]) == bool(G[i, 1]), (c["name"], i)


def test_cloth_feasible_region_verdicts_exact():
    """The cloth oracle's feasible_region against the reference's own checkFeasibleRegion (TriMeshFEM) with checkEdgeFeasibleRegion /
    checkVertexFeasibleRegion (TriMeshCollisionGeometry.cpp:7-42, 91-186; oracle/ref_shim/refclothfeasible.cpp) on 1500 points
    around a wrinkled open cloth: every closest-point type incl. boundary edges and corners, verdicts identical."""
    c = cloth_feasible_region_cases()
    G = GOLD["cloth_feasible"]
    o = c["scene"]
    for t in range(7):
        assert (G[:, 0] == t).sum() > 50 and 0 < G[G[:, 0] == t, 1].sum()
    assert (G[G[:, 0] != 6, 1] == 0).sum() > 500
    for i, (f, q) in enumerate(c["pts"]):
        assert o.check_feasible_region(0, q, f, int(G[i, 0]), c["eps"]) == bool(G[i, 1]), i


def test_initial_step_and_velocity_update_bit_exact():
    """apply_initial_step / update_velocities of the oracle against the reference's own VBDBaseTetMesh::applyInitialStep
    (computeInertia + forwardStep, VBD_BaseMaterial.h:201-221, 244-259; .cpp:7-132) and VBDPhysics::updateVelocities with
    dampVelocities (VBDPhysics.cpp:2486-2561), compiled into oracle/_ref (oracle/ref_shim/refstep.cpp): four substeps for each
    initial-guess type 0-6 (5 = forwardStepSimplecticEuler, .cpp:134-147), with exponential / constant damping, the velocity cap in
    and out of the no-gravity zone and fixed points. Elementwise float arithmetic: bit-exact (initial guess, inertia, velocities)."""
    c = step_sequence_cases()
    assert [t for t, _ in c["cases"]] == [0, 1, 2, 3, 4, 5, 6]
    for init_type, kw in c["cases"]:
        o, _ = c["make"](init_type, kw)
        forces = []
        init, inert, vel, solved = run_oracle_step_sequence(o, c, forces)
        G = GOLD[f"step_init{init_type}"]
        if init_type == 5:
            # forwardStepSimplecticEuler (VBD_BaseMaterial.cpp:134-147) was given the oracle's elastic forces (pinned in
            # test_elastic_force_alone_equals_the_force_half...) and 1 / mass: the same forces here, and a guess that they moved
            assert np.array_equal(np.array(forces, np.float32), GOLD["step_forces5"])
            assert np.abs(init[0] - c["pos0"]).max() > 1e-3 and not np.array_equal(init[0][c["fixed"]], c["pos0"][c["fixed"]])   # fixed vertices move too, as there
        assert np.array_equal(solved, GOLD[f"step_solved{init_type}"])
        assert np.array_equal(init, G[0]) and np.array_equal(inert, G[1]) and np.array_equal(vel, G[2]), (init_type, kw)
        assert np.abs(G[2]).max() > 1.0 and np.abs(G[0][-1] - c["pos0"]).max() > 1e-3
    # the guesses really differ between the types (the history-dependent ones from the second substep on)
    assert not np.array_equal(GOLD["step_init0"][0][1], GOLD["step_init2"][0][1]) and not np.array_equal(GOLD["step_init4"][0][1], GOLD["step_init6"][0][1])


def test_cloth_initial_step_velocity_update_and_truncation_bit_exact():
    """The cloth oracle's apply_initial_step / update_velocity / truncate_to_bound against the reference's own
    VBDBaseTriMesh::applyInitialStep + ::forwardStep_intertia (VBDBaseTriMesh.cpp:58-178), VBDClothSimulationFramework::updateVelocity
    (VBDClothPhysics.cpp:2318-2359) and ::applyConservativeBoundTruncation (:1223-1246), compiled into oracle/_ref
    (oracle/ref_shim/refclothstep.cpp): four substeps per initial-guess type 0-3 with damping, velocity cap, fixed points and
    conservative bounds, and 200 truncation candidates. Elementwise float arithmetic: bit-exact."""
    c = cloth_step_sequence_cases()
    for init_type, kw in c["cases"]:
        o, _ = c["make"](init_type, kw)
        init, inert, vel, solved = run_cloth_oracle_step_sequence(o, c)
        G = GOLD[f"cloth_step_init{init_type}"]
        assert np.array_equal(solved, GOLD[f"cloth_step_solved{init_type}"])
        assert np.array_equal(init, G[0]) and np.array_equal(inert, G[1]) and np.array_equal(vel, G[2]), (init_type, kw)
    assert not np.array_equal(GOLD["cloth_step_init0"][0][1], GOLD["cloth_step_init3"][0][1])      # the bound clips type 3's guess
    o, _ = c["make"](0, {})
    o.set_state(0, c["pos0"], None); o.set_bounds(0, c["bounds"], c["pos0"])
    flags = GOLD["cloth_truncate_flag"]
    assert 50 < flags.sum() < 190
    for i, (v, p) in enumerate(zip(c["trunc_verts"], c["trunc_pos"])):
        t, q = o.probe_truncate(0, int(v), p)
        assert t == bool(flags[i]) and np.array_equal(q, GOLD["cloth_truncate_pos"][i]), i


def test_golden_file_matches_a_fresh_run_of_the_reference():
    """The committed fixture is what the reference's code produces today (guards against a stale .npz)."""
    ref = load_ref()
    Ms, Bs = solve3x3_cases()
    for i in range(0, len(Ms), 7):
        x = np.zeros(3, np.float32)
        ok = ref.ref_solve3x3(fp(Ms[i]), fp(Bs[i]), fp(x))
        assert ok == GOLD["solve_ok"][i] and np.array_equal(x, GOLD["solve_x"][i], equal_nan=True)
    x0, x1, p0, p1 = ccd_cases()
    for i in range(0, len(x0), 5):
        t = C.c_float(-1); b = np.zeros(3, np.float32)
        hit = ref.ref_ccd_tri_point(fp(x0[i]), fp(x1[i]), fp(p0[i]), fp(p1[i]), C.byref(t), fp(b))
        assert hit == GOLD["ccd_hit"][i] and (not hit or t.value == GOLD["ccd_t"][i])


def test_cloth_sewing_and_attachment_terms():
    """Sewing and vertex-face attachments of the cloth path against the reference's own lines (VBDBaseTriMesh::
    accumlateSewingForceAndHessian, VBDBaseTriMesh.cpp:320-344; the attachment set-up loop and ::accumulateConstraintForceAndHessian,
    VBDClothPhysics.cpp:161-220, 1464-1498; oracle/ref_shim/refclothconstraint.cpp): every vertex of both simulated cloths, seams with
    ratios 0 and 1, attachments to a fixed point / the other cloth / a collider / the own cloth, a zero barycentric coordinate,
    given and default stiffness. Which vertices carry a term must agree exactly; force / Hessian 1e-6 of the largest entry."""
    from golden_inputs import build_cloth_constraint_scene, cloth_constraint_case
    from oracle.binding_cloth import OracleClothScene
    c = cloth_constraint_case()
    n0 = len(c["meshes"][0][0]); n1 = len(c["meshes"][1][0])
    for tag, with_k in (("", True), ("_default_k", False)):
        o = build_cloth_constraint_scene(OracleClothScene, c, with_stiffness=with_k, threads=1)
        sw = np.stack([np.concatenate(o.sewing_force_hessian(0, v)) for v in range(n0)])
        at = np.stack([np.concatenate(o.constraint_force_hessian(m, v)) for m, n in ((0, n0), (1, n1)) for v in range(n)])
        G, A = GOLD["cloth_sewing"], GOLD["cloth_attach" + tag]
        assert np.array_equal(np.abs(sw).max(1) > 0, np.abs(G).max(1) > 0) and (np.abs(G).max(1) > 0).sum() == 14
        assert np.array_equal(np.abs(at).max(1) > 0, GOLD["cloth_attach_count" + tag] > 0) and (GOLD["cloth_attach_count" + tag] > 0).sum() == 13
        assert relmax(sw[:, :3], G[:, :3]) < 1e-6 and relmax(sw[:, 3:], G[:, 3:]) < 1e-6
        assert relmax(at[:, :3], A[:, :3]) < 1e-6 and relmax(at[:, 3:], A[:, 3:]) < 1e-6
        # both terms are part of the vertex's total: the total moves by exactly these amounts when they are switched off
        f_all = np.stack([np.concatenate([o.vertex_force_hessian(0, v)[0], o.vertex_force_hessian(0, v)[1].T.reshape(-1)]) for v in range(n0)])
        plain = build_cloth_constraint_scene(OracleClothScene, dict(c, seams=np.zeros((0, 3), np.int32), ratios=np.zeros(0, np.float32),
                                                                    v_infos=c["v_infos"][:0], f_infos=c["f_infos"][:0]), threads=1)
        f_plain = np.stack([np.concatenate([plain.vertex_force_hessian(0, v)[0], plain.vertex_force_hessian(0, v)[1].T.reshape(-1)]) for v in range(n0)])
        assert relmax(f_all - f_plain, sw + at[:n0]) < 1e-5


def test_elastic_force_alone_equals_the_force_half_of_the_pinned_routine():
    """neohookean_force (VBDTetMeshNeoHookean::accumlateMaterialForce, VBD_NeoHookean.cpp:781-870: what initial-guess type 5 sums)
    is the force half of neohookean_force_hessian with the damping off - the routine pinned to the reference's compiled code in
    test_neohookean_force_hessian_plain: same F, cofactors and weights, summed in the same order, so bit for bit."""
    c = neohookean_case(seed=11, dampH=0.0, dampD=0.0)
    o = c["scene"]
    o.set_state(0, c["pos"], None); o.set_prev_positions(0, c["prev"])
    f = o.neohookean_forces(0)
    ref = np.stack([o.vertex_material_force_hessian(0, v)[0] for v in range(len(c["pos"]))])
    assert np.abs(f).max() > 1.0 and np.array_equal(f, ref)
    assert relmax(f, GOLD["nh_plain_force"]) < 2e-5


def test_bowl_ground_force_hessian():
    """useBowlGround - the inside of a sphere as the ground - against the reference's own accumlateBoundaryForceAndHessian
    (VBDPhysics.cpp:2824-2864 of the function compiled in oracle/ref_shim/refboundarycpu.cpp): 160 points inside, just outside and far
    outside the sphere, below and above its centre, with and without the cap and friction. Zeros in the same places; 2e-6 of the
    largest entry (the sphere normal goes through a square root and a division)."""
    from golden_inputs import bowl_cases
    lib = load_oracle()
    lib.oracle_boundary_bowl_force_hessian.argtypes = [FP, FP, FP, C.c_float, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, FP, FP]
    bw = bowl_cases()
    G = GOLD["boundary_bowl_cpu"]
    for cap in (0, 1):
        for fr in (0, 1):
            got = np.zeros_like(G[cap, fr])
            for i in range(len(bw["p"])):
                f = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
                lib.oracle_boundary_bowl_force_hessian(fp(bw["p"][i]), fp(bw["pPrev"][i]), fp(bw["center"]), bw["radius"], cap, bw["kB"], bw["muB"], bw["epsVB"],
                                                       bw["dt"], fr, fp(f), fp(h))
                got[i] = np.concatenate([f, h])
            hit = np.abs(G[cap, fr]).max(1) > 0
            assert np.array_equal(np.abs(got).max(1) > 0, hit)
            assert relmax(got[:, :3], G[cap, fr][:, :3]) < 2e-6 and relmax(got[:, 3:], G[cap, fr][:, 3:]) < 2e-6
    n_open, n_cap = (np.abs(G[0, 1]).max(1) > 0).sum(), (np.abs(G[1, 1]).max(1) > 0).sum()
    assert 30 < n_open < n_cap < len(bw["p"])                  # the cap adds the points above the centre
    assert not np.array_equal(G[1, 0], G[1, 1])                 # friction matters


def test_cloth_conservative_bounds_fold_bit_exact():
    """The cloth oracle's conservative bounds and active-collision mask against the reference's own loop (the end of
    VBDClothSimulationFramework::applyCollisionDetection, VBDClothPhysics.cpp:775-819, compiled in oracle/ref_shim/refclothbounds.cpp):
    the same query leftovers - smallest distance and number of contacts per vertex / face / edge, some closer than the
    thickness - folded over each vertex's incident faces and edges; a mask that is only ever raised. Minima, one subtraction,
    one product: bit for bit. (The incidence lists are the oracle's own; a numpy recomputation checks them here.)"""
    from golden_inputs import cloth_bounds_case
    c = cloth_bounds_case()
    o = c["scene"]
    nV, nF, nE = c["sizes"]
    f_off, f_adj, e_off, e_adj = o.incidence(0)
    faces = o.topology(0)[0]; edges = o.edges(0)
    for v in range(nV):          # incidence = exactly the faces / edges that contain the vertex
        fs = sorted(f_adj[f_off[v]:f_off[v + 1]].tolist()); es = sorted(e_adj[e_off[v]:e_off[v + 1]].tolist())
        assert es == sorted(np.nonzero((edges[:, 0] == v) | (edges[:, 1] == v))[0].tolist())
        assert len(fs) == f_off[v + 1] - f_off[v] and len(set(fs)) == len(fs)
    b, m = o.probe_bounds(0, c["vf_min"], c["vf_cnt"], c["face_min"], c["face_cnt"], c["ee_min"], c["ee_cnt"], c["mask"])
    assert np.array_equal(b, GOLD["cloth_bounds"]) and np.array_equal(m, GOLD["cloth_bounds_mask"])
    assert (b == 0).sum() >= 3 and (b > 0).sum() > 20 and len(np.unique(b)) > 15 and b.max() <= np.float32(c["relax"]) * np.float32(c["R"] - c["thick"]) * 1.000001
    assert (m >= c["mask"]).all() and m.sum() > c["mask"].sum() and (m == 0).sum() > 5


def test_step_schedule_is_the_reference_loop():
    """WHICH operations the step runs, in which order, on which (mesh, vertex), with which flags: the oracle's run_step traced
    event by event against the reference's own loop and detection passes - VBDPhysics::runStep_hybridCollisionHandling
    (VBDPhysics.cpp:1634-1707) with ::dcd, ::intermediateDCD, ::ccd, ::intermediateCCD, ::intermediateCollisionDetection (:1958-2172),
    compiled against stand-in members and detectors that record their calls (oracle/ref_shim/refsteploop.cpp). Three frames of a
    three-block scene with fixed vertices, contacts, a block that appears in frame 1, friction from iteration 2 on, intermediate
    detection every 3 iterations and tree rebuilds every 2 / 1 / 2 frames; one frame each with DCD alone (the intermediate pass
    then re-queries every vertex) and CCD alone (the results are cleared instead). Covered: when the trees are rebuilt, when the
    penetrated mask is cleared, which surface vertices each pass queries (DCD: all; intermediate DCD: the penetrated ones; CCD: the
    others), closest-point queries after hits only, contact updates before the solves of a group, every vertex solve with its
    friction flag. The outcome of a DCD query and the active lists are data the stand-in repeats from the oracle's recording."""
    from golden_inputs import step_schedule_case
    for tag, dcd, ccd, frames in (("", 1, 1, 3), ("_dcd", 1, 0, 1), ("_ccd", 0, 1, 1)):
        c = step_schedule_case(dcd, ccd)
        o = c["scene"]
        seen = set()
        for frame in range(frames):
            o.trace_begin(); o.run_frames(1)
            tr = o.trace_read()
            G = GOLD[f"step_trace{tag}_f{frame}"]
            assert tr.shape == G.shape and np.array_equal(tr, G), (tag, frame)
            seen |= set(G[:, 0].tolist())
            assert (G[:, 0] == 1).sum() == 2 and (G[:, 0] == 9).sum() == 2 and (G[:, 0] == 8).sum() == 2 * 2          # substeps; 6 // 3 intermediate detections each
            assert (G[:, 0] == 2).sum() == 2 * int(c["active"][frame].sum())
            solves = G[G[:, 0] == 7]
            assert set(solves[:, 3].tolist()) == {0, 1} and (2 in solves[:, 1]) == (frame >= 1)                       # friction switches on; block 2 only once it has appeared
            n_free = sum(int((c["mask"][c["fixed_off"][m]:c["fixed_off"][m + 1]] == 0).sum()) for m in range(3) if c["active"][frame][m])
            assert len(solves) == 2 * 6 * n_free                                                                       # every free vertex of every active mesh, every iteration
            if dcd and ccd:
                rebuilds = G[G[:, 0] == 12][:, 1:3].tolist()
                assert rebuilds[0] == [int(frame == 2), int(frame >= 1)] and all(r == [0, 0] for r in rebuilds[1:])     # first substep of a rebuild frame only
                assert G[G[:, 0] == 15][:, 1].tolist()[0] == int(frame == 2) and G[G[:, 0] == 15][:, 1].sum() == int(frame == 2)
        if dcd and ccd:
            assert seen == {1, 2, 3, 4, 5, 6, 7, 8, 9, 11, 12, 13, 14, 15, 16, 18}
        elif dcd:
            assert seen == {1, 2, 3, 4, 5, 6, 7, 8, 9, 11, 12, 13, 14, 18}
            assert (G[:, 0] == 13).sum() == 2 * 3 * int(np.diff(c["sv_off"])[:2].sum())       # the intermediate passes re-query every surface vertex of the active meshes
        else:
            assert 17 in seen and 13 not in seen and 16 in seen


def test_active_lists_and_relations_are_the_reference_construction():
    """Row a15 on the oracle's side: the active collision list of every parallel group, the face-side relations of every vertex
    and the active-collision mask, built from the same detection results by the reference's own code - VBDPhysics::
    prepareCollisionDataCPU (VBDPhysics.cpp:1709-1810) with ActiveCollisionList::clear / addToActiveCollisionList
    (ActiveCollisionList.cpp:6-110, .h:9-53) and MeshFrame2's CPArray, compiled in oracle/ref_shim/refprepare.cpp. Four scene
    states with DCD and CCD contacts; ORDER and all: entry for entry. (The product's lists are compared with the oracle's bit for
    bit, after the canonical sort the ABI returns them in, in tests/test_gpu_collision_parity.py.)"""
    from golden_inputs import prepared_collision_cases, prepared_collision_inputs
    total = 0
    for tag, o in prepared_collision_cases():
        q = prepared_collision_inputs(o)
        a_off, a_pairs = GOLD[f"prep_{tag}_aoff"], GOLD[f"prep_{tag}_apairs"]
        for g, mine in enumerate(q["active"]):
            assert np.array_equal(mine, a_pairs[a_off[g]:a_off[g + 1]]), (tag, g)
        r_off, rows, mask = GOLD[f"prep_{tag}_roff"], GOLD[f"prep_{tag}_rows"], GOLD[f"prep_{tag}_mask"]
        for m, (off, rw, mk, _) in enumerate(q["rel"]):
            b = q["v_off"][m]
            assert np.array_equal(off + r_off[b], r_off[b:b + len(off)]) and np.array_equal(rw, rows[r_off[b]:r_off[b] + len(rw)]), (tag, m)
            # the mask: the oracle follows the rule of the reference's GPU step, where the accelerator reads it (own contact with a
            # path, or any relation: VBDPhysicsCompute.cu:1740-1770). prepareCollisionDataCPU clears each mesh's mask when it COMES
            # to that mesh (:1719), wiping what earlier meshes' contacts marked on it - its mask is a subset, read by nothing on
            # this path. Both statements are checked against the reference's own output here.
            ref_mask = mask[b:b + len(mk)]
            own = np.zeros(len(mk), bool)
            sv = slice(q["sv_off"][m], q["sv_off"][m + 1])
            own[q["svids"][sv][(q["has"][sv] & q["path"][sv]) != 0]] = True
            gpu_rule = own | (np.diff(r_off[b:b + len(off)]) > 0)
            assert np.array_equal(mk != 0, gpu_rule) and not (ref_mask & ~mk.astype(bool)).any()
        assert len(rows) == 3 * int((q["has"] & q["path"]).sum()) and a_off[-1] >= int(q["has"].sum())
        total += int(q["has"].sum())
    assert total > 60


def test_vertex_step_tail_bit_exact():
    """What the vertex step does once its terms are summed - the force threshold, the 3x3 solve, the gradient fall-back with
    stepSizeGD when the solve fails, x += stepSize * direction - against the reference's own VBDPhysics::VBDStepWithCollision
    (VBDPhysics.cpp:2402-2484) with its own CuMatrix::solve3x3_psd_stable, compiled in oracle/ref_shim/refvertexstep.cu: 240 (force,
    Hessian, position) triples with forces around the threshold and singular / indefinite / zero Hessians. New positions bit for
    bit; all three outcomes occur. The same compile records the order in which that function sums its terms - material, inertia,
    world box, contacts - which is the order vbd_step_with_collision restates."""
    from golden_inputs import vertex_step_tail_cases
    lib = load_oracle()
    lib.oracle_vertex_step_tail.argtypes = [C.c_int, FP, FP, FP, C.c_float, C.c_float, C.POINTER(C.c_int32)]
    c = vertex_step_tail_cases()
    x = c["x"].copy(); flags = np.zeros(len(x), np.int32)
    lib.oracle_vertex_step_tail(len(x), fp(c["f"]), fp(c["h"]), fp(x), c["step"], c["step_gd"], flags.ctypes.data_as(C.POINTER(C.c_int32)))
    G = GOLD["vertex_step_tail_x"]
    assert np.array_equal(x, G)
    moved = (G != c["x"]).any(1)
    assert not moved[flags == 0].any() and moved[flags > 0].sum() > 150        # (a step below the ulp of the position leaves it unchanged)
    assert (flags == 0).sum() >= 10 and (flags == 1).sum() >= 100 and (flags == 2).sum() >= 10
    below = (c["f"].astype(np.float32) ** 2).sum(1, dtype=np.float32) <= np.float32(1e-5) * np.float32(1e-5)
    assert below.sum() >= 10 and np.array_equal(flags == 0, below)                # the threshold is |f|^2 > CMP_EPSILON^2, nothing else stops a step
    assert GOLD["vertex_step_order"].tolist() == [0, 1, 2, 3]


def test_inertia_term_bit_exact():
    """Row a4: m (y - x) / dt^2 and (m / dt^2) I added onto accumulators that already hold the material term, against the reference's
    own VBDBaseTetMesh::accumlateInertiaForceAndHessian (VBD_BaseMaterial.h:359-366, compiled in oracle/ref_shim/refvertexstep.cu):
    200 vertices, masses over five decades. Bit for bit."""
    from golden_inputs import inertia_cases
    lib = load_oracle()
    lib.oracle_inertia_terms.argtypes = [C.c_int, FP, FP, FP, C.c_float, FP, FP]
    c = inertia_cases()
    f = c["f0"].copy(); h = c["h0"].copy()
    lib.oracle_inertia_terms(len(f), fp(c["mass"]), fp(c["inertia"]), fp(c["x"]), c["dtsq"], fp(f), fp(h))
    assert np.array_equal(f, GOLD["inertia_f"]) and np.array_equal(h, GOLD["inertia_h"])
    assert not np.array_equal(f, c["f0"]) and np.array_equal(h[:, [1, 2, 3, 5, 6, 7]], c["h0"][:, [1, 2, 3, 5, 6, 7]]) and (h[:, 0] != c["h0"][:, 0]).any()


def test_chebyshev_weights_bit_exact():
    """Row a10: the accelerator's weights omega_1 .. omega_40 of the tet oracle against VBDPhysics::getAcceleratorOmega
    (VBDPhysics.cpp:3558-3574) and of the cloth oracle against ChebysevAccelerator::getAcceleratorOmega (M/SolverUtils/
    ChebysevAccelerator.h:6-38), both compiled in oracle/ref_shim/refaccel.cpp, for six values of rho (0 included). Bit for bit."""
    from oracle.binding_cloth import _lib as cloth_lib
    lib = load_oracle(); cl = cloth_lib()
    lib.oracle_accelerator_omegas.argtypes = [C.c_float, C.c_int, FP]; lib.oracle_accelerator_omegas.restype = None
    cl.oracle_cloth_accelerator_omegas.argtypes = [C.c_float, C.c_int, FP]; cl.oracle_cloth_accelerator_omegas.restype = None
    W = GOLD["accel_omegas"]
    for i, rho in enumerate(GOLD["accel_rhos"]):
        a = np.zeros(40, np.float32); b = np.zeros(40, np.float32)
        lib.oracle_accelerator_omegas(float(rho), 40, fp(a)); cl.oracle_cloth_accelerator_omegas(float(rho), 40, fp(b))
        assert np.array_equal(a, W[0, i]) and np.array_equal(b, W[1, i]), rho
    assert W[0, 0].max() == 1.0 and W[0, -1, -1] > 1.5 and np.array_equal(W[0], W[1])      # rho 0: no acceleration; the two recurrences agree
    # the position update y_{k-1} + omega (y_{k+1} - y_{k-1}): the body of VBDSolveParallelGroup_applyAcceleration_kernel (VBDPhysicsCompute.cu:932-935)
    lib.oracle_accelerate.argtypes = [C.c_int, FP, FP, C.c_float]; lib.oracle_accelerate.restype = None
    v = GOLD["accel_y"].copy()
    lib.oracle_accelerate(len(v), fp(v), fp(np.ascontiguousarray(GOLD["accel_ypp"])), 1.7321)
    assert np.array_equal(v, GOLD["accel_out"]) and not np.array_equal(v, GOLD["accel_y"])


def test_tet_deformers_bit_exact():
    """Row f3: the oracle's Rotator and Translator against the reference's own DeformerRotator / DeformerTranslater (M/VBD/
    VBDDeformer.h:45-179, compiled in oracle/ref_shim/refdeformer.cpp against the stand-in <Eigen/Dense>): the driven vertices of a
    beam after each of six frames, at the oracle's own clock (float accumulation of dt) - axis not normalised and off the centre,
    rotation and translation that stop inside the run. Bit for bit (double trigonometry from the same libm, cast as there)."""
    from golden_inputs import deformer_cases, run_oracle_deformer
    c = deformer_cases()
    for i, case in enumerate(c["cases"]):
        times, pos, _ = run_oracle_deformer(case, c)
        assert np.array_equal(times, GOLD[f"deformer{i}_times"])
        assert np.array_equal(pos, GOLD[f"deformer{i}_pos"]), i
        moved = np.abs(pos[-1] - c["verts"][case[1]]).max()
        assert moved > 0.005
        if case[5] > 0:
            assert np.array_equal(pos[-1], pos[-2]) and not np.array_equal(pos[0], pos[1])      # stopped at its end time


def test_cloth_deformers_bit_exact():
    """The cloth oracle's Rotator and Translator against the reference's own ClothDeformerRotator / ClothDeformerTranslater
    (M/VBDCloth/VBDClothDeformer.h:43-199, compiled in oracle/ref_shim/refclothdeformer.cpp), applied once per substep at the
    oracle's own clock: the driven edge of a cloth after each of five frames; a rotation that stops, a translation that runs only
    between its start and end times. Bit for bit."""
    from golden_inputs import cloth_deformer_cases, run_cloth_oracle_deformer
    c = cloth_deformer_cases()
    for i, case in enumerate(c["cases"]):
        _, _, pos = run_cloth_oracle_deformer(case, c)
        assert np.array_equal(pos, GOLD[f"cloth_deformer{i}_pos"]), i
        assert np.abs(pos[-1] - c["verts"][case[1]]).max() > 0.005
        if case[6] > 0:
            assert np.array_equal(pos[-1], pos[-2])      # stopped


def test_collider_key_frame_sequence_bit_exact():
    """The cloth oracle's key-frame collider against the reference's own ColliderTrimeshSequence::update / ::set_state (M/SpatialQuery/
    ColiiderTriMeshSequence.h:59-124, compiled in oracle/ref_shim/refcollider.cpp): every call of ten frames x 3 substeps x 4 iterations
    for three interpolation settings - which calls move the collider (the flag that forces a re-detection) and where to. Bit for bit."""
    from golden_inputs import collider_sequence_cases, run_cloth_oracle_collider
    c = collider_sequence_cases()
    for i, (isub, iit) in enumerate(c["settings"]):
        pos, upd = run_cloth_oracle_collider(c, isub, iit)
        assert np.array_equal(upd, GOLD[f"collider_seq{i}_upd"]), i
        assert np.array_equal(pos, GOLD[f"collider_seq{i}_pos"]), i
        assert upd[0] == 1 and 0 < upd.sum() < len(upd)
    assert GOLD["collider_seq0_upd"].sum() > GOLD["collider_seq1_upd"].sum() > GOLD["collider_seq2_upd"].sum()


def test_parallel_group_merge_is_the_reference_construction():
    """Row a1: the merge of every mesh's colour classes into the global parallel groups - each class of a mesh joins the currently
    smallest group that mesh has not used yet - against the reference's own findSmallestParallelGroup and the merge loop of
    VBDPhysics::initialize (VBDPhysics.cpp:99-122, 236-274, compiled in oracle/ref_shim/refgroups.cpp): five meshes with 2 to 5 classes
    of uneven sizes. Groups entry for entry, order included. (The size sort before it is the stated stable rule: the reference's
    std::sort leaves ties to the library.)"""
    from gaia_b200 import capi
    from golden_inputs import group_merge_case
    from oracle.binding import OracleScene
    o = OracleScene(capi.default_physics_params(), capi.default_collision_params(allowDCD=0, allowCCD=0), threads=1)
    for verts, tets, colors in group_merge_case():
        o.add_tet_mesh(verts, tets, capi.default_object_params(), colors)
    o.finalize()
    groups = o.parallel_groups()
    off, pairs = GOLD["group_merge_off"], GOLD["group_merge_pairs"]
    assert len(groups) == len(off) - 1 == 5
    for g, mine in enumerate(groups):
        assert np.array_equal(mine, pairs[off[g] // 2:off[g + 1] // 2]), g          # (offsets count ints: two per pair)
    sizes = np.diff(off)
    assert sizes.min() > 0 and len(set(pairs[:off[1] // 2, 0].tolist())) >= 4      # the groups mix several meshes


def test_cloth_parallel_group_merge_is_the_reference_construction():
    """The cloth oracle's parallel groups against the reference's own VBDClothSimulationFramework::initializeParallelGroups and
    ::findSmallestParallelGroup (VBDClothPhysics.cpp:72-138, compiled in oracle/ref_shim/refclothgroups.cpp): four cloths with 9, 6, 9
    and 4 classes of uneven sizes, taken in file order, a used group not retired (two classes of one mesh may share a group). Groups
    entry for entry, order included."""
    from gaia_b200 import cloth
    from golden_inputs import cloth_group_merge_case
    from oracle.binding_cloth import OracleClothScene
    o = OracleClothScene(cloth.default_cloth_physics_params(handleCollision=0), threads=1)
    for v, f, cols in cloth_group_merge_case():
        o.add_tri_mesh(v, f, cloth.default_cloth_object_params(), cols)
    o.finalize()
    groups = o.parallel_groups()
    off, pairs = GOLD["cloth_group_merge_off"], GOLD["cloth_group_merge_pairs"]
    assert len(groups) == len(off) - 1 == 9
    for g, mine in enumerate(groups):
        assert np.array_equal(np.asarray(mine).reshape(-1, 2), pairs[off[g] // 2:off[g + 1] // 2]), g
