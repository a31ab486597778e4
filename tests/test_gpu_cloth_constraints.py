"""GPU: sewing and vertex-face attachment constraints of the cloth path (gaia_b200_cloth_set_sewing / _add_attachment; the
reference's VBDBaseTriMesh::accumlateSewingForceAndHessian, M/VBDCloth/VBDBaseTriMesh.cpp:320-344, and
VBDClothSimulationFramework::accumulateConstraintForceAndHessian, VBDClothPhysics.cpp:1464-1498) against the CPU oracle, whose
two terms are pinned to those reference lines in tests/test_oracle_vs_ref.py. Through the C ABI.

The colourings below keep every pair of vertices that a seam or an attachment couples in different colours (as the reference
needs from its colour files), so the oracle's sequential sweep and the GPU's parallel one solve the same problem; tolerances as
in test_gpu_cloth_parity.py (the constraint terms join the contact pre-pass slot, so sums associate differently)."""
import numpy as np
import pytest

from gaia_b200 import cloth
from gaia_b200.capi import GaiaError
from oracle.binding_cloth import OracleClothScene

from conftest import rel_err

pytestmark = pytest.mark.gpu


def shifted_grid_coloring(n, shift):
    """grid_cloth_coloring with the row class shifted: still 9 classes of equal size for n = 9"""
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    c = (((i + shift) % 3) * 3 + (j % 3)).ravel()
    return [np.nonzero(c == k)[0].astype(np.int32) for k in range(9)]


# seams of cloth 0 and attachments (mesh, vertex, target mesh or -1, three target vertices, barycentrics / position, stiffness)
SEAMS = np.array([[40, 44, 53], [2, 20, 21], [76, 58, 59]], np.int32)
RATIOS = np.array([0.4, 0.0, 1.0], np.float32)
ATTACH = [(1, 10, 0, (0, 1, 9), (0.5, 0.25, 0.25), 4e3), (1, 16, 0, (7, 8, 17), (0.2, 0.8, 0.0), 4e3), (1, 64, 0, (63, 72, 54), (0.3, 0.3, 0.4), 4e3),
          (1, 70, 0, (71, 62, 80), (1.0 / 3, 1.0 / 3, 1.0 / 3), 4e3), (1, 40, -1, None, (0.0, 0.75, 0.0), 2e3), (1, 22, 2, (0, 1, 4), (0.3, 0.3, 0.4), 5e2),
          (0, 31, -1, None, (0.1, 1.05, -0.1), 1e3)]


def hanging_pair(cls, kw, handle_collision=0, sew_k=1.5e3):
    """cloth 0 pinned at two corners with three seams; cloth 1 hangs from vertex triples of cloth 0 by four attachments, one of
    its vertices is tied to a fixed point and one to a face of a static collider sheet.
    Colours: the cloth path merges the colours of all simulated meshes into parallel groups (initializeParallelGroups: colour c
    of a mesh joins the smallest group so far, first one on ties). Both cloths get 9 classes of 9 vertices, so class c of either
    lands in group c; cloth 1's classes are shifted by one row, and the seams / attachments are chosen so that no two vertices
    one of them couples share a group: the three vertices of a seam; an attached vertex and its three targets; and - because
    the reference computes a target vertex's reaction from x(V-side mesh, that vertex's id) - vertex a of cloth 0 and vertex a
    of cloth 1 (tests/test_multigpu_cpu.py-style check of exactly this in test_no_coupled_pair_shares_a_group)."""
    n = 9
    v0, f0 = cloth.grid_cloth(n, 1.0, origin=(0.0, 1.0, 0.0))
    v1, f1 = cloth.grid_cloth(n, 1.0, origin=(0.0, 0.8, 0.0))
    vc, fc = cloth.grid_cloth(4, 1.4, origin=(0.0, 0.3, 0.0))
    rng = np.random.RandomState(3)
    v0 = v0 + rng.uniform(-0.004, 0.004, v0.shape).astype(np.float32); v1 = v1 + rng.uniform(-0.004, 0.004, v1.shape).astype(np.float32)
    s = cls(cloth.default_cloth_physics_params(numSubsteps=4, iterations=8, handleCollision=handle_collision, contactRadius=0.02, maxQueryDis=0.03,
                                               contactStiffness=1e5), **kw)
    ok = dict(density=0.2, miu=2e3, bendingStiffness=0.5, dampingStVK=1e-6, dampingBending=1e-6)
    ok["lambda"] = 2e3
    s.add_tri_mesh(v0, f0, cloth.default_cloth_object_params(**ok), shifted_grid_coloring(n, 0), fixed=[0, n - 1])
    s.add_tri_mesh(v1, f1, cloth.default_cloth_object_params(**ok), shifted_grid_coloring(n, 1))
    s.add_tri_mesh(vc, fc, cloth.default_cloth_object_params(), simulated=False)
    s.set_sewing(0, SEAMS, RATIOS, sew_k)
    for mv, v, mf, fv, b, k in ATTACH:
        s.add_attachment(mv, v, mf, fv, b, k)
    s.finalize()
    return s


def both(build, **kw):
    return build(cloth.ClothScene, {}, **kw), build(OracleClothScene, dict(threads=4), **kw)


# Calibration of the bars below (oracle against itself with every input coordinate moved by 1 / 8 ulp, this scene, seams at
# k = 300): after one iteration 0.9e-6 / 6.6e-6; after two frames 1.0e-6 / 2.8e-6 without collisions; after three frames
# 2.9e-6 / 1.5e-5 with them (7.3e-5 on the way). Without collisions the cloths pass through each other and the third frame is
# chaotic (1 ulp -> 0.1), as are seams at k = 1500 from the first frame on: hence two frames and the softer seams for the
# trajectories. A wrong weight, sign or index in either term shows as >= 1e-3 after ONE iteration.
@pytest.mark.parametrize("variant", ["0", "1", "3"])
def test_sewing_and_attachments_without_collisions(variant, monkeypatch):
    """No contact pre-pass at all: the constraint kernels alone fill the slot the sweep reads (every sweep variant)."""
    if variant != "0":
        monkeypatch.setenv("GAIA_B200_CLOTH_VARIANT", variant)
    g, o = both(hanging_pair)                                          # stiff seams: one iteration only
    g.run_iterations(1); o.run_iterations(1)
    for m in range(2):
        assert rel_err(g.positions(m), o.positions(m)) < 2e-5
    g, o = both(hanging_pair, sew_k=300.0)
    x1 = o.positions(1).copy()
    g.run_frames(2); o.run_frames(2)
    for m in range(2):
        pg, po = g.positions(m), o.positions(m)
        assert np.isfinite(pg).all()
        assert np.abs(pg - po).max() < 5e-4, (m, np.abs(pg - po).max())
    # not vacuous: in 1/30 s gravity alone moves a vertex by 6 mm; the seam has pulled vertex 40 of cloth 0 a quarter of a metre
    x0, p0 = cloth.grid_cloth(9, 1.0, origin=(0.0, 1.0, 0.0))[0], o.positions(0)
    seam_gap = lambda x: np.linalg.norm(x[40] - (0.6 * x[44] + 0.4 * x[53]))
    assert seam_gap(p0) < 0.6 * seam_gap(x0) and x1[22, 1] - o.positions(1)[22, 1] > 0.03
    sg, so = g.stats(), o.stats()
    assert sg["vertexSolves"] == so["vertexSolves"] and sg["colourSweeps"] == so["colourSweeps"]


@pytest.mark.parametrize("exact", ["0", "1"])
def test_sewing_and_attachments_with_collisions(exact, monkeypatch):
    """With the collision pipeline on, vertices hold contact terms AND constraint terms in the same slot (the two cloths touch
    where they are tied together; the lower one is tied to the collider sheet)."""
    if exact == "1":
        monkeypatch.setenv("GAIA_B200_CLOTH_EXACT_ORDER", "1")
    g, o = both(hanging_pair, handle_collision=1)
    g.run_iterations(1); o.run_iterations(1)
    for m in range(2):
        assert rel_err(g.positions(m), o.positions(m)) < 2e-5
    g, o = both(hanging_pair, handle_collision=1, sew_k=300.0)
    g.run_frames(3); o.run_frames(3)
    for m in range(2):
        pg, po = g.positions(m), o.positions(m)
        assert np.isfinite(pg).all()
        assert np.abs(pg - po).max() < 5e-4, (m, np.abs(pg - po).max())
    so = o.stats()
    assert so["numberCollisionDetection"] > 0 and so["numContactsVF"] + so["numContactsEE"] > 0


def test_no_coupled_pair_shares_a_group():
    """the premise of the two tests above, checked on the groups the library itself built"""
    g = hanging_pair(cloth.ClothScene, {})
    gid = {}
    for gi, grp in enumerate(g.parallel_groups()):
        for m, v in np.asarray(grp).reshape(-1, 2):
            gid[(int(m), int(v))] = gi
    for s3 in SEAMS:
        assert len({gid[(0, int(v))] for v in s3}) == 3
    for mv, v, mf, fv, bary, _ in ATTACH:
        if mf < 0 or mf >= 2:
            continue
        assert len({gid[(mv, v)]} | {gid[(mf, a)] for a in fv}) == 4
        for a, b in zip(fv, bary):
            if b != 0.0:
                assert gid[(mf, a)] != gid[(mv, a)]


def test_seam_filtered_pairs_contact_sets_bit_exact():
    """gaia_b200_cloth_set_filtered_pairs (the reference's filteredPairsPath): on a sheet folded onto itself every 7th detected
    self contact is declared a filtered pair; contact sets and conservative bounds stay bit-identical to the oracle's, the pairs
    are gone, and without seams the lists are inert (tests/test_oracle_cloth.py checks the oracle's side of this)."""
    from cloth_filter_scene import build, pick_pairs
    base = build(OracleClothScene, dict(threads=2))
    base.detect_contacts()
    vf0, ee0 = base.vf_contacts(), base.ee_contacts()
    vf_pairs, ee_pairs = pick_pairs(vf0, ee0)
    g = build(cloth.ClothScene, {}, vf_pairs=vf_pairs, ee_pairs=ee_pairs)
    o = build(OracleClothScene, dict(threads=2), vf_pairs=vf_pairs, ee_pairs=ee_pairs)
    g.detect_contacts(); o.detect_contacts()
    vf1, ee1 = o.vf_contacts(), o.ee_contacts()
    assert 50 < len(vf1) < len(vf0) and 50 < len(ee1) < len(ee0)
    assert np.array_equal(g.vf_contacts(), vf1) and np.array_equal(g.ee_contacts(), ee1)
    assert np.array_equal(g.conservative_bounds(0), o.conservative_bounds(0))
    have = {(int(r[1]), int(r[3])) for r in g.vf_contacts()}
    assert not any((int(v), int(f)) in have for v, f in vf_pairs)
    g2 = build(cloth.ClothScene, {}, seams=False, vf_pairs=vf_pairs, ee_pairs=ee_pairs)
    g2.detect_contacts()
    assert np.array_equal(g2.vf_contacts(), vf0) and np.array_equal(g2.ee_contacts(), ee0)
    assert np.array_equal(g2.conservative_bounds(0), base.conservative_bounds(0))


def test_constraint_arguments_are_checked():
    v, f = cloth.grid_cloth(4, 1.0)
    for cls, kw in ((cloth.ClothScene, {}), (OracleClothScene, dict(threads=1))):
        s = cls(cloth.default_cloth_physics_params(), **kw)
        s.add_tri_mesh(v, f, cloth.default_cloth_object_params(), cloth.grid_cloth_coloring(4))
        cv, cf = cloth.grid_cloth(3, 0.5, origin=(2.0, 0.0, 0.0))
        s.add_tri_mesh(cv, cf, cloth.default_cloth_object_params(), simulated=False)
        with pytest.raises(GaiaError):
            s.set_sewing(0, [[0, 1, 99]], [0.5], 1e3)                 # seam vertex out of range
        with pytest.raises(GaiaError):
            s.set_sewing(1, [[0, 1, 2]], [0.5], 1e3)                  # a collider is not sewn
        with pytest.raises(GaiaError):
            s.add_attachment(1, 0, -1, None, [0, 0, 0])               # the vertex side is a simulated mesh
        with pytest.raises(GaiaError):
            s.add_attachment(0, 16, -1, None, [0, 0, 0])              # vertex out of range
        with pytest.raises(GaiaError):
            s.add_attachment(0, 3, 1, [0, 1, 9], [0.3, 0.3, 0.4])     # face vertex out of range
        with pytest.raises(GaiaError):
            s.set_filtered_pairs(0, [[0, 18]], [])                    # the 4 x 4 sheet has 18 faces: 0 .. 17
        with pytest.raises(GaiaError):
            s.set_filtered_pairs(0, [], [[0, 33]])                    # ... and 33 edges
        with pytest.raises(GaiaError):
            s.set_filtered_pairs(1, [[0, 0]], [])                     # colliders have no seams
        s.set_filtered_pairs(0, [[0, 17]], [[0, 32]])
        s.add_attachment(0, 3, 1, [0, 1, 2], [0.3, 0.3, 0.4])
        s.set_sewing(0, [[0, 5, 6]], [0.5], 1e3)
        s.finalize()
        with pytest.raises(GaiaError):
            s.set_sewing(0, [[0, 5, 6]], [0.5], 1e3)                  # after finalize
        with pytest.raises(GaiaError):
            s.add_attachment(0, 3, -1, None, [0, 0, 0])
