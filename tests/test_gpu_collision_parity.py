"""-m gpu: GPU LBVH collision detection (through the C ABI) against the CPU oracle.
Integer work — contact lists (query vertex, hit tet, closest face, type flags) and the Morton leaf order of every BVH —
must be bit-exact on identical states; trajectories with contact are compared within a stated displacement tolerance."""
import numpy as np
import pytest

from conftest import make_contact_scene, contacts_table

pytestmark = pytest.mark.gpu


def _pair(**kw):
    from gaia_b200.capi import Scene
    from oracle.binding import OracleScene
    return make_contact_scene(Scene, **kw), make_contact_scene(OracleScene, threads=2, **kw)


def _same_state(g, o, n):
    for i in range(n):
        assert np.array_equal(g.positions(i), o.positions(i))
        assert np.array_equal(g.prev_positions(i), o.prev_positions(i))


def test_static_contact_sets_dcd_are_bit_exact():
    g, o = _pair(n_objects=3, ccd=0)
    for s in (g, o):
        s.detect_collisions(True, False)
    cg, co = g.contacts(), o.contacts()
    assert len(co) > 10 and (co["shortestPathFound"] == 1).any()
    assert np.array_equal(contacts_table(cg), contacts_table(co))


def test_contact_sets_after_initial_step_dcd_and_ccd():
    # DCD at the overlapping rest state, initial step (bit-identical y = x + dt v), CCD over the substep's motion
    g, o = _pair(n_objects=4, v_rel=(0.5, -9.0, 0.3))
    for s in (g, o):
        s.run_iterations(0)
    _same_state(g, o, 4)
    cg, co = g.contacts(), o.contacts()
    assert (co["fromCCD"] == 1).any() and (co["fromCCD"] == 0).any()
    assert np.array_equal(contacts_table(cg), contacts_table(co))
    sg, so = g.stats(), o.stats()
    assert sg["numContactsDCD"] == so["numContactsDCD"] and sg["numContactsCCD"] == so["numContactsCCD"]


def test_ccd_only_fast_impact():
    g, o = _pair(n_objects=2, dcd=0, gap=0.02, v_rel=(0, -30.0, 0))
    for s in (g, o):
        s.run_iterations(0)
    _same_state(g, o, 2)
    co = o.contacts()
    assert len(co) > 0 and (co["fromCCD"] == 1).all()
    assert np.array_equal(contacts_table(g.contacts()), contacts_table(co))


def test_self_collision_of_a_folded_beam():
    from gaia_b200 import capi, meshio
    from gaia_b200.capi import Scene
    from oracle.binding import OracleScene
    verts, tets = meshio.kuhn_beam(16, 2, 2, h=0.1, origin=(0, 1, 0))
    colors = meshio.kuhn_coloring(16, 2, 2)
    # fold the beam back onto itself: the second half is mirrored about x = 0.8 and lowered into the first half
    folded = verts.copy()
    hi = verts[:, 0] > 0.8
    folded[hi, 0] = 1.6 - verts[hi, 0] + 0.013
    folded[hi, 1] = verts[hi, 1] + 0.12
    folded[hi, 2] = verts[hi, 2] + 0.017
    res = []
    for cls, kw in ((Scene, {}), (OracleScene, dict(threads=2))):
        s = cls(capi.default_physics_params(numSubsteps=2, iterations=4), capi.default_collision_params(allowCCD=0), **kw)
        s.add_tet_mesh(verts, tets, capi.default_object_params(miu=1e5, lmbd=1e6, density=50), colors)
        s.finalize()
        s.set_state(0, folded, np.zeros_like(folded))
        s.detect_collisions(True, False)
        res.append(s.contacts())
    assert len(res[1]) > 5 and (res[1]["hitMesh"] == 0).all()
    assert np.array_equal(contacts_table(res[0]), contacts_table(res[1]))
    # with self collision switched off nothing is reported
    for cls, kw in ((Scene, {}), (OracleScene, dict(threads=1))):
        s = cls(capi.default_physics_params(), capi.default_collision_params(allowCCD=0, handleSelfCollision=0), **kw)
        s.add_tet_mesh(verts, tets, capi.default_object_params(), colors)
        s.finalize()
        s.set_state(0, folded, np.zeros_like(folded))
        s.detect_collisions(True, False)
        assert len(s.contacts()) == 0


def test_bvh_leaf_order_is_bit_exact():
    g, o = _pair(n_objects=4, v_rel=(0.5, -9.0, 0.3))
    for s in (g, o):
        s.run_iterations(0)
    for which in (0, 1, 2):
        lg, lo = g.bvh_leaf_order(which), o.bvh_leaf_order(which)
        assert len(lg) == len(lo) and len(lo) > 50
        # the GPU trees were built at the state of their first detection: tet/surface scenes before the initial step
        # (positions = rest state = the oracle's prev), the CCD scene after it
    # rebuild on identical states and compare exactly
    g2, o2 = _pair(n_objects=4)
    for s in (g2, o2):
        s.detect_collisions(True, True)
    for which in (0, 1):
        assert np.array_equal(g2.bvh_leaf_order(which), o2.bvh_leaf_order(which))


def test_inactive_mesh_is_invisible_until_it_appears():
    from gaia_b200 import capi, meshio
    from gaia_b200.capi import Scene
    from oracle.binding import OracleScene
    scenes = []
    for cls, kw in ((Scene, {}), (OracleScene, dict(threads=2))):
        s = cls(capi.default_physics_params(numSubsteps=2, iterations=4, worldBoundsMin=[-10, 0, -10], worldBoundsMax=[10, 20, 10]),
                capi.default_collision_params(), **kw)
        a = meshio.kuhn_beam(3, 2, 2, h=0.25, origin=(0, 0.2, 0))
        b = meshio.kuhn_beam(2, 2, 2, h=0.2, origin=(0.1, 0.6, 0.05))
        s.add_tet_mesh(a[0], a[1], capi.default_object_params(miu=1e5, lmbd=1e6, density=50), meshio.kuhn_coloring(3, 2, 2))
        s.add_tet_mesh(b[0], b[1], capi.default_object_params(miu=1e5, lmbd=1e6, density=50, frameToAppear=2), meshio.kuhn_coloring(2, 2, 2))
        s.finalize()
        s.run_frames(2)
        assert len(s.contacts()) == 0            # the second block does not exist yet
        scenes.append(s)
    g, o = scenes
    assert np.abs(g.positions(0) - o.positions(0)).max() < 1e-4
    # put both on the SAME state (the GPU's), then let frame 2 begin: the second block appears, overlapping the first
    p, v = g.positions(0), g.velocities(0)
    for s in scenes:
        s.set_state(0, p, v)
        s.run_iterations(0)
    cg, co = contacts_table(g.contacts()), contacts_table(o.contacts())
    assert len(co) > 0 and np.array_equal(cg, co)


def test_stack_trajectory_with_contact_and_friction():
    # three blocks settle on each other and on the ground: 40 frames; tolerance 2e-3 (1 % of the smallest block edge)
    g, o = _pair(n_objects=3, substeps=3, iterations=8, gap=0.01, v_rel=(0.2, -1.0, 0.1))
    for s in (g, o):
        s.run_frames(40)
    worst = 0.0
    for i in range(3):
        worst = max(worst, float(np.abs(g.positions(i) - o.positions(i)).max()))
    assert o.stats()["numContactsDCD"] + o.stats()["numContactsCCD"] > 0
    assert o.positions(2)[:, 1].min() > 0.2          # the top block rests on the others, it did not fall through
    assert worst < 2e-3, worst


def test_intermediate_collision_detection_and_accelerator():
    g, o = _pair(n_objects=2, substeps=2, iterations=9, gap=-0.03, phys_kw=dict(intermediateCollisionIterations=3, useAccelerator=1, acceleratorRho=0.8))
    for s in (g, o):
        s.run_frames(6)
    for i in range(2):
        assert np.abs(g.positions(i) - o.positions(i)).max() < 2e-3


@pytest.mark.gpu
def test_contact_runs_are_bit_deterministic():
    """Two runs of the same stacked-contact scene give identical bits (a race in the relation sort once made vertices that
    sit on several contact faces accumulate their face-side terms in a run-dependent order)."""
    from gaia_b200 import capi
    from conftest import make_contact_scene
    runs = []
    for rep in range(2):
        g = make_contact_scene(capi.Scene, n_objects=4, substeps=2, iterations=6)
        hist, most = [], 0
        for f in range(25):
            g.run_frames(1)
            hist.append(np.concatenate([g.positions(m) for m in range(4)]))
            most = max(most, g.stats()["numContactsDCD"] + g.stats()["numContactsCCD"])
        assert most > 0
        runs.append(hist)
        g.close()
    for a, b in zip(*runs):
        assert np.array_equal(a, b)
