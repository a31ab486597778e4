"""-m gpu: the CUDA VBD step (through the C ABI) against the CPU oracle on the same seeded inputs.
Tolerances: integer work bit-exact; positions after one iteration <= 1e-5 relative (north_star); trajectories within the
max-displacement tolerance stated per test."""
import numpy as np
import pytest

from conftest import make_beam_scene, rel_err

pytestmark = pytest.mark.gpu


def _pair(**kw):
    from gaia_b200.capi import Scene
    from oracle.binding import OracleScene
    return make_beam_scene(Scene, **kw), make_beam_scene(OracleScene, threads=2, **kw)


def test_setup_is_bit_exact():
    g, o = _pair(obj_kw=dict(rotation=[0.3, -0.2, 0.9], scale=[1.5, 0.7, 1.1], translation=[0.2, 3.0, -1.0]))
    assert np.array_equal(g.positions(0), o.positions(0))            # transform
    assert np.array_equal(g.vertex_masses(0), o.vertex_masses(0))    # lumped mass
    assert np.array_equal(g.surface_vertex_ids(0), o.surface_vertex_ids(0))
    assert np.array_equal(g.surface_faces(0), o.surface_faces(0))
    for a, b in zip(g.parallel_groups(), o.parallel_groups()):
        assert np.array_equal(a, b)
    assert g.mesh_sizes(0) == o.mesh_sizes(0)


@pytest.mark.parametrize("init_type", [0, 1, 2, 3, 4, 6])
def test_positions_after_one_iteration(init_type):
    g, o = _pair(obj_kw=dict(initializationType=init_type))
    for s in (g, o):
        s.run_frames(1)          # give type 2/4/6 a velocity history
        s.run_iterations(1)
    assert np.array_equal(g.prev_positions(0), o.prev_positions(0)) or rel_err(g.prev_positions(0), o.prev_positions(0)) < 1e-5
    assert rel_err(g.inertia(0), o.inertia(0)) < 1e-5
    assert rel_err(g.positions(0), o.positions(0)) < 1e-5


def test_symplectic_euler_initial_guess_type_5():
    """initializationType 5 (VBDBaseTetMesh::forwardStepSimplecticEuler, M/VBD/VBD_BaseMaterial.cpp:134-147): the elastic force of
    every vertex first, then x += dt (v_prev + dt f / m). The oracle's version is bit-identical to those reference lines
    (tests/test_oracle_vs_ref.py). A softer, heavier beam than the other types' test: the explicit guess is only tame for
    dt^2 k / m << 1 - on the default beam it throws vertices 0.35 m and a 1-ulp change of the input moves the result by 5e-5;
    here the guess moves them by up to 1.2 cm and 1 / 8 ulp move the result by 4e-7."""
    kw = dict(obj_kw=dict(initializationType=5, miu=1e4, lmbd=1e5, density=200), substeps=4)
    g, o = _pair(**kw)
    for s in (g, o):
        s.run_frames(1)
        s.run_iterations(1)
    assert rel_err(g.prev_positions(0), o.prev_positions(0)) < 1e-5
    assert rel_err(g.inertia(0), o.inertia(0)) < 1e-5
    assert rel_err(g.positions(0), o.positions(0)) < 1e-5
    g.run_frames(3); o.run_frames(3)
    assert np.abs(g.positions(0) - o.positions(0)).max() < 1e-4
    # not the inertia guess in disguise
    from oracle.binding import OracleScene
    o0 = make_beam_scene(OracleScene, threads=2, obj_kw=dict(initializationType=0, miu=1e4, lmbd=1e5, density=200), substeps=4)
    o0.run_frames(1); o0.run_iterations(1); o0.run_frames(3)
    assert np.abs(o0.positions(0) - o.positions(0)).max() > 1e-3


def test_first_iteration_from_rest_state():
    g, o = _pair()
    g.run_iterations(1); o.run_iterations(1)
    assert np.array_equal(g.inertia(0), o.inertia(0))       # y = x + dt*(v + dt*g): elementwise, no contraction
    assert rel_err(g.positions(0), o.positions(0)) < 1e-5


def test_ground_contact_and_friction_trajectory():
    # 30 frames: free fall, impact on the ground plane with friction, rebound. Tolerance: 1e-3 of the beam length
    g, o = _pair(v0=(1.0, -5.0, 0.5), substeps=4, iterations=10)
    for s in (g, o):
        s.run_frames(30)
    pg, po = g.positions(0), o.positions(0)
    assert po[:, 1].min() < 0.05                              # it reached the ground
    assert np.abs(pg - po).max() < 1e-3 * 1.5
    assert rel_err(g.velocities(0), o.velocities(0)) < 5e-2


def test_damping_and_velocity_damping():
    kw = dict(obj_kw=dict(dampingHydrostatic=1e-4, dampingDeviatoric=1e-4, exponentialVelDamping=0.9, constantVelDamping=0.1,
                          maxVelocityMagnitude=25.0))
    g, o = _pair(**kw)
    for s in (g, o):
        s.run_frames(5)
    assert np.abs(g.positions(0) - o.positions(0)).max() < 2e-4
    assert np.abs(g.velocities(0) - o.velocities(0)).max() < 2e-2


def test_fixed_points_and_greedy_colouring():
    from gaia_b200 import capi, meshio
    from gaia_b200.capi import Scene
    from oracle.binding import OracleScene
    verts, tets = meshio.kuhn_beam(8, 2, 2, h=0.2, origin=(0, 2.0, 0))
    colors = meshio.greedy_coloring(len(verts), tets)
    assert meshio.check_coloring(len(verts), tets, colors)
    fixed = np.nonzero(verts[:, 0] < 1e-6)[0].astype(np.int32)
    phys = capi.default_physics_params(numSubsteps=3, iterations=8)
    col = capi.default_collision_params(allowDCD=0, allowCCD=0)
    obj = capi.default_object_params(miu=5e4, lmbd=5e5, density=100)
    out = []
    for cls, kw in ((Scene, {}), (OracleScene, dict(threads=2))):
        s = cls(phys, col, **kw)
        s.add_tet_mesh(verts, tets, obj, colors, fixed)
        s.finalize()
        s.run_frames(10)
        out.append((s.positions(0), s.velocities(0)))
    (pg, vg), (po, vo) = out
    assert np.array_equal(pg[fixed], verts[fixed]) and np.array_equal(po[fixed], verts[fixed])
    assert (vg[fixed] == 0).all()
    assert po[:, 1].min() < 1.99                                # the cantilever sagged
    assert np.abs(pg - po).max() < 2e-4


def test_two_meshes_group_merge_and_late_appearance():
    from gaia_b200 import capi, meshio
    from gaia_b200.capi import Scene
    from oracle.binding import OracleScene
    a = meshio.kuhn_beam(4, 2, 2, h=0.25, origin=(0, 1.0, 0)); ca = meshio.kuhn_coloring(4, 2, 2)
    b = meshio.kuhn_beam(3, 3, 2, h=0.2, origin=(3, 0.5, 0)); cb = meshio.greedy_coloring(len(b[0]), b[1])
    phys = capi.default_physics_params(numSubsteps=2, iterations=6, worldBoundsMin=[-10, 0, -10], worldBoundsMax=[10, 20, 10])
    col = capi.default_collision_params(allowDCD=0, allowCCD=0)
    oa = capi.default_object_params(miu=1e5, lmbd=1e6, density=50)
    ob = capi.default_object_params(miu=2e5, lmbd=8e5, density=20, frameToAppear=3, initialVelocity=[0, -2, 0])
    res = []
    for cls, kw in ((Scene, {}), (OracleScene, dict(threads=2))):
        s = cls(phys, col, **kw)
        s.add_tet_mesh(a[0], a[1], oa, ca); s.add_tet_mesh(b[0], b[1], ob, cb)
        s.finalize()
        groups = s.parallel_groups()
        s.run_frames(2)
        p_early = s.positions(1)
        s.run_frames(6)
        res.append((groups, p_early, s.positions(0), s.positions(1), s.stats()["vertexSolves"]))
    for x, y in zip(res[0][0], res[1][0]):
        assert np.array_equal(x, y)
    assert np.array_equal(res[0][1], res[1][1]) and np.array_equal(res[0][1], b[0])   # mesh 1 untouched before frame 3
    assert np.abs(res[0][2] - res[1][2]).max() < 2e-4 and np.abs(res[0][3] - res[1][3]).max() < 2e-4
    assert res[0][4] == res[1][4]                                                     # same vertex-solve count


def test_chebyshev_accelerator_matches_oracle():
    g, o = _pair(phys_kw=dict(useAccelerator=1, acceleratorRho=0.9), iterations=12)
    for s in (g, o):
        s.run_frames(4)
    assert np.abs(g.positions(0) - o.positions(0)).max() < 2e-4


def test_state_roundtrip_and_whole_scene_io():
    g, o = _pair()
    g.run_frames(2)
    p, v = g.positions(0), g.velocities(0)
    assert np.array_equal(g.all_positions(), p)
    g.set_state(0, p + 0.25, v * 0.5)
    assert np.array_equal(g.positions(0), (p + 0.25).astype(np.float32))
    assert np.array_equal(g.velocities(0), (v * 0.5).astype(np.float32))
    sp = g.surface_positions(0)
    assert np.array_equal(sp, g.positions(0)[g.surface_vertex_ids(0)])


@pytest.mark.gpu
def test_registered_host_buffers_take_the_direct_dma_path():
    from gaia_b200 import capi
    from conftest import make_beam_scene
    g = make_beam_scene(capi.Scene)
    g.run_frames(1)
    ref = g.all_positions()
    buf = np.zeros_like(ref)
    g.register_host_buffer(buf)
    assert np.array_equal(g.all_positions(out=buf), ref)          # D2H straight into the page-locked array
    buf += np.float32(0.125)
    vel = np.ones_like(buf)
    g.set_all_state(buf, vel)                                      # pos from the registered array, vel through staging
    assert np.array_equal(g.all_positions(), ref + np.float32(0.125))
    g.unregister_host_buffer(buf)
    with pytest.raises(capi.GaiaError):
        g.unregister_host_buffer(buf)
    assert np.array_equal(g.all_positions(out=buf), ref + np.float32(0.125))
    g.close()


@pytest.mark.gpu
def test_hundred_frame_trajectory_config1_shape():
    """north_star: "trajectories over 100 frames within a stated max-displacement tolerance". Config-1 shape (one tet mesh
    dropping onto the ground plane, ground contact + friction only, S01 teapot material, 4 substeps x 20 iterations, 100
    frames = 1.67 s: fall, impact, rebound, sliding to rest). Stated tolerance: max |x_gpu - x_oracle| <= 5e-3 at every
    checkpoint (0.33 % of the 1.5-unit beam; the ground-friction Hessian is discontinuous at zero sliding, so rounding
    differences grow at each impact) and <= 1e-4 while the body is still in free flight."""
    g, o = _pair(nx=6, ny=3, nz=3, y0=2.0, v0=(0.6, -1.0, 0.3), substeps=4, iterations=20,
                 obj_kw=dict(miu=1e5, lmbd=1e6, density=50), phys_kw=dict(boundaryFrictionDynamic=0.1, boundaryCollisionStiffness=1e5))
    worst = []
    for chunk in range(10):
        for s in (g, o):
            s.run_frames(10)
        pg, po = g.positions(0), o.positions(0)
        assert np.isfinite(pg).all()
        worst.append(float(np.abs(pg - po).max()))
        if po[:, 1].min() > 0.05:
            assert worst[-1] < 1e-4, (chunk, worst)
    assert o.positions(0)[:, 1].min() < 0.02                      # it landed
    assert max(worst) < 5e-3, worst


@pytest.mark.gpu
def test_solver_lanes_option_changes_association_only():
    """gaia_b200_set_solver_lanes: 8 lanes per vertex instead of 4 is the same arithmetic per tet with a different tree of
    partial sums - one iteration agrees with the oracle to the same 1e-5 bar, and the two settings differ by rounding only."""
    from gaia_b200 import capi
    g4, o = _pair()
    g8, _ = _pair()
    g8.set_solver_lanes(8)
    with pytest.raises(capi.GaiaError):
        g8.set_solver_lanes(3)
    for s in (g4, g8, o):
        s.run_frames(1)
        s.run_iterations(1)
    assert rel_err(g4.positions(0), o.positions(0)) < 1e-5 and rel_err(g8.positions(0), o.positions(0)) < 1e-5
    assert np.abs(g4.positions(0) - g8.positions(0)).max() < 1e-5
    g8.set_solver_lanes(0)
    g4.run_frames(2); g8.run_frames(2)          # back on the default: the two runs only differ by the one 8-lane frame
    assert np.abs(g4.positions(0) - g8.positions(0)).max() < 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["twist", "twist_dcd", "ground_friction", "pile_before_impact"])
def test_multi_colour_launch_is_bit_identical_to_one_launch_per_colour(case):
    """gaia_b200_set_multi_colour_launch: contact-free stretches of the colour loop as ONE cooperative launch (grid barrier
    between colours, next colour's stream copies started before the barrier) against one launch per colour - same bits,
    with rotators, with intermediate detection cutting the stretches, with the world-box friction switching on at
    frictionStartIter, and on a many-mesh scene (unequal, partly empty colours per block)."""
    from gaia_b200 import capi, workloads
    if case == "twist":
        wl = workloads.WORKLOADS["twist_beam_small"]()
    elif case == "twist_dcd":
        wl = workloads.workload_twist_beam(24, 12, 12, collisions=True)
    elif case == "ground_friction":
        wl = None
    else:
        wl = workloads.workload_drop_pile(8, grid=(2, 2), preroll=0)

    def build():
        if wl is None:
            return make_beam_scene(capi.Scene, v0=(1.0, -5.0, 0.5), substeps=4, iterations=10, phys_kw=dict(frictionStartIter=3))
        return workloads.build_scene(capi.Scene, wl)
    a, b = build(), build()
    a.set_multi_colour_launch(2)          # everywhere, also where the default would replay a graph of per-colour launches
    b.set_multi_colour_launch(False)
    frames = 12 if case != "pile_before_impact" else 6
    for s in (a, b):
        s.run_frames(frames)
    n_mesh = len(wl["meshes"]) if wl is not None else 1
    moved = 0.0
    for m in range(n_mesh):
        assert np.array_equal(a.positions(m), b.positions(m)), (case, m)
        assert np.array_equal(a.velocities(m), b.velocities(m)), (case, m)
    sa, sb = a.stats(), b.stats()
    assert sa["vertexSolves"] == sb["vertexSolves"] and sa["colourSweeps"] == sb["colourSweeps"]
    assert sa["kernelLaunches"] < sb["kernelLaunches"]            # the stretches really were single launches
