"""ORACLE — TEST INFRASTRUCTURE ONLY. ctypes binding of the cloth part of oracle/liboracle.so (CPU restatement of the
reference VBD cloth step). Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs only."""
import ctypes as C

import numpy as np

IPT = C.POINTER(C.c_int32)

from gaia_b200.capi import _f32, _fp
from gaia_b200.cloth import ClothPhysicsParams, ClothSceneBase, bind_cloth_api, default_cloth_physics_params
from .binding import load_oracle, use_speed_build  # noqa: F401

_bound = False


def _lib():
    global _bound
    lib = load_oracle()
    if not _bound:
        P, FP, IP = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int32)
        lib.oracle_cloth_create.argtypes = [C.POINTER(ClothPhysicsParams), C.c_int]
        lib.oracle_cloth_create.restype = C.c_void_p
        lib.oracle_cloth_destroy.argtypes = [P]; lib.oracle_cloth_destroy.restype = None
        lib.oracle_cloth_num_threads.argtypes = [P]
        lib.oracle_cloth_set_brute_force.argtypes = [P, C.c_int]; lib.oracle_cloth_set_brute_force.restype = None
        lib.oracle_cloth_set_prev_positions.argtypes = [P, C.c_int, FP]
        lib.oracle_cloth_get_topology.argtypes = [P, C.c_int, IP, IP, IP, C.POINTER(C.c_int64)]
        lib.oracle_cloth_check_feasible_region.argtypes = [P, C.c_int, FP, C.c_int, C.c_int, C.c_float]
        lib.oracle_cloth_probe_initial_step.argtypes = [P]; lib.oracle_cloth_probe_update_velocity.argtypes = [P]
        lib.oracle_cloth_probe_truncate.argtypes = [P, C.c_int, C.c_int, FP]
        lib.oracle_cloth_get_inertia.argtypes = [P, C.c_int, FP]
        lib.oracle_cloth_set_bounds.argtypes = [P, C.c_int, FP, FP]
        lib.oracle_cloth_probe_vf.argtypes = [P] + [C.c_int] * 5 + [FP, FP]; lib.oracle_cloth_probe_vf.restype = C.c_float
        lib.oracle_cloth_probe_ee.argtypes = [P] + [C.c_int] * 5 + [FP, FP]; lib.oracle_cloth_probe_ee.restype = C.c_float
        lib.oracle_cloth_probe_boundary.argtypes = [P, C.c_int, C.c_int, FP, FP]; lib.oracle_cloth_probe_boundary.restype = None
        lib.oracle_cloth_material_force_hessian.argtypes = [P, C.c_int, C.c_int, FP, FP]
        lib.oracle_cloth_vertex_force_hessian.argtypes = [P, C.c_int, C.c_int, FP, FP]
        lib.oracle_cloth_elastic_energy.argtypes = [P, C.c_int]; lib.oracle_cloth_elastic_energy.restype = C.c_float
        lib.oracle_cloth_contact_repulsion.argtypes = [P, C.c_float, FP, FP]; lib.oracle_cloth_contact_repulsion.restype = None
        lib.oracle_cloth_get_dminv.argtypes = [P, C.c_int, FP, FP]
        lib.oracle_cloth_get_bending_forms.argtypes = [P, C.c_int, FP]
        lib.oracle_cloth_closest_points_segments.argtypes = [FP, FP, FP, FP, FP]; lib.oracle_cloth_closest_points_segments.restype = None
        bind_cloth_api(lib, "oracle_cloth_")
        _bound = True
    return lib


class OracleClothScene(ClothSceneBase):
    _prefix = "oracle_cloth_"

    def __init__(self, phys=None, threads=0):
        self._lib = _lib()
        self.phys = phys if phys is not None else default_cloth_physics_params()
        self.h = C.c_void_p(self._lib.oracle_cloth_create(C.byref(self.phys), int(threads)))

    @property
    def num_threads(self):
        return self._lib.oracle_cloth_num_threads(self.h)

    def set_brute_force(self, on=True):
        self._lib.oracle_cloth_set_brute_force(self.h, 1 if on else 0)

    def set_prev_positions(self, mesh, prev):
        p = _f32(prev)
        self._lib.oracle_cloth_set_prev_positions(self.h, mesh, _fp(p))

    def material_force_hessian(self, mesh, v):
        f = np.empty(3, np.float32); h = np.empty(9, np.float32)
        self._lib.oracle_cloth_material_force_hessian(self.h, mesh, v, _fp(f), _fp(h))
        return f, h.reshape(3, 3).T.copy()          # column-major -> H[r, c]

    def vertex_force_hessian(self, mesh, v):
        f = np.empty(3, np.float32); h = np.empty(9, np.float32)
        self._lib.oracle_cloth_vertex_force_hessian(self.h, mesh, v, _fp(f), _fp(h))
        return f, h.reshape(3, 3).T.copy()

    def topology(self, mesh):
        """(edge id of each face edge [nF,3], CSR offsets [nV+1], neighbour vertices)"""
        nV, nF, nE = self.mesh_sizes(mesh)
        n = C.c_int64()
        self._lib.oracle_cloth_get_topology(self.h, mesh, None, None, None, C.byref(n))
        f3 = np.empty((nF, 3), np.int32); off = np.empty(nV + 1, np.int32); nb = np.empty(max(n.value, 1), np.int32)
        self._lib.oracle_cloth_get_topology(self.h, mesh, f3.ctypes.data_as(IPT), off.ctypes.data_as(IPT), nb.ctypes.data_as(IPT), C.byref(n))
        return f3, off, nb[: n.value]

    def check_feasible_region(self, mesh, p, face, point_type, eps):
        q = _f32(p)
        return bool(self._lib.oracle_cloth_check_feasible_region(self.h, mesh, _fp(q), int(face), int(point_type), float(eps)))

    def probe_initial_step(self):
        self._lib.oracle_cloth_probe_initial_step(self.h)

    def probe_update_velocity(self):
        self._lib.oracle_cloth_probe_update_velocity(self.h)

    def probe_truncate(self, mesh, v, new_pos):
        p = _f32(new_pos).copy()
        t = self._lib.oracle_cloth_probe_truncate(self.h, int(mesh), int(v), _fp(p))
        return bool(t), p

    def inertia(self, mesh):
        a = np.empty((self.mesh_sizes(mesh)[0], 3), np.float32)
        self._lib.oracle_cloth_get_inertia(self.h, int(mesh), _fp(a))
        return a

    def set_bounds(self, mesh, bounds, pos_at_cd):
        b = _f32(bounds); p = _f32(pos_at_cd)
        self._lib.oracle_cloth_set_bounds(self.h, int(mesh), _fp(b), _fp(p))

    def probe_vf(self, v_mesh, v, f_mesh, f, order):
        """one vertex-face contact in isolation (closest point recomputed, friction on): (distance, force, H[r, c])"""
        fo = np.empty(3, np.float32); h = np.empty(9, np.float32)
        fn = self._lib.oracle_cloth_probe_vf; fn.restype = C.c_float
        d = fn(self.h, int(v_mesh), int(v), int(f_mesh), int(f), int(order), _fp(fo), _fp(h))
        return float(d), fo, h.reshape(3, 3).T.copy()

    def probe_ee(self, m1, e1, m2, e2, order):
        fo = np.empty(3, np.float32); h = np.empty(9, np.float32)
        fn = self._lib.oracle_cloth_probe_ee; fn.restype = C.c_float
        d = fn(self.h, int(m1), int(e1), int(m2), int(e2), int(order), _fp(fo), _fp(h))
        return float(d), fo, h.reshape(3, 3).T.copy()

    def probe_boundary(self, mesh, v):
        fo = np.empty(3, np.float32); h = np.empty(9, np.float32)
        fn = self._lib.oracle_cloth_probe_boundary; fn.restype = None
        fn(self.h, int(mesh), int(v), _fp(fo), _fp(h))
        return fo, h.reshape(3, 3).T.copy()

    def incidence(self, mesh):
        """vertex -> incident faces / edges of `mesh` (CSR: fOff, fAdj, eOff, eAdj) as the conservative-bounds fold walks them"""
        nV, nF, nE = self.mesh_sizes(mesh)
        f_off = np.zeros(nV + 1, np.int32); e_off = np.zeros(nV + 1, np.int32); f_adj = np.zeros(3 * nF, np.int32); e_adj = np.zeros(2 * nE, np.int32)
        ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int32))
        fn = self._lib.oracle_cloth_get_incidence; fn.restype = C.c_int
        fn(self.h, int(mesh), ip(f_off), ip(f_adj), ip(e_off), ip(e_adj))
        return f_off, f_adj, e_off, e_adj

    def probe_bounds(self, mesh, vf_min, vf_cnt, face_min, face_cnt, ee_min, ee_cnt, mask_before):
        """the last loop of applyCollisionDetection alone: (conservative bounds, mask after)"""
        nV = self.mesh_sizes(mesh)[0]
        b = np.zeros(nV, np.float32); m = np.zeros(nV, np.uint8)
        ip = lambda a: np.ascontiguousarray(a, np.int32).ctypes.data_as(C.POINTER(C.c_int32))
        up = lambda a: a.ctypes.data_as(C.POINTER(C.c_uint8))
        mb = np.ascontiguousarray(mask_before, np.uint8)
        fn = self._lib.oracle_cloth_probe_bounds; fn.restype = C.c_int
        rc = fn(self.h, int(mesh), _fp(_f32(vf_min)), ip(vf_cnt), _fp(_f32(face_min)), ip(face_cnt), _fp(_f32(ee_min)), ip(ee_cnt), up(mb), _fp(b), up(m))
        assert rc == 0
        return b, m

    def sewing_force_hessian(self, mesh, v):
        """VBDBaseTriMesh::accumlateSewingForceAndHessian alone: (force, H column-major 9)"""
        fo = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
        fn = self._lib.oracle_cloth_sewing_force_hessian; fn.restype = None
        fn(self.h, int(mesh), int(v), _fp(fo), _fp(h))
        return fo, h

    def constraint_force_hessian(self, mesh, v):
        """accumulateConstraintForceAndHessian (attachments) alone: (force, H column-major 9)"""
        fo = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
        fn = self._lib.oracle_cloth_constraint_force_hessian; fn.restype = None
        fn(self.h, int(mesh), int(v), _fp(fo), _fp(h))
        return fo, h

    def elastic_energy(self, mesh):
        return float(self._lib.oracle_cloth_elastic_energy(self.h, mesh))

    def contact_repulsion(self, dis):
        a = np.empty(1, np.float32); b = np.empty(1, np.float32)
        self._lib.oracle_cloth_contact_repulsion(self.h, float(dis), _fp(a), _fp(b))
        return float(a[0]), float(b[0])

    def dminv(self, mesh):
        nF = self.mesh_sizes(mesh)[1]
        d = np.empty((nF, 4), np.float32); a = np.empty(nF, np.float32)
        self._lib.oracle_cloth_get_dminv(self.h, mesh, _fp(d), _fp(a))
        return d, a

    def bending_forms(self, mesh):
        q = np.empty((self.mesh_sizes(mesh)[2], 4, 4), np.float32)
        self._lib.oracle_cloth_get_bending_forms(self.h, mesh, _fp(q))
        return q

    def close(self):
        if self.h:
            self._lib.oracle_cloth_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def closest_points_segments(p0, p1, q0, q1):
    out = np.empty(8, np.float32)
    _lib().oracle_cloth_closest_points_segments(_fp(_f32(p0)), _fp(_f32(p1)), _fp(_f32(q0)), _fp(_f32(q1)), _fp(out))
    return out[:3], out[3:6], float(out[6]), float(out[7])
