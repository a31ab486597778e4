"""ORACLE — TEST INFRASTRUCTURE ONLY. ctypes binding of oracle/liboracle.so (CPU restatement of the reference VBD step)
and of oracle/_ref/libgaia_ref.so (the reference's own math compiled from /root/reference, when present).
Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs only."""
import ctypes as C
import os
import numpy as np
from gaia_b200.capi import (SceneBase, bind_scene_api, PhysicsParams, CollisionParams, default_physics_params,
                            default_collision_params, _fp, _ip, _f32, _i32)

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_LIB = os.path.join(_HERE, "liboracle.so")
REF_LIB = os.path.join(_HERE, "_ref", "libgaia_ref.so")
_lib = None
_ref = None


FAST_LIB = os.path.join(_HERE, "liboracle_fast.so")
_use_fast = False


def use_speed_build():
    """bench.py's CPU arm: load the -O3 / AVX2 / FMA build of the same sources (oracle/Makefile: liboracle_fast.so) instead of
    the parity build. Must be called before the first scene is created; returns False when that build is absent or the host
    CPU lacks AVX2 + FMA (the parity build is used then)."""
    global _use_fast
    if _lib is not None:
        return _use_fast
    ok = os.path.exists(FAST_LIB)
    try:
        with open("/proc/cpuinfo") as f:
            flags = f.read()
        ok = ok and (" avx2" in flags and " fma" in flags)
    except OSError:
        ok = False
    _use_fast = ok
    return ok


def load_oracle():
    global _lib
    if _lib is None:
        path = FAST_LIB if _use_fast else ORACLE_LIB
        if not os.path.exists(path):
            raise RuntimeError(f"{path} not built (make -C oracle)")
        lib = C.CDLL(path)
        lib.oracle_create.argtypes = [C.POINTER(PhysicsParams), C.POINTER(CollisionParams), C.c_int]
        lib.oracle_create.restype = C.c_void_p
        lib.oracle_destroy.argtypes = [C.c_void_p]; lib.oracle_destroy.restype = None
        bind_scene_api(lib, "oracle_")
        P, FP, IP = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int32)
        lib.oracle_num_threads.argtypes = [P]
        for f in ("phase_begin_substep", "phase_prepare", "phase_intermediate_cd", "phase_end_substep"):
            getattr(lib, "oracle_" + f).argtypes = [P]
        lib.oracle_phase_solve_group.argtypes = [P, C.c_int, C.c_int]
        lib.oracle_set_prev_positions.argtypes = [P, C.c_int, FP]
        lib.oracle_vertex_merit_energy.argtypes = [P, C.c_int, C.c_int, FP, FP, FP]
        lib.oracle_get_dminv.argtypes = [P, C.c_int, FP]
        lib.oracle_get_rest_volumes.argtypes = [P, C.c_int, FP]
        lib.oracle_get_vertex_tet_csr.argtypes = [P, C.c_int, IP, IP, IP]
        lib.oracle_get_tet_topology.argtypes = [P, C.c_int, IP, IP, IP]
        lib.oracle_get_surface_topology.argtypes = [P, C.c_int, IP, IP, IP, IP, C.POINTER(C.c_int64)]
        lib.oracle_check_feasible_region.argtypes = [P, C.c_int, FP, C.c_int, C.c_int, C.c_float]
        lib.oracle_tet_traverse.argtypes = [P, C.c_int, FP, FP, C.c_float, C.c_int, C.c_int, C.c_int, C.c_float, IP]
        lib.oracle_set_brute_force.argtypes = [P, C.c_int]
        lib.oracle_vertex_material_force_hessian.argtypes = [P, C.c_int, C.c_int, FP, FP]
        lib.oracle_contact_force_hessian.argtypes = [FP, FP, FP, FP, C.c_int] + [C.c_float] * 6 + [FP, FP, FP]
        lib.oracle_boundary_force_hessian.argtypes = [FP, FP, FP, FP] + [C.c_float] * 4 + [C.c_int, FP, FP]
        lib.oracle_solve3x3.argtypes = [FP, FP, FP]
        lib.oracle_ccd_tri_point.argtypes = [FP, FP, FP, FP, FP, FP]
        lib.oracle_cubic_roots.argtypes = [FP, C.c_float, C.c_float, FP]
        lib.oracle_point_in_tet.argtypes = [FP] * 5
        lib.oracle_closest_point_triangle.argtypes = [FP] * 6
        _lib = lib
    return _lib


def ref_available():
    return os.path.exists(REF_LIB)


def load_ref():
    global _ref
    if _ref is None:
        lib = C.CDLL(REF_LIB)
        FP, IP = C.POINTER(C.c_float), C.POINTER(C.c_int32)
        lib.ref_neohookean_vertex.argtypes = [C.c_int, C.c_int, FP, FP, IP, FP, FP, IP, IP, IP, C.c_float, C.c_float,
                                              C.c_float, C.c_float, C.c_float, C.c_int, FP, FP]
        lib.ref_neohookean_vertex.restype = None
        lib.ref_collision_per_collision.argtypes = [FP, FP, FP, FP, C.c_int] + [C.c_float] * 6 + [FP, FP]
        lib.ref_collision_per_collision.restype = None
        lib.ref_boundary.argtypes = [FP, FP, FP, C.c_float, C.c_float, C.c_float, C.c_float, FP, FP]
        lib.ref_boundary.restype = None
        lib.ref_solve3x3.argtypes = [FP, FP, FP]
        lib.ref_damp_velocity.argtypes = [FP, C.c_float, C.c_float, C.c_float]; lib.ref_damp_velocity.restype = None
        if hasattr(lib, "ref_cloth_material_vertex"):
            lib.ref_cloth_bending_forms.argtypes = [C.c_int, FP, C.c_int, IP, C.c_int, IP, FP]; lib.ref_cloth_bending_forms.restype = None
            lib.ref_cloth_material_vertex.argtypes = [C.c_int, FP, FP, C.c_int, IP, FP, FP, IP, IP, IP, C.c_int, IP, FP, IP, IP, IP] + \
                [C.c_float] * 7 + [C.c_int, C.c_int, FP, FP]
            lib.ref_cloth_material_vertex.restype = None
        if hasattr(lib, "ref_boundary_bowl_cpu"):
            lib.ref_boundary_bowl_cpu.argtypes = [FP, FP, FP, C.c_float, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, FP, FP]; lib.ref_boundary_bowl_cpu.restype = None
        if hasattr(lib, "ref_boundary_cpu"):
            lib.ref_boundary_cpu.argtypes = [FP, FP, FP, FP, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, FP, FP]; lib.ref_boundary_cpu.restype = None
        if hasattr(lib, "ref_cloth_merge_groups"):
            lib.ref_cloth_merge_groups.argtypes = [C.c_int, IP, IP, IP, IP, IP, IP, IP]; lib.ref_cloth_merge_groups.restype = C.c_int
        if hasattr(lib, "ref_merge_groups"):
            lib.ref_merge_groups.argtypes = [C.c_int, IP, IP, IP, IP, IP, IP]; lib.ref_merge_groups.restype = C.c_int
        if hasattr(lib, "ref_collider_sequence"):
            lib.ref_collider_sequence.argtypes = [C.c_int, C.c_int, IP, FP, C.c_int, C.c_int, C.c_int, IP, C.c_int, C.c_int, FP, IP]; lib.ref_collider_sequence.restype = None
        if hasattr(lib, "ref_cloth_deformer"):
            lib.ref_cloth_deformer.argtypes = [C.c_int, C.c_int, FP, C.c_int, IP, FP, FP, C.c_double, C.c_float, C.c_float, C.c_int, FP, C.c_float, FP]; lib.ref_cloth_deformer.restype = None
        if hasattr(lib, "ref_deformer"):
            lib.ref_deformer.argtypes = [C.c_int, C.c_int, FP, C.c_int, IP, FP, FP, C.c_double, C.c_double, C.c_int, C.POINTER(C.c_double), FP]; lib.ref_deformer.restype = None
        if hasattr(lib, "ref_accelerate"):
            lib.ref_accelerate.argtypes = [C.c_int, FP, FP, C.c_float]; lib.ref_accelerate.restype = None
        if hasattr(lib, "ref_accelerator_omegas"):
            lib.ref_accelerator_omegas.argtypes = [C.c_int, C.c_float, C.c_int, FP]; lib.ref_accelerator_omegas.restype = None
        if hasattr(lib, "ref_inertia_terms"):
            lib.ref_inertia_terms.argtypes = [C.c_int, FP, FP, FP, C.c_float, FP, FP]; lib.ref_inertia_terms.restype = None
        if hasattr(lib, "ref_vertex_step_tail"):
            lib.ref_vertex_step_tail.argtypes = [C.c_int, FP, FP, FP, C.c_float, C.c_float, C.c_int, IP]; lib.ref_vertex_step_tail.restype = None
        if hasattr(lib, "ref_prepare_collision_data"):
            UP8 = C.POINTER(C.c_uint8)
            lib.ref_prepare_collision_data.argtypes = [C.c_int, IP, IP, IP, IP, IP, IP, UP8, UP8, UP8, IP, IP, C.c_int, IP, C.c_float, IP, IP, C.c_int64, IP, IP, C.c_int64, UP8]
            lib.ref_prepare_collision_data.restype = C.c_int64
        if hasattr(lib, "ref_step_trace"):
            lib.ref_step_trace.argtypes = [IP, C.c_int, IP, IP, C.POINTER(C.c_uint8), C.c_int, IP, IP, IP, C.c_int64, IP, C.c_int64, IP, IP, C.POINTER(C.c_uint8)]; lib.ref_step_trace.restype = C.c_int64
        if hasattr(lib, "ref_cloth_bounds"):
            UP = C.POINTER(C.c_uint8)
            lib.ref_cloth_bounds.argtypes = [C.c_int, C.c_int, C.c_int, IP, IP, IP, IP, C.c_float, C.c_float, C.c_float, FP, IP, FP, IP, FP, IP, UP, FP, UP]
            lib.ref_cloth_bounds.restype = None
        if hasattr(lib, "ref_cloth_sewing"):
            lib.ref_cloth_sewing.argtypes = [C.c_int, FP, C.c_int, IP, FP, C.c_float, C.c_int, FP, FP]; lib.ref_cloth_sewing.restype = None
            lib.ref_cloth_attachments.argtypes = [C.c_int, C.c_int, IP, FP, C.c_int, IP, IP, FP, FP, FP, IP]; lib.ref_cloth_attachments.restype = None
        if hasattr(lib, "ref_contact_cpu"):
            lib.ref_contact_cpu.argtypes = [FP, FP, FP, FP, C.c_int] + [C.c_float] * 6 + [FP, FP]; lib.ref_contact_cpu.restype = C.c_float
        if hasattr(lib, "ref_cloth_step_sequence"):
            lib.ref_cloth_step_sequence.argtypes = [C.c_int, FP, FP, FP, C.c_int, IP, FP, C.c_int, FP, FP, FP, FP]; lib.ref_cloth_step_sequence.restype = None
            lib.ref_cloth_truncate.argtypes = [C.c_int, FP, FP, C.c_int, IP, FP, IP]; lib.ref_cloth_truncate.restype = None
        if hasattr(lib, "ref_step_sequence"):
            lib.ref_step_sequence.argtypes = [C.c_int, FP, FP, FP, C.c_int, IP, C.c_int, FP, FP, FP, FP, FP, FP]; lib.ref_step_sequence.restype = None
        if hasattr(lib, "ref_cloth_check_feasible_region"):
            lib.ref_cloth_check_feasible_region.argtypes = [C.c_int, FP, C.c_int, IP, IP, C.c_int, IP, IP, IP, FP, C.c_int, C.c_int, C.c_float]
        if hasattr(lib, "ref_check_feasible_region"):
            lib.ref_check_feasible_region.argtypes = [C.c_int, FP, C.c_int, IP, IP, IP, C.c_int, IP, IP, FP, C.c_int, C.c_int, C.c_float]
        if hasattr(lib, "ref_tet_traverse"):
            lib.ref_tet_traverse.argtypes = [C.c_int, FP, C.c_int, IP, IP, IP, FP, FP, C.c_float, C.c_int, C.c_int, C.c_int, C.c_float, IP, IP]
        if hasattr(lib, "ref_cloth_vf_contact"):
            lib.ref_cloth_vf_contact.argtypes = [FP, FP, FP, FP, C.c_int, FP, FP, IP, C.c_int, FP, FP, C.c_int, C.c_int, FP, FP]
            lib.ref_cloth_vf_contact.restype = C.c_float
            lib.ref_cloth_ee_contact.argtypes = [FP, FP, FP, FP, C.c_int, IP, C.c_int, FP, FP, C.c_int, IP, C.c_int, FP, FP, C.c_float, C.c_float,
                                                 C.c_int, C.c_int, FP, FP]
            lib.ref_cloth_ee_contact.restype = C.c_float
            lib.ref_cloth_boundary.argtypes = [FP, FP, FP, FP, FP, C.c_int, FP, FP]; lib.ref_cloth_boundary.restype = None
        if hasattr(lib, "ref_neohookean_tet_energy"):
            lib.ref_neohookean_tet_energy.argtypes = [C.c_float, C.c_float, C.c_float, C.c_int, IP, FP, FP]; lib.ref_neohookean_tet_energy.restype = C.c_float
        lib.ref_ccd_tri_point.argtypes = [FP, FP, FP, FP, FP, FP]
        lib.ref_cubic_roots.argtypes = [FP, C.c_float, C.c_float, FP]
        lib.ref_point_in_tet.argtypes = [FP] * 5
        lib.ref_closest_point_triangle.argtypes = [FP] * 6
        lib.ref_closest_points_segments.argtypes = [FP] * 5; lib.ref_closest_points_segments.restype = None
        lib.ref_cloth_contact_repulsion.argtypes = [C.c_float] * 4 + [C.c_int, FP, FP]; lib.ref_cloth_contact_repulsion.restype = None
        _ref = lib
    return _ref


class OracleScene(SceneBase):
    _prefix = "oracle_"

    def __init__(self, phys=None, col=None, threads=0):
        self._lib = load_oracle()
        phys = phys or default_physics_params(); col = col or default_collision_params()
        self.h = C.c_void_p(self._lib.oracle_create(C.byref(phys), C.byref(col), threads))

    def close(self):
        if self.h is not None:
            self._lib.oracle_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def num_threads(self):
        return self._lib.oracle_num_threads(self.h)

    def prepared_raw(self):
        """the prepared collision data in the order prepare_collision_data built it: (active pairs per group [list of [n, 2]],
        per mesh (relOff [nV + 1], rows [n, 5], mask [nV], group of every vertex [nV]), per mesh the detection results of its
        surface vertices (has, fromCCD, pathFound, hitMesh, hitFace))"""
        L = self._lib
        IPc = C.POINTER(C.c_int32); UPc = C.POINTER(C.c_uint8)
        L.oracle_get_active_list_raw.restype = C.c_int64; L.oracle_get_active_list_raw.argtypes = [C.c_void_p, C.c_int, IPc, C.c_int64]
        L.oracle_get_relations_raw.restype = C.c_int64; L.oracle_get_relations_raw.argtypes = [C.c_void_p, C.c_int, IPc, IPc, C.c_int64, UPc, IPc]
        L.oracle_get_col_results.argtypes = [C.c_void_p, C.c_int, UPc, UPc, UPc, IPc, IPc]
        ip = lambda a: a.ctypes.data_as(IPc); up = lambda a: a.ctypes.data_as(UPc)
        active = []
        for g in range(self._call("num_parallel_groups")):
            n = L.oracle_get_active_list_raw(self.h, g, None, 0)
            a = np.zeros(max(2 * n, 1), np.int32)
            L.oracle_get_active_list_raw(self.h, g, ip(a), 2 * n)
            active.append(a[: 2 * n].reshape(-1, 2))
        rel, res = [], []
        for m in range(self.num_meshes):
            nV, nS = self.mesh_sizes(m)[0], self.mesh_sizes(m)[2]
            n = L.oracle_get_relations_raw(self.h, m, None, None, 0, None, None)
            off = np.zeros(nV + 1, np.int32); rows = np.zeros((max(n, 1), 5), np.int32); mask = np.zeros(nV, np.uint8); grp = np.zeros(nV, np.int32)
            L.oracle_get_relations_raw(self.h, m, ip(off), ip(rows), n, up(mask), ip(grp))
            rel.append((off, rows[:n], mask, grp))
            r = [np.zeros(nS, np.uint8) for _ in range(3)] + [np.zeros(nS, np.int32) for _ in range(2)]
            L.oracle_get_col_results(self.h, m, up(r[0]), up(r[1]), up(r[2]), ip(r[3]), ip(r[4]))
            res.append(tuple(r))
        return active, rel, res

    @property
    def cur_time(self):
        fn = self._lib.oracle_cur_time; fn.restype = C.c_double; fn.argtypes = [C.c_void_p]
        return float(fn(self.h))

    def trace_begin(self):
        """start recording the schedule of run_step (threads=1 scenes only): see oracle/vbd_oracle.h, Scene::trace"""
        rc = self._lib.oracle_trace_begin(self.h)
        if rc != 0:
            raise RuntimeError("oracle_trace_begin: the scene must be single-threaded")

    def trace_read(self):
        """the events recorded since trace_begin: int32 [n, 4] = (code, a, b, c)"""
        fn = self._lib.oracle_trace_read; fn.restype = C.c_int64; fn.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int64]
        n = fn(self.h, None, 0)
        out = np.zeros(max(n, 1), np.int32)
        fn(self.h, out.ctypes.data_as(C.POINTER(C.c_int32)), n)
        return out[:n].reshape(-1, 4)

    def neohookean_forces(self, mesh):
        """elastic force of every vertex at the current positions (accumlateMaterialForce: no damping)"""
        out = np.zeros((self.mesh_sizes(mesh)[0], 3), np.float32)
        fn = self._lib.oracle_neohookean_forces; fn.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_float)]; fn.restype = C.c_int
        fn(self.h, int(mesh), out.ctypes.data_as(C.POINTER(C.c_float)))
        return out

    def phase_begin_substep(self):
        self._call("phase_begin_substep")

    def phase_prepare(self):
        self._call("phase_prepare")

    def phase_solve_group(self, g, apply_friction=True):
        self._call("phase_solve_group", int(g), int(apply_friction))

    def phase_intermediate_cd(self):
        self._call("phase_intermediate_cd")

    def phase_end_substep(self):
        self._call("phase_end_substep")

    def set_prev_positions(self, mesh, prev):
        p = _f32(prev)
        self._call("set_prev_positions", mesh, _fp(p))

    def set_brute_force(self, on=True):
        self._call("set_brute_force", int(on))

    def vertex_merit_energy(self, mesh, v, x):
        """(inertia, elastic) parts of evaluateVertexMeritEnergy with vertex v at x"""
        x = _f32(x); a = C.c_float(); b = C.c_float()
        self._call("vertex_merit_energy", mesh, int(v), _fp(x), C.byref(a), C.byref(b))
        return a.value, b.value

    def dminv(self, mesh):
        a = np.empty((self.mesh_sizes(mesh)[1], 9), np.float32)
        self._call("get_dminv", mesh, _fp(a)); return a

    def rest_volumes(self, mesh):
        a = np.empty(self.mesh_sizes(mesh)[1], np.float32)
        self._call("get_rest_volumes", mesh, _fp(a)); return a

    def vertex_tet_csr(self, mesh):
        nV, nT = self.mesh_sizes(mesh)[:2]
        infos = np.empty(2 * nV, np.int32); t = np.empty(4 * nT, np.int32); o = np.empty(4 * nT, np.int32)
        self._call("get_vertex_tet_csr", mesh, _ip(infos), _ip(t), _ip(o))
        return infos, t, o

    def tet_topology(self, mesh):
        """(tets [nT,4], neighbour tet across face i [nT,4], xor of each tet's vertex ids [nT])"""
        nT = self.mesh_sizes(mesh)[1]
        t = np.empty((nT, 4), np.int32); nb = np.empty((nT, 4), np.int32); x = np.empty(nT, np.int32)
        self._call("get_tet_topology", mesh, _ip(t), _ip(nb), _ip(x))
        return t, nb, x

    def surface_topology(self, mesh):
        """(neighbour face across each edge [nF,3], tet vertex -> surface vertex [nV], CSR offsets [nSV+1], neighbour tet-mesh vertex ids)"""
        nV, nT, nSV, nSF = self.mesh_sizes(mesh)
        n = C.c_int64()
        self._call("get_surface_topology", mesh, None, None, None, None, C.byref(n))
        f3 = np.empty((nSF, 3), np.int32); t2s = np.empty(nV, np.int32); off = np.empty(nSV + 1, np.int32); nb = np.empty(max(n.value, 1), np.int32)
        self._call("get_surface_topology", mesh, _ip(f3), _ip(t2s), _ip(off), _ip(nb), C.byref(n))
        return f3, t2s, off, nb[: n.value]

    def check_feasible_region(self, mesh, p, face, point_type, eps):
        q = _f32(p)
        return bool(self._call("check_feasible_region", mesh, _fp(q), int(face), int(point_type), float(eps)))

    def tet_traverse(self, mesh, origin, direction, max_dis, start_tet, start_face, target_tet, eps):
        """tetrahedralTraverseTo: (reached, tets traversed)"""
        o = _f32(origin); d = _f32(direction); n = np.zeros(1, np.int32)
        r = self._call("tet_traverse", mesh, _fp(o), _fp(d), float(max_dis), int(start_tet), int(start_face), int(target_tet), float(eps), _ip(n))
        return bool(r), int(n[0])

    def vertex_material_force_hessian(self, mesh, v):
        f = np.empty(3, np.float32); h = np.empty(9, np.float32)
        self._call("vertex_material_force_hessian", mesh, v, _fp(f), _fp(h))
        return f, h
