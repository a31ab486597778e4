"""ctypes binding of the C ABI declared in include/gaia_b200.h (libgaia_b200.so).

This is the reference-side binding a maintainer would use from Python; the C++ host adapter
(gaia_b200/host/VBDPhysicsB200.h) binds the same symbols.  There is no CPU fallback: if the
shared library (built in-tree by __graft_entry__.build()) is missing, loading raises.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libgaia_b200.so")

c_f3 = C.c_float * 3


class PhysicsParams(C.Structure):
    """gaia_physics_params (BasePhysicsParams + VBDPhysicsParameters of the reference)."""
    _fields_ = [
        ("gravity", c_f3), ("timeStep", C.c_float), ("numSubsteps", C.c_int32), ("iterations", C.c_int32),
        ("usePlaneGround", C.c_int32), ("useBowlGround", C.c_int32), ("checkAndUpdateWorldBounds", C.c_int32),
        ("worldBoundsMin", c_f3), ("worldBoundsMax", c_f3), ("boundaryFrictionDynamic", C.c_float),
        ("dcdTetMeshSceneBVHRebuildSteps", C.c_int32), ("dcdSurfaceSceneBVHRebuildSteps", C.c_int32),
        ("ccdBVHRebuildSteps", C.c_int32),
        ("stepSize", C.c_float), ("stepSizeGD", C.c_float), ("associateGravityWithInertia", C.c_int32),
        ("collisionStiffness", C.c_float), ("intermediateCollisionIterations", C.c_int32),
        ("boundaryCollisionStiffness", C.c_float), ("boundaryFrictionEpsV", C.c_float),
        ("frictionStartIter", C.c_int32), ("useAccelerator", C.c_int32), ("acceleratorRho", C.c_float),
        ("collisionOffHeight", C.c_float),
    ]


class CollisionParams(C.Structure):
    """gaia_collision_params (CollisionDetectionParamters of the reference)."""
    _fields_ = [
        ("allowCCD", C.c_int32), ("allowDCD", C.c_int32), ("checkFeasibleRegion", C.c_int32),
        ("checkTetTraverse", C.c_int32), ("handleSelfCollision", C.c_int32),
        ("stopTraversingAfterPassingQueryPoint", C.c_int32), ("tetrahedralTraverseForNonSelfIntersection", C.c_int32),
        ("useStaticTraverse", C.c_int32), ("loopLessTraverse", C.c_int32), ("shiftQueryPointToCenter", C.c_int32),
        ("centerShiftLevel", C.c_float), ("maxNumberOfBVHQuery", C.c_int32), ("feasibleRegionEpsilon", C.c_float),
        ("rayTriIntersectionEPSILON", C.c_float), ("maxSearchDistanceMultiplier", C.c_float),
    ]


class ObjectParams(C.Structure):
    """gaia_object_params (ObjectParams + ObjectParamsVBD + ObjectParametersVBDNeoHookean)."""
    _fields_ = [
        ("scale", c_f3), ("rotation", c_f3), ("translation", c_f3), ("initialVelocity", c_f3),
        ("density", C.c_float), ("hasNoGravZone", C.c_int32), ("noGravZoneThreshold", C.c_float),
        ("maxVelocityMagnitude", C.c_float), ("frameToAppear", C.c_int32),
        ("frictionDynamic", C.c_float), ("frictionEpsV", C.c_float), ("exponentialVelDamping", C.c_float),
        ("constantVelDamping", C.c_float), ("initializationType", C.c_int32), ("initRatio_g", C.c_float),
        ("miu", C.c_float), ("lmbd", C.c_float), ("dampingHydrostatic", C.c_float), ("dampingDeviatoric", C.c_float),
    ]


class Contact(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("queryMesh", "queryVertex", "hitMesh", "hitTet", "hitFace", "fromCCD",
                                         "shortestPathFound", "closestPointType")]


class Stats(C.Structure):
    _fields_ = [(n, C.c_double) for n in ("timeCsmpInitialStep", "timeCsmpMaterialSolve", "timeCsmpUpdateVelocity",
                                          "timeCsmpUpdatingBVHDCD", "timeCsmpColDetectDCD", "timeCsmpUpdatingBVHCCD",
                                          "timeCsmpColDetectCCD", "timeCsmpFrame", "solveKernelMs")] + \
               [(n, C.c_int64) for n in ("vertexSolves", "colourSweeps", "substeps", "frames", "kernelLaunches",
                                         "algorithmicBytes", "numContactsDCD", "numContactsCCD")]


CONTACT_DTYPE = np.dtype([(n, np.int32) for n, _ in Contact._fields_])


def default_physics_params(**kw):
    p = PhysicsParams()
    p.gravity = c_f3(0, -10, 0); p.timeStep = 0.01666666666666666666; p.numSubsteps = 20; p.iterations = 1
    p.usePlaneGround = 1; p.useBowlGround = 0; p.checkAndUpdateWorldBounds = 0
    p.worldBoundsMin = c_f3(-1e7, -1e7, -1e7); p.worldBoundsMax = c_f3(1e7, 1e7, 1e7)
    p.boundaryFrictionDynamic = 0.1
    p.dcdTetMeshSceneBVHRebuildSteps = 16; p.dcdSurfaceSceneBVHRebuildSteps = 3; p.ccdBVHRebuildSteps = 7
    p.stepSize = 1.0; p.stepSizeGD = 1e-3; p.associateGravityWithInertia = 1
    p.collisionStiffness = 1e5; p.intermediateCollisionIterations = -1
    p.boundaryCollisionStiffness = 1e5; p.boundaryFrictionEpsV = 1.0; p.frictionStartIter = 0
    p.useAccelerator = 0; p.acceleratorRho = 0.93; p.collisionOffHeight = 1e7
    _apply(p, kw)
    return p


def default_collision_params(**kw):
    c = CollisionParams()
    c.allowCCD = 1; c.allowDCD = 1; c.checkFeasibleRegion = 1; c.checkTetTraverse = 1; c.handleSelfCollision = 1
    c.stopTraversingAfterPassingQueryPoint = 1; c.tetrahedralTraverseForNonSelfIntersection = 1
    c.useStaticTraverse = 1; c.loopLessTraverse = 0; c.shiftQueryPointToCenter = 1; c.centerShiftLevel = 0.01
    c.maxNumberOfBVHQuery = 500; c.feasibleRegionEpsilon = 1e-2; c.rayTriIntersectionEPSILON = 1e-10
    c.maxSearchDistanceMultiplier = 1.8
    _apply(c, kw)
    return c


def default_object_params(**kw):
    o = ObjectParams()
    o.scale = c_f3(1, 1, 1); o.rotation = c_f3(0, 0, 0); o.translation = c_f3(0, 0, 0); o.initialVelocity = c_f3(0, 0, 0)
    o.density = 1.0; o.hasNoGravZone = 0; o.noGravZoneThreshold = 0.0; o.maxVelocityMagnitude = -1.0
    o.frameToAppear = -1; o.frictionDynamic = 0.1; o.frictionEpsV = 0.01; o.exponentialVelDamping = 1.0
    o.constantVelDamping = 0.0; o.initializationType = 0; o.initRatio_g = 0.5
    o.miu = 1e4; o.lmbd = 1e5; o.dampingHydrostatic = 0.0; o.dampingDeviatoric = 0.0
    _apply(o, kw)
    return o


def _apply(st, kw):
    names = {n for n, _ in st._fields_}
    for k, v in kw.items():
        if k not in names:
            raise KeyError(f"{type(st).__name__} has no field {k!r}")
        if isinstance(v, (list, tuple, np.ndarray)):
            setattr(st, k, c_f3(*[float(x) for x in v]))
        else:
            setattr(st, k, v)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


class GaiaError(RuntimeError):
    pass


def bind_scene_api(lib, prefix):
    """Declare argtypes for the functions shared by the product ABI and the oracle's shadow ABI."""
    P = C.c_void_p
    FP, IP = C.POINTER(C.c_float), C.POINTER(C.c_int32)
    sig = {
        "add_tet_mesh": ([P, C.c_int32, C.c_int32, FP, IP, C.POINTER(ObjectParams), C.c_int32, IP, IP, C.c_int32, IP], C.c_int),
        "add_rotator": ([P, C.c_int, IP, C.c_int32, FP, FP, C.c_double, C.c_double], C.c_int),
        "add_translator": ([P, C.c_int, IP, C.c_int32, FP, C.c_double], C.c_int),
        "finalize": ([P], C.c_int), "run_frames": ([P, C.c_int32], C.c_int), "run_iterations": ([P, C.c_int32], C.c_int),
        "frame_id": ([P], C.c_int), "num_meshes": ([P], C.c_int),
        "mesh_sizes": ([P, C.c_int, IP, IP, IP, IP], C.c_int),
        "get_positions": ([P, C.c_int, FP], C.c_int), "get_prev_positions": ([P, C.c_int, FP], C.c_int),
        "get_velocities": ([P, C.c_int, FP], C.c_int), "get_inertia": ([P, C.c_int, FP], C.c_int),
        "set_state": ([P, C.c_int, FP, FP], C.c_int), "get_surface_positions": ([P, C.c_int, FP], C.c_int),
        "get_world_bounds": ([P, FP, FP], C.c_int), "num_parallel_groups": ([P], C.c_int),
        "get_parallel_group": ([P, C.c_int, IP, IP], C.c_int),
        "get_surface_vertex_ids": ([P, C.c_int, IP], C.c_int), "get_surface_faces": ([P, C.c_int, IP], C.c_int),
        "get_vertex_masses": ([P, C.c_int, FP], C.c_int),
        "detect_collisions": ([P, C.c_int, C.c_int], C.c_int),
        "get_contacts": ([P, C.POINTER(Contact), C.c_int64, C.POINTER(C.c_int64)], C.c_int),
        "enable_mesh": ([P, C.c_int, C.c_int, C.c_int], C.c_int),
        "get_stats": ([P, C.POINTER(Stats)], C.c_int), "reset_stats": ([P], C.c_int),
        "get_bvh_leaf_order": ([P, C.c_int, IP, C.c_int64, C.POINTER(C.c_int64)], C.c_int),
        "get_active_list": ([P, C.c_int, IP, C.c_int64, C.POINTER(C.c_int64)], C.c_int),
        "get_relations": ([P, C.c_int, IP, C.c_int64, C.POINTER(C.c_int64)], C.c_int),
        "set_partition_mode": ([P, C.c_int32, C.c_int32], C.c_int), "set_ghosts": ([P, C.c_int, IP, C.c_int32], C.c_int),
        "set_line_search": ([P, C.c_int32, C.c_float, C.c_float, C.c_int32], C.c_int), "line_search_shrinks": ([P], C.c_int64),
        "set_bowl_ground": ([P, FP, C.c_float, C.c_int32], C.c_int),
    }
    for name, (args, res) in sig.items():
        fn = getattr(lib, prefix + name)
        fn.argtypes = args
        fn.restype = res


class SceneBase:
    """Calls shared by the product (`gaia_b200_*`) and the oracle shadow ABI (`oracle_*`)."""
    _lib = None
    _prefix = ""
    h = None

    def _call(self, name, *args):
        rc = getattr(self._lib, self._prefix + name)(self.h, *args)
        if rc < 0:
            raise GaiaError(f"{self._prefix}{name} failed: {rc} {self.last_error()}")
        return rc

    def last_error(self):
        return ""

    def add_tet_mesh(self, xyz, tets, obj, colors, fixed=()):
        xyz = _f32(xyz).reshape(-1, 3); tets = _i32(tets).reshape(-1, 4)
        offs = np.zeros(len(colors) + 1, np.int32)
        offs[1:] = np.cumsum([len(c) for c in colors])
        cv = _i32(np.concatenate([np.asarray(c, np.int32) for c in colors]) if len(colors) else np.zeros(0, np.int32))
        fx = _i32(np.asarray(fixed, np.int32))
        return self._call("add_tet_mesh", xyz.shape[0], tets.shape[0], _fp(xyz), _ip(tets), C.byref(obj),
                          len(colors), _ip(offs), _ip(cv), len(fx), _ip(fx))

    def add_rotator(self, mesh, verts, origin, axis, angular_velocity, end_time=-1.0):
        v = _i32(verts); o = _f32(origin); a = _f32(axis)
        self._call("add_rotator", mesh, _ip(v), len(v), _fp(o), _fp(a), float(angular_velocity), float(end_time))

    def add_translator(self, mesh, verts, speed, end_time=-1.0):
        v = _i32(verts); sp = _f32(speed)
        self._call("add_translator", mesh, _ip(v), len(v), _fp(sp), float(end_time))

    def set_partition_mode(self, rank, world):
        self._call("set_partition_mode", int(rank), int(world))

    def set_ghosts(self, mesh, verts):
        v = _i32(verts)
        self._call("set_ghosts", mesh, _ip(v), len(v))

    def finalize(self):
        self._call("finalize")

    def run_frames(self, n=1):
        self._call("run_frames", n)

    def set_line_search(self, enabled=True, tau=0.5, tau_gd=0.1, max_iters=16):
        """backtracking line search on the vertex merit energy (APPLY_LOCAL_LINE_SEARCH, M/VBD/VBDPhysics.cpp:25, 2467-2474)"""
        self._call("set_line_search", int(bool(enabled)), float(tau), float(tau_gd), int(max_iters))

    def line_search_shrinks(self):
        """trial steps the line search rejected so far (each one shrank alpha by tau)"""
        return int(getattr(self._lib, self._prefix + "line_search_shrinks")(self.h))

    def set_bowl_ground(self, center, radius, cap=False):
        """centre, radius and cap of the bowl ground of a scene created with useBowlGround=1 (the reference's bowlCenter / bowlRadius / bowlCap)"""
        c = _f32(np.asarray(center, np.float32))
        self._call("set_bowl_ground", _fp(c), float(radius), int(bool(cap)))

    def run_iterations(self, n=1):
        self._call("run_iterations", n)

    @property
    def frame_id(self):
        return self._call("frame_id")

    @property
    def num_meshes(self):
        return self._call("num_meshes")

    def mesh_sizes(self, mesh):
        v = [C.c_int32() for _ in range(4)]
        self._call("mesh_sizes", mesh, *[C.byref(x) for x in v])
        return tuple(x.value for x in v)

    def _getv(self, name, mesh, n):
        out = np.empty((n, 3), np.float32)
        self._call(name, mesh, _fp(out))
        return out

    def positions(self, mesh):
        return self._getv("get_positions", mesh, self.mesh_sizes(mesh)[0])

    def prev_positions(self, mesh):
        return self._getv("get_prev_positions", mesh, self.mesh_sizes(mesh)[0])

    def velocities(self, mesh):
        return self._getv("get_velocities", mesh, self.mesh_sizes(mesh)[0])

    def inertia(self, mesh):
        return self._getv("get_inertia", mesh, self.mesh_sizes(mesh)[0])

    def surface_positions(self, mesh):
        return self._getv("get_surface_positions", mesh, self.mesh_sizes(mesh)[2])

    def set_state(self, mesh, pos=None, vel=None):
        p = _f32(pos) if pos is not None else None
        v = _f32(vel) if vel is not None else None
        self._call("set_state", mesh, _fp(p) if p is not None else None, _fp(v) if v is not None else None)

    def world_bounds(self):
        mn = np.empty(3, np.float32); mx = np.empty(3, np.float32)
        self._call("get_world_bounds", _fp(mn), _fp(mx))
        return mn, mx

    def parallel_groups(self):
        out = []
        for g in range(self._call("num_parallel_groups")):
            n = C.c_int32()
            self._call("get_parallel_group", g, None, C.byref(n))
            a = np.empty(2 * n.value, np.int32)
            self._call("get_parallel_group", g, _ip(a), C.byref(n))
            out.append(a.reshape(-1, 2))
        return out

    def surface_vertex_ids(self, mesh):
        a = np.empty(self.mesh_sizes(mesh)[2], np.int32)
        self._call("get_surface_vertex_ids", mesh, _ip(a))
        return a

    def surface_faces(self, mesh):
        a = np.empty((self.mesh_sizes(mesh)[3], 3), np.int32)
        self._call("get_surface_faces", mesh, _ip(a))
        return a

    def vertex_masses(self, mesh):
        a = np.empty(self.mesh_sizes(mesh)[0], np.float32)
        self._call("get_vertex_masses", mesh, _fp(a))
        return a

    def detect_collisions(self, dcd=True, ccd=True):
        self._call("detect_collisions", int(dcd), int(ccd))

    def contacts(self):
        n = C.c_int64()
        self._call("get_contacts", None, 0, C.byref(n))
        arr = np.zeros(max(n.value, 1), CONTACT_DTYPE)
        self._call("get_contacts", arr.ctypes.data_as(C.POINTER(Contact)), n.value, C.byref(n))
        return arr[: n.value]

    def bvh_leaf_order(self, which):
        n = C.c_int64()
        self._call("get_bvh_leaf_order", which, None, 0, C.byref(n))
        a = np.empty(n.value, np.int32)
        self._call("get_bvh_leaf_order", which, _ip(a), n.value, C.byref(n))
        return a

    def active_list(self, group):
        """(n, 2) int32 (queryMesh, queryVertex) of the contacts active in one parallel group, ascending (row a15)."""
        n = C.c_int64()
        self._call("get_active_list", group, None, 0, C.byref(n))
        a = np.empty((n.value, 2), np.int32)
        if n.value:
            self._call("get_active_list", group, _ip(a), n.value, C.byref(n))
        return a

    def relations(self, mesh):
        """(n, 4) int32 (vertex, queryMesh, queryVertex, face-vertex order): face-side contact relations of a mesh (row a15)."""
        n = C.c_int64()
        self._call("get_relations", mesh, None, 0, C.byref(n))
        a = np.empty((n.value, 4), np.int32)
        if n.value:
            self._call("get_relations", mesh, _ip(a), n.value, C.byref(n))
        return a

    def enable_mesh(self, mesh, collide=True, solve=True):
        self._call("enable_mesh", mesh, int(collide), int(solve))

    def stats(self):
        s = Stats()
        self._call("get_stats", C.byref(s))
        return {n: getattr(s, n) for n, _ in Stats._fields_}

    def reset_stats(self):
        self._call("reset_stats")


_lib = None


def load_library(path=None):
    """Load libgaia_b200.so; raises if the CUDA extension has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise GaiaError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(there is no CPU fallback)")
    lib = C.CDLL(path)
    lib.gaia_b200_abi_version.restype = C.c_int
    lib.gaia_b200_last_error.restype = C.c_char_p
    lib.gaia_b200_last_error.argtypes = [C.c_void_p]
    lib.gaia_b200_create.argtypes = [C.POINTER(PhysicsParams), C.POINTER(CollisionParams), C.c_int, C.POINTER(C.c_void_p)]
    lib.gaia_b200_create.restype = C.c_int
    lib.gaia_b200_destroy.argtypes = [C.c_void_p]
    lib.gaia_b200_destroy.restype = None
    bind_scene_api(lib, "gaia_b200_")
    P = C.c_void_p
    FP, IP = C.POINTER(C.c_float), C.POINTER(C.c_int32)
    lib.gaia_b200_synchronize.argtypes = [P]
    lib.gaia_b200_total_vertices.argtypes = [P]; lib.gaia_b200_total_vertices.restype = C.c_int64
    lib.gaia_b200_get_all_positions.argtypes = [P, FP]
    lib.gaia_b200_set_all_state.argtypes = [P, FP, FP]
    lib.gaia_b200_get_sorted_ids.argtypes = [P, C.c_int, IP, C.c_int32, IP]
    lib.gaia_b200_halo_ipc_handles.argtypes = [P, C.c_void_p]
    lib.gaia_b200_halo_add_peer.argtypes = [P, C.c_int32, C.c_void_p, C.c_int32, IP, IP]
    lib.gaia_b200_halo_finalize.argtypes = [P]
    lib.gaia_b200_halo_set_world.argtypes = [P, C.c_int32, C.c_void_p]
    lib.gaia_b200_set_time.argtypes = [P, C.c_int32, C.c_double]
    lib.gaia_b200_cur_time.argtypes = [P]; lib.gaia_b200_cur_time.restype = C.c_double
    lib.gaia_b200_set_profiling.argtypes = [P, C.c_int]
    lib.gaia_b200_set_solver_lanes.argtypes = [P, C.c_int]
    lib.gaia_b200_set_multi_colour_launch.argtypes = [P, C.c_int]
    lib.gaia_b200_set_lazy_refit_margin.argtypes = [P, C.c_float]
    lib.gaia_b200_get_layout_bytes.argtypes = [P, C.POINTER(C.c_int64)]
    lib.gaia_b200_total_surface_vertices.argtypes = [P]; lib.gaia_b200_total_surface_vertices.restype = C.c_int64
    lib.gaia_b200_surface_snapshot_begin.argtypes = [P, C.c_int]
    lib.gaia_b200_surface_snapshot_wait.argtypes = [P, C.c_int, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_int64)]
    lib.gaia_b200_color_graph.argtypes = [C.c_int32, C.c_int64, IP, C.c_int32, C.c_float, IP, IP, IP]
    lib.gaia_b200_color_tet_mesh.argtypes = [C.c_int32, C.c_int32, IP, C.c_int32, C.c_float, IP, IP, IP]
    lib.gaia_b200_color_tri_mesh.argtypes = [C.c_int32, C.c_int32, IP, C.c_int32, C.c_int32, C.c_float, IP, IP, IP]
    lib.gaia_b200_color_graph_device.argtypes = [C.c_int, C.c_int32, C.c_int64, IP, C.c_int32, C.c_float, IP, IP, IP]
    lib.gaia_b200_color_graph_device_noalloc.argtypes = [C.c_int, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int32]
    I64P, U8P = C.POINTER(C.c_int64), C.POINTER(C.c_uint8)
    lib.gaia_b200_partition_slab_owner.argtypes = [C.c_int32, FP, C.c_int32, C.c_int32, IP]
    lib.gaia_b200_partition_create.argtypes = [C.c_int32, C.c_int64, IP, IP, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]
    lib.gaia_b200_partition_destroy.argtypes = [C.c_void_p]; lib.gaia_b200_partition_destroy.restype = None
    lib.gaia_b200_partition_sizes.argtypes = [C.c_void_p, IP, I64P, IP]
    lib.gaia_b200_partition_local_verts.argtypes = [C.c_void_p, IP, U8P]
    lib.gaia_b200_partition_local_tets.argtypes = [C.c_void_p, IP, I64P]
    lib.gaia_b200_partition_peer.argtypes = [C.c_void_p, C.c_int32, IP, IP, IP]
    lib.gaia_b200_partition_peer_lists.argtypes = [C.c_void_p, C.c_int32, IP, IP]
    lib.gaia_b200_partition_restrict_ids.argtypes = [C.c_void_p, C.c_int32, IP, IP]
    lib.gaia_b200_partition_restrict_colors.argtypes = [C.c_void_p, C.c_int32, IP, IP, IP, IP]
    lib.gaia_b200_tet_surface.argtypes = [C.c_int32, C.c_int32, IP, IP, IP, IP, IP]
    lib.gaia_b200_halo_local_pointers.argtypes = [P, C.c_void_p]
    lib.gaia_b200_halo_add_peer_pointers.argtypes = [P, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, IP, IP]
    lib.gaia_b200_cuda_stream.argtypes = [P]; lib.gaia_b200_cuda_stream.restype = C.c_void_p
    for f in ("default_physics_params", "default_collision_params", "default_object_params"):
        getattr(lib, "gaia_b200_" + f).restype = None
    _lib = lib
    return lib


class Scene(SceneBase):
    """A VBD scene on one B200 (mirror of VBDPhysics: initialize -> runStep loop)."""
    _prefix = "gaia_b200_"

    def __init__(self, phys=None, col=None, device=0):
        self._lib = load_library()
        phys = phys or default_physics_params(); col = col or default_collision_params()
        h = C.c_void_p()
        rc = self._lib.gaia_b200_create(C.byref(phys), C.byref(col), device, C.byref(h))
        if rc < 0:
            raise GaiaError(f"gaia_b200_create failed: {rc} {self._lib.gaia_b200_last_error(None).decode()}")
        self.h = h

    def last_error(self):
        return self._lib.gaia_b200_last_error(self.h).decode()

    def close(self):
        if self.h is not None:
            self._lib.gaia_b200_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def synchronize(self):
        self._call("synchronize")

    @property
    def total_vertices(self):
        return self._lib.gaia_b200_total_vertices(self.h)

    def all_positions(self, out=None):
        if out is None:
            out = np.empty((self.total_vertices, 3), np.float32)
        self._call("get_all_positions", _fp(out))
        return out

    def set_all_state(self, pos, vel):
        self._call("set_all_state", _fp(pos), _fp(vel))

    def register_host_buffer(self, arr):
        """Page-lock a C-contiguous numpy array: get / set calls on it then DMA without the staging copy. Keep `arr` alive."""
        assert arr.flags["C_CONTIGUOUS"]
        fn = self._lib.gaia_b200_register_host_buffer
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        rc = fn(self.h, arr.ctypes.data, arr.nbytes)
        if rc < 0:
            raise GaiaError(f"gaia_b200_register_host_buffer failed: {rc} {self.last_error()}")

    def unregister_host_buffer(self, arr):
        fn = self._lib.gaia_b200_unregister_host_buffer
        fn.argtypes = [C.c_void_p, C.c_void_p]
        rc = fn(self.h, arr.ctypes.data)
        if rc < 0:
            raise GaiaError(f"gaia_b200_unregister_host_buffer failed: {rc} {self.last_error()}")

    def sorted_ids(self, mesh, verts):
        v = _i32(verts); out = np.empty(len(v), np.int32)
        self._call("get_sorted_ids", mesh, _ip(v), len(v), _ip(out))
        return out

    def halo_ipc_handles(self):
        buf = C.create_string_buffer(128)
        self._call("halo_ipc_handles", buf)
        return buf.raw

    def halo_add_peer(self, peer_rank, handles, send_mesh_verts, remote_slots):
        mv = _i32(send_mesh_verts).reshape(-1, 2); rs = _i32(remote_slots)
        assert len(mv) == len(rs)
        self._call("halo_add_peer", int(peer_rank), C.create_string_buffer(handles, 128), len(rs), _ip(mv), _ip(rs))

    def halo_set_world(self, handles_by_rank):
        """handles of ALL ranks (rank 0 first): lets the ranks agree, after every detection, whether anyone holds a contact"""
        blob = b"".join(bytes(h) for h in handles_by_rank)
        assert len(blob) == 128 * len(handles_by_rank)
        self._call("halo_set_world", len(handles_by_rank), C.create_string_buffer(blob, len(blob)))

    def halo_finalize(self):
        self._call("halo_finalize")

    def set_time(self, frame_id, cur_time):
        self._call("set_time", int(frame_id), float(cur_time))

    @property
    def cur_time(self):
        return self._lib.gaia_b200_cur_time(self.h)

    def set_profiling(self, on=True):
        self._call("set_profiling", int(on))

    def surface_snapshot_begin(self, slot):
        self._call("surface_snapshot_begin", int(slot))

    def surface_snapshot_wait(self, slot):
        """[nSurfaceVertices, 3] view-copy of ring slot `slot` (all meshes concatenated, writeAllToBinary order)"""
        ptr = C.POINTER(C.c_float)()
        n = C.c_int64()
        self._call("surface_snapshot_wait", int(slot), C.byref(ptr), C.byref(n))
        return np.ctypeslib.as_array(ptr, shape=(n.value,)).reshape(-1, 3).copy()

    def layout_bytes(self):
        a = (C.c_int64 * 6)()
        self._call("get_layout_bytes", a)
        return dict(zip(("records", "neighbours", "vertex_state", "tiles", "pos_reads", "incidences"), [int(x) for x in a]))

    def set_lazy_refit_margin(self, edges):
        """margin of the DCD trees' boxes in typical tet edges (0 = refit at every detection; negative = the adaptive default); before finalize"""
        self._call("set_lazy_refit_margin", float(edges))

    def set_multi_colour_launch(self, enabled):
        """contact-free stretches of the colour loop as one cooperative launch: False / 0 never, True / 1 (default) where the
        host issues the loop, 2 everywhere (no graph replay of per-colour launches); results do not change"""
        self._call("set_multi_colour_launch", int(enabled))

    def set_solver_lanes(self, lanes):
        """0 (default 4), 4 or 8 lanes per vertex; runs compared bit for bit must use the same value"""
        self._call("set_solver_lanes", int(lanes))

    @property
    def cuda_stream(self):
        return self._lib.gaia_b200_cuda_stream(self.h)


# ---------------------------------------------------------------------------------------------- vertex colouring (host code)
def _categories(fn, n, *args):
    colors = np.empty(n, np.int32); off = np.empty(n + 1, np.int32); nodes = np.empty(n, np.int32)
    nc = fn(*args, _ip(colors), _ip(off), _ip(nodes))
    if nc < 0:
        raise GaiaError(f"colouring failed: {nc}")
    return [nodes[off[c]:off[c + 1]].copy() for c in range(nc)]


def _color_flags(balance, algorithm):
    if algorithm not in ("greedy", "mcs"):
        raise ValueError("algorithm: 'greedy' or 'mcs'")
    return (1 if balance else 0) | (2 if algorithm == "mcs" else 0)


def color_graph_nodes(num_nodes, edges, balance=True, goal_max_min_ratio=1.05, algorithm="greedy"):
    """colour id per node (the array the reference tool's GraphColor::get_coloring returns)"""
    e = _i32(np.asarray(edges, np.int32).reshape(-1, 2))
    col = np.zeros(max(int(num_nodes), 1), np.int32)
    rc = load_library().gaia_b200_color_graph(int(num_nodes), len(e), _ip(e), _color_flags(balance, algorithm), float(goal_max_min_ratio),
                                              _ip(col), None, None)
    if rc < 0:
        raise GaiaError(f"gaia_b200_color_graph failed: {rc}")
    return col[:int(num_nodes)]


def color_graph(num_nodes, edges, balance=True, goal_max_min_ratio=1.05, algorithm="greedy"):
    """The reference tool's `greedy` colouring (+ balancing) of an undirected graph given as an [nE, 2] edge list; returns the
    colour categories in the tool's order (what it writes to *.vertexColoring.json)."""
    e = _i32(np.asarray(edges, np.int32).reshape(-1, 2))
    return _categories(load_library().gaia_b200_color_graph, int(num_nodes), int(num_nodes), len(e), _ip(e), _color_flags(balance, algorithm),
                       float(goal_max_min_ratio))


def color_graph_device(num_nodes, edges, balance=True, goal_max_min_ratio=1.05, device=0):
    """colour categories computed on the GPU (gaia_b200_color_graph_device): valid, few, balanced colours - not the tool's"""
    e = _i32(np.asarray(edges, np.int32).reshape(-1, 2))
    fn = load_library().gaia_b200_color_graph_device
    return _categories(lambda *a: fn(int(device), *a), int(num_nodes), int(num_nodes), len(e), _ip(e), 1 if balance else 0, float(goal_max_min_ratio))


NOALLOC_COLOR_MAX_NODES = 1 << 20


def color_graph_device_noalloc(num_nodes, d_offsets, d_adjacency, d_colors, device=0):
    """gaia_b200_color_graph_device_noalloc: d_* are DEVICE addresses (ints) of the caller's CSR graph and colour output; returns
    the number of colours. Allocates nothing on the device (fixed scratch inside the library, NOALLOC_COLOR_MAX_NODES nodes)."""
    msg = C.create_string_buffer(256)
    rc = load_library().gaia_b200_color_graph_device_noalloc(int(device), int(num_nodes), C.c_void_p(int(d_offsets)), C.c_void_p(int(d_adjacency)),
                                                            C.c_void_p(int(d_colors)), msg, 256)
    if rc < 0:
        e = GaiaError(f"gaia_b200_color_graph_device_noalloc failed: {rc} {msg.value.decode()}")
        e.code = rc
        raise e
    return rc


def tet_mesh_edges(n_verts, tets):
    """unique undirected edges of the tet-mesh vertex graph (M/GraphColoring/TetMeshVertexGraph.h)"""
    t = np.asarray(tets, np.int64).reshape(-1, 4)
    e = np.concatenate([t[:, [a, b]] for a in range(4) for b in range(a + 1, 4)])
    e = np.sort(e, axis=1)
    key = np.unique(e[:, 0] * int(n_verts) + e[:, 1])
    return np.stack([key // int(n_verts), key % int(n_verts)], 1).astype(np.int32)


def color_tet_mesh(n_verts, tets, balance=True, goal_max_min_ratio=1.05, algorithm="greedy"):
    t = _i32(np.asarray(tets, np.int32).reshape(-1, 4))
    return _categories(load_library().gaia_b200_color_tet_mesh, int(n_verts), int(n_verts), len(t), _ip(t), _color_flags(balance, algorithm),
                       float(goal_max_min_ratio))


def color_tri_mesh(n_verts, faces, edge_energy=True, balance=True, goal_max_min_ratio=1.05, algorithm="greedy"):
    f = _i32(np.asarray(faces, np.int32).reshape(-1, 3))
    return _categories(load_library().gaia_b200_color_tri_mesh, int(n_verts), int(n_verts), len(f), _ip(f), 1 if edge_energy else 0,
                       _color_flags(balance, algorithm), float(goal_max_min_ratio))


# --------------------------------------------------------------------------------------------- slab partitioning (csrc/partition.cpp)
def partition_slab_owner(verts, world, axis=-1):
    v = _f32(np.asarray(verts, np.float32).reshape(-1, 3))
    owner = np.zeros(len(v), np.int32)
    rc = load_library().gaia_b200_partition_slab_owner(len(v), _fp(v), int(world), int(axis), _ip(owner))
    if rc < 0:
        raise GaiaError(f"gaia_b200_partition_slab_owner failed: {rc}")
    return owner


class SlabPartition:
    """What rank `rank` holds of one global tet mesh: binding of gaia_b200_partition_* (see include/gaia_b200.h)."""

    def __init__(self, n_verts, tets, owner, rank, world):
        lib = load_library()
        t = _i32(np.asarray(tets, np.int32).reshape(-1, 4)); o = _i32(np.asarray(owner, np.int32))
        self._p = C.c_void_p()
        rc = lib.gaia_b200_partition_create(int(n_verts), len(t), _ip(t), _ip(o), int(rank), int(world), C.byref(self._p))
        if rc < 0:
            raise GaiaError(f"gaia_b200_partition_create failed: {rc}")
        self._lib, self.rank, self.world = lib, rank, world
        nv, nt, npeer = C.c_int32(), C.c_int64(), C.c_int32()
        lib.gaia_b200_partition_sizes(self._p, C.byref(nv), C.byref(nt), C.byref(npeer))
        self.local_verts = np.zeros(nv.value, np.int32); owned = np.zeros(nv.value, np.uint8)
        lib.gaia_b200_partition_local_verts(self._p, _ip(self.local_verts), owned.ctypes.data_as(C.POINTER(C.c_uint8)))
        self.owned = owned.astype(bool)
        self.tets = np.zeros((nt.value, 4), np.int32); self.local_tets = np.zeros(nt.value, np.int64)
        lib.gaia_b200_partition_local_tets(self._p, _ip(self.tets), self.local_tets.ctypes.data_as(C.POINTER(C.c_int64)))
        self.send, self.recv = {}, {}
        for i in range(npeer.value):
            q, ns, nr = C.c_int32(), C.c_int32(), C.c_int32()
            lib.gaia_b200_partition_peer(self._p, i, C.byref(q), C.byref(ns), C.byref(nr))
            s_, r_ = np.zeros(max(ns.value, 1), np.int32), np.zeros(max(nr.value, 1), np.int32)
            lib.gaia_b200_partition_peer_lists(self._p, i, _ip(s_), _ip(r_))
            if ns.value:
                self.send[q.value] = s_[:ns.value]
            if nr.value:
                self.recv[q.value] = r_[:nr.value]

    def restrict_ids(self, ids):
        g = _i32(np.asarray(ids, np.int32))
        out = np.zeros(max(len(g), 1), np.int32)
        k = self._lib.gaia_b200_partition_restrict_ids(self._p, len(g), _ip(g), _ip(out))
        if k < 0:
            raise GaiaError(f"gaia_b200_partition_restrict_ids failed: {k}")
        return out[:k]

    def restrict_colors(self, colors):
        offs = np.zeros(len(colors) + 1, np.int32)
        offs[1:] = np.cumsum([len(c) for c in colors])
        flat = _i32(np.concatenate([np.asarray(c, np.int32) for c in colors]) if len(colors) else np.zeros(0, np.int32))
        oo = np.zeros(len(colors) + 1, np.int32); ov = np.zeros(max(len(self.local_verts), 1), np.int32)
        rc = self._lib.gaia_b200_partition_restrict_colors(self._p, len(colors), _ip(offs), _ip(flat), _ip(oo), _ip(ov))
        if rc < 0:
            raise GaiaError(f"gaia_b200_partition_restrict_colors failed: {rc}")
        return [ov[oo[c]:oo[c + 1]].copy() for c in range(len(colors))]

    def close(self):
        if self._p:
            self._lib.gaia_b200_partition_destroy(self._p)
            self._p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def tet_surface(n_verts, tets):
    """(surfaceVIds, surfaceFaces[nF,3]) of a tet mesh in the reference's orderings, no device needed"""
    lib = load_library()
    t = _i32(np.asarray(tets, np.int32).reshape(-1, 4))
    nv, nf = C.c_int32(), C.c_int32()
    rc = lib.gaia_b200_tet_surface(int(n_verts), len(t), _ip(t), C.byref(nv), None, C.byref(nf), None)
    if rc < 0:
        raise GaiaError(f"gaia_b200_tet_surface failed: {rc}")
    ids = np.zeros(nv.value, np.int32); faces = np.zeros((nf.value, 3), np.int32)
    lib.gaia_b200_tet_surface(int(n_verts), len(t), _ip(t), C.byref(nv), _ip(ids), C.byref(nf), _ip(faces))
    return ids, faces
