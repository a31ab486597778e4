"""ctypes binding of the cloth C ABI (include/gaia_b200_cloth.h, libgaia_b200.so) and helpers to build cloth scenes.

`ClothSceneBase` holds the calls shared by the product (`gaia_b200_cloth_*`) and the oracle's shadow ABI
(`oracle_cloth_*`, oracle/binding_cloth.py), so parity tests drive both with the same code.
"""
import ctypes as C

import numpy as np

from .capi import GaiaError, _f32, _fp, _i32, _ip, c_f3, load_library


class ClothPhysicsParams(C.Structure):
    """gaia_cloth_physics_params (BasePhysicsParams + VBDClothPhysicsParameters + ClothContactDetectorParameters)."""
    _fields_ = [
        ("gravity", c_f3), ("timeStep", C.c_float), ("numSubsteps", C.c_int32), ("iterations", C.c_int32),
        ("worldBoundsMin", c_f3), ("worldBoundsMax", c_f3), ("boundaryFrictionDynamic", C.c_float),
        ("stepSize", C.c_float), ("associateGravityWithInertia", C.c_int32), ("applyAcceleration", C.c_int32),
        ("acceleratorRho", C.c_float), ("handleCollision", C.c_int32), ("contactStiffness", C.c_float),
        ("contactRadius", C.c_float), ("thickness", C.c_float), ("conservativeStepRelaxation", C.c_float),
        ("boundaryContactStiffness", C.c_float), ("boundaryFrictionEpsValue", C.c_float), ("contactEnergyTypes", C.c_int32),
        ("contactBVHReconstructionSteps", C.c_int32), ("degenerateTriangleThres", C.c_float), ("maxQueryDis", C.c_float),
    ]


class ClothObjectParams(C.Structure):
    """gaia_cloth_object_params (ObjectParams + TriMeshParams + VBDObjectParamsTriMesh[StVk])."""
    _fields_ = [
        ("scale", c_f3), ("rotation", c_f3), ("translation", c_f3), ("density", C.c_float), ("hasNoGravZone", C.c_int32),
        ("noGravZoneThreshold", C.c_float), ("maxVelocityMagnitude", C.c_float), ("frictionDynamic", C.c_float),
        ("frictionEpsV", C.c_float), ("bendingStiffness", C.c_float), ("exponentialVelDamping", C.c_float),
        ("constantVelDamping", C.c_float), ("initializationType", C.c_int32), ("miu", C.c_float), ("lambda_", C.c_float),
        ("dampingStVK", C.c_float), ("dampingBending", C.c_float),
    ]


class ClothVFContact(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("vertexMesh", "vertex", "faceMesh", "face", "primitiveType", "primitiveId")]


class ClothEEContact(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("mesh1", "edge1", "mesh2", "edge2")]


class ClothStats(C.Structure):
    _fields_ = [("timeCsmpUpdatingCollisionInfoDCD", C.c_double), ("timeCsmpInitialStep", C.c_double),
                ("timeCsmpMaterialSolve", C.c_double), ("solveKernelMs", C.c_double)] + \
               [(n, C.c_int64) for n in ("vertexSolves", "colourSweeps", "substeps", "frames", "kernelLaunches", "algorithmicBytes",
                                         "numberCollisionDetection", "numContactsVF", "numContactsEE")]


def _set(struct, kw):
    for k, v in kw.items():
        k = "lambda_" if k == "lambda" else k
        cur = getattr(struct, k)
        if isinstance(cur, C.Array):
            setattr(struct, k, type(cur)(*[float(x) for x in v]))
        else:
            setattr(struct, k, v)
    return struct


def default_cloth_physics_params(**kw):
    p = ClothPhysicsParams()
    p.gravity = c_f3(0, -10, 0); p.timeStep = 0.01666666666666666666; p.numSubsteps = 20; p.iterations = 1
    p.worldBoundsMin = c_f3(-1e7, -1e7, -1e7); p.worldBoundsMax = c_f3(1e7, 1e7, 1e7); p.boundaryFrictionDynamic = 0.1
    p.stepSize = 1.0; p.associateGravityWithInertia = 1; p.applyAcceleration = 0; p.acceleratorRho = 0.95
    p.handleCollision = 1; p.contactStiffness = 1e4; p.contactRadius = 1.0; p.thickness = 0.0
    p.conservativeStepRelaxation = 0.45; p.boundaryContactStiffness = 1e4; p.boundaryFrictionEpsValue = 1e-2
    p.contactEnergyTypes = 1; p.contactBVHReconstructionSteps = 32; p.degenerateTriangleThres = 1e-6; p.maxQueryDis = 0.4
    return _set(p, kw)


def default_cloth_object_params(**kw):
    o = ClothObjectParams()
    o.scale = c_f3(1, 1, 1); o.rotation = c_f3(0, 0, 0); o.translation = c_f3(0, 0, 0); o.density = 1.0
    o.hasNoGravZone = 0; o.noGravZoneThreshold = 0.0; o.maxVelocityMagnitude = -1.0; o.frictionDynamic = 0.1
    o.frictionEpsV = 0.1; o.bendingStiffness = 0.0; o.exponentialVelDamping = 1.0; o.constantVelDamping = 0.0
    o.initializationType = 0; o.miu = 1e4; o.lambda_ = 1e4; o.dampingStVK = 0.0; o.dampingBending = 0.0
    return _set(o, kw)


def bind_cloth_api(lib, prefix):
    P, FP, IP = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int32)
    sig = {
        "add_tri_mesh": ([P, C.c_int32, C.c_int32, C.c_int32, FP, IP, C.POINTER(ClothObjectParams), C.c_int32, IP, IP, C.c_int32, IP], C.c_int),
        "add_rotator": ([P, C.c_int, IP, C.c_int32, FP, FP, C.c_double, C.c_double], C.c_int),
        "add_translator": ([P, C.c_int, IP, C.c_int32, FP, C.c_float, C.c_float], C.c_int),
        "set_collider_sequence": ([P, C.c_int, C.c_int32, IP, FP, C.c_int32, C.c_int32], C.c_int),
        "set_sewing": ([P, C.c_int, C.c_int32, IP, FP, C.c_float], C.c_int),
        "add_attachment": ([P, C.c_int, C.c_int32, C.c_int, IP, FP, C.c_float], C.c_int),
        "set_filtered_pairs": ([P, C.c_int, C.c_int32, IP, C.c_int32, IP], C.c_int),
        "finalize": ([P], C.c_int), "run_frames": ([P, C.c_int32], C.c_int), "run_iterations": ([P, C.c_int32], C.c_int),
        "frame_id": ([P], C.c_int), "cur_time": ([P], C.c_double), "set_time": ([P, C.c_int32, C.c_double], C.c_int),
        "synchronize": ([P], C.c_int), "num_meshes": ([P], C.c_int), "mesh_sizes": ([P, C.c_int, IP, IP, IP], C.c_int),
        "get_positions": ([P, C.c_int, FP], C.c_int), "get_velocities": ([P, C.c_int, FP], C.c_int),
        "set_state": ([P, C.c_int, FP, FP], C.c_int), "total_vertices": ([P], C.c_int64),
        "get_all_positions": ([P, FP], C.c_int), "set_all_state": ([P, FP, FP], C.c_int),
        "get_edges": ([P, C.c_int, IP], C.c_int), "num_parallel_groups": ([P], C.c_int),
        "get_parallel_group": ([P, C.c_int, IP, IP], C.c_int), "get_vertex_masses": ([P, C.c_int, FP], C.c_int),
        "detect_contacts": ([P], C.c_int),
        "get_vf_contacts": ([P, C.POINTER(ClothVFContact), C.c_int64, C.POINTER(C.c_int64)], C.c_int),
        "get_ee_contacts": ([P, C.POINTER(ClothEEContact), C.c_int64, C.POINTER(C.c_int64)], C.c_int),
        "get_conservative_bounds": ([P, C.c_int, FP], C.c_int),
        "get_stats": ([P, C.POINTER(ClothStats)], C.c_int), "reset_stats": ([P], C.c_int), "set_profiling": ([P, C.c_int], C.c_int),
    }
    for name, (args, res) in sig.items():
        fn = getattr(lib, prefix + name)
        fn.argtypes = args
        fn.restype = res


class ClothSceneBase:
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

    def add_tri_mesh(self, xyz, faces, obj, colors=(), fixed=(), simulated=True):
        xyz = _f32(xyz).reshape(-1, 3); faces = _i32(faces).reshape(-1, 3)
        offs = np.zeros(len(colors) + 1, np.int32)
        offs[1:] = np.cumsum([len(c) for c in colors])
        cv = _i32(np.concatenate([np.asarray(c, np.int32) for c in colors]) if len(colors) else np.zeros(0, np.int32))
        fx = _i32(np.asarray(fixed, np.int32))
        return self._call("add_tri_mesh", 1 if simulated else 0, xyz.shape[0], faces.shape[0], _fp(xyz), _ip(faces), C.byref(obj),
                          len(colors), _ip(offs), _ip(cv), len(fx), _ip(fx))

    def add_rotator(self, mesh, verts, center, axis, angular_velocity, rotation_end_time=-1.0):
        """ClothDeformerRotator on `verts` of `mesh` (they become fixed points); before finalize."""
        v = _i32(np.asarray(verts, np.int32)); c = _f32(np.asarray(center, np.float32)); a = _f32(np.asarray(axis, np.float32))
        return self._call("add_rotator", mesh, _ip(v), len(v), _fp(c), _fp(a), float(angular_velocity), float(rotation_end_time))

    def add_translator(self, mesh, verts, speed, start_time=0.0, end_time=-1.0):
        """ClothDeformerTranslater on `verts` of `mesh` (they become fixed points); before finalize."""
        v = _i32(np.asarray(verts, np.int32)); sp = _f32(np.asarray(speed, np.float32))
        return self._call("add_translator", mesh, _ip(v), len(v), _fp(sp), float(start_time), float(end_time))

    def set_collider_sequence(self, mesh, key_positions, key_frames=None, interpolate_substep=True, interpolate_iter=1):
        """ColliderTrimeshSequence: key-frame meshes [nKeys, nV, 3] of collider `mesh` (added with simulated=False); before finalize."""
        kp = _f32(np.asarray(key_positions, np.float32))
        kf = _i32(np.asarray(key_frames, np.int32)) if key_frames is not None else None
        return self._call("set_collider_sequence", mesh, kp.shape[0], _ip(kf) if kf is not None else None, _fp(kp),
                          1 if interpolate_substep else 0, int(interpolate_iter))

    def set_sewing(self, mesh, seams, ratios, stiffness):
        """Seams [[v0, v1, v2], ...] of simulated `mesh` with their ratios (the reference's sewingPath JSON) and
        ObjectParams::sewingStiffness; before finalize."""
        v = _i32(np.asarray(seams, np.int32).reshape(-1, 3)); r = _f32(np.asarray(ratios, np.float32).reshape(-1))
        if len(v) != len(r):
            raise ValueError("one ratio per seam")
        return self._call("set_sewing", mesh, len(r), _ip(v), _fp(r), float(stiffness))

    def set_filtered_pairs(self, mesh, vf_pairs=(), ee_pairs=()):
        """filteredPairsPath of the reference: (vertex, face) and (edge, edge) pairs of `mesh` that never form a contact; they take
        effect only if the mesh also has seams; before finalize."""
        vf = _i32(np.asarray(vf_pairs, np.int32).reshape(-1, 2)); ee = _i32(np.asarray(ee_pairs, np.int32).reshape(-1, 2))
        return self._call("set_filtered_pairs", mesh, len(vf), _ip(vf), len(ee), _ip(ee))

    def add_attachment(self, mesh_v, vertex, mesh_f, face_verts, bary_or_pos, stiffness=1e4):
        """One VertexFaceAttachmentConstraint: vertex of simulated mesh_v tied to bary . face_verts of mesh_f, or (mesh_f = -1)
        to the fixed position bary_or_pos; before finalize."""
        fv = _i32(np.asarray(face_verts if mesh_f >= 0 else [-1, -1, -1], np.int32)); b = _f32(np.asarray(bary_or_pos, np.float32))
        return self._call("add_attachment", int(mesh_v), int(vertex), int(mesh_f), _ip(fv), _fp(b), float(stiffness))

    def finalize(self):
        self._call("finalize")

    def run_frames(self, n=1):
        self._call("run_frames", n)

    def run_iterations(self, n=1):
        self._call("run_iterations", n)

    def synchronize(self):
        self._call("synchronize")

    @property
    def frame_id(self):
        return int(getattr(self._lib, self._prefix + "frame_id")(self.h))

    @property
    def cur_time(self):
        return float(getattr(self._lib, self._prefix + "cur_time")(self.h))

    def set_time(self, frame_id, cur_time):
        self._call("set_time", int(frame_id), float(cur_time))

    @property
    def num_meshes(self):
        return self._call("num_meshes")

    def mesh_sizes(self, mesh):
        v = [C.c_int32() for _ in range(3)]
        self._call("mesh_sizes", mesh, *[C.byref(x) for x in v])
        return tuple(x.value for x in v)

    def positions(self, mesh):
        out = np.empty((self.mesh_sizes(mesh)[0], 3), np.float32)
        self._call("get_positions", mesh, _fp(out))
        return out

    def velocities(self, mesh):
        out = np.empty((self.mesh_sizes(mesh)[0], 3), np.float32)
        self._call("get_velocities", mesh, _fp(out))
        return out

    def set_state(self, mesh, pos=None, vel=None):
        p = _f32(pos) if pos is not None else None
        v = _f32(vel) if vel is not None else None
        self._call("set_state", mesh, _fp(p) if p is not None else None, _fp(v) if v is not None else None)

    @property
    def total_vertices(self):
        return int(getattr(self._lib, self._prefix + "total_vertices")(self.h))

    def all_positions(self, out=None):
        if out is None:
            out = np.empty((self.total_vertices, 3), np.float32)
        self._call("get_all_positions", _fp(out))
        return out

    def set_all_state(self, pos=None, vel=None):
        p = _f32(pos) if pos is not None else None
        v = _f32(vel) if vel is not None else None
        self._call("set_all_state", _fp(p) if p is not None else None, _fp(v) if v is not None else None)

    def edges(self, mesh):
        out = np.empty((self.mesh_sizes(mesh)[2], 6), np.int32)
        self._call("get_edges", mesh, _ip(out))
        return out

    def parallel_groups(self):
        groups = []
        for g in range(self._call("num_parallel_groups")):
            n = C.c_int32()
            self._call("get_parallel_group", g, None, C.byref(n))
            out = np.empty((n.value, 2), np.int32)
            self._call("get_parallel_group", g, _ip(out), C.byref(n))
            groups.append(out)
        return groups

    def vertex_masses(self, mesh):
        out = np.empty(self.mesh_sizes(mesh)[0], np.float32)
        self._call("get_vertex_masses", mesh, _fp(out))
        return out

    def detect_contacts(self):
        self._call("detect_contacts")

    def vf_contacts(self):
        n = C.c_int64()
        self._call("get_vf_contacts", None, 0, C.byref(n))
        buf = (ClothVFContact * max(1, n.value))()
        self._call("get_vf_contacts", buf, n.value, C.byref(n))
        return np.array([[c.vertexMesh, c.vertex, c.faceMesh, c.face, c.primitiveType, c.primitiveId] for c in buf[:n.value]], np.int32).reshape(-1, 6)

    def ee_contacts(self):
        n = C.c_int64()
        self._call("get_ee_contacts", None, 0, C.byref(n))
        buf = (ClothEEContact * max(1, n.value))()
        self._call("get_ee_contacts", buf, n.value, C.byref(n))
        return np.array([[c.mesh1, c.edge1, c.mesh2, c.edge2] for c in buf[:n.value]], np.int32).reshape(-1, 4)

    def conservative_bounds(self, mesh):
        out = np.empty(self.mesh_sizes(mesh)[0], np.float32)
        self._call("get_conservative_bounds", mesh, _fp(out))
        return out

    def stats(self):
        st = ClothStats()
        self._call("get_stats", C.byref(st))
        return {n: getattr(st, n) for n, _ in ClothStats._fields_}

    def reset_stats(self):
        self._call("reset_stats")

    def set_profiling(self, on):
        self._call("set_profiling", 1 if on else 0)


_bound = False


class ClothScene(ClothSceneBase):
    """The product: gaia_b200_cloth_* of libgaia_b200.so (CUDA, no CPU fallback)."""
    _prefix = "gaia_b200_cloth_"

    def __init__(self, phys=None, device=0):
        global _bound
        lib = load_library()
        if not _bound:
            lib.gaia_b200_cloth_create.argtypes = [C.POINTER(ClothPhysicsParams), C.c_int, C.POINTER(C.c_void_p)]
            lib.gaia_b200_cloth_destroy.argtypes = [C.c_void_p]; lib.gaia_b200_cloth_destroy.restype = None
            lib.gaia_b200_cloth_last_error.argtypes = [C.c_void_p]; lib.gaia_b200_cloth_last_error.restype = C.c_char_p
            lib.gaia_b200_cloth_cuda_stream.argtypes = [C.c_void_p]; lib.gaia_b200_cloth_cuda_stream.restype = C.c_void_p
            bind_cloth_api(lib, "gaia_b200_cloth_")
            _bound = True
        self._lib = lib
        self.phys = phys if phys is not None else default_cloth_physics_params()
        h = C.c_void_p()
        rc = lib.gaia_b200_cloth_create(C.byref(self.phys), int(device), C.byref(h))
        if rc < 0:
            msg = lib.gaia_b200_cloth_last_error(None)
            raise GaiaError(f"gaia_b200_cloth_create failed: {rc} {msg.decode() if msg else ''}")
        self.h = h

    def last_error(self):
        msg = self._lib.gaia_b200_cloth_last_error(self.h)
        return msg.decode() if msg else ""

    @property
    def cuda_stream(self):
        return self._lib.gaia_b200_cloth_cuda_stream(self.h)

    def close(self):
        if self.h:
            self._lib.gaia_b200_cloth_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ------------------------------------------------------------------------------------------------ synthetic cloth
def grid_cloth(n, size=1.0, origin=(0.0, 0.0, 0.0), axes="xz"):
    """n x n vertex grid (the reference's C{N}_horizontal.obj family: SURVEY.md §8(d) cfg 3), two triangles per cell with
    alternating diagonals; returns (verts [n*n,3], faces [2(n-1)^2,3])."""
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    u = (i.ravel() / (n - 1) - 0.5) * size
    w = (j.ravel() / (n - 1) - 0.5) * size
    verts = np.zeros((n * n, 3), np.float32)
    a, b = "xyz".index(axes[0]), "xyz".index(axes[1])
    verts[:, a] = u; verts[:, b] = w
    verts += np.asarray(origin, np.float32)
    vid = lambda ii, jj: ii * n + jj  # noqa: E731
    faces = []
    for ii in range(n - 1):
        for jj in range(n - 1):
            v00, v10, v01, v11 = vid(ii, jj), vid(ii + 1, jj), vid(ii, jj + 1), vid(ii + 1, jj + 1)
            if (ii + jj) % 2 == 0:
                faces += [(v00, v11, v10), (v00, v01, v11)]
            else:
                faces += [(v00, v01, v10), (v10, v01, v11)]
    return verts, np.asarray(faces, np.int32)


def grid_cloth_fast(n, size=1.0, origin=(0.0, 0.0, 0.0)):
    """Vectorised grid_cloth (xz plane) for the million-vertex bench scene."""
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    verts = np.zeros((n * n, 3), np.float32)
    verts[:, 0] = (i.ravel() / (n - 1) - 0.5) * size
    verts[:, 2] = (j.ravel() / (n - 1) - 0.5) * size
    verts += np.asarray(origin, np.float32)
    ii, jj = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="ij")
    ii = ii.ravel(); jj = jj.ravel()
    v00 = ii * n + jj; v10 = v00 + n; v01 = v00 + 1; v11 = v10 + 1
    even = ((ii + jj) % 2 == 0)
    f0 = np.where(even[:, None], np.stack([v00, v11, v10], 1), np.stack([v00, v01, v10], 1))
    f1 = np.where(even[:, None], np.stack([v00, v01, v11], 1), np.stack([v10, v01, v11], 1))
    faces = np.empty((2 * len(ii), 3), np.int32)
    faces[0::2] = f0; faces[1::2] = f1
    return verts, faces


def cloth_vertex_graph(n_verts, faces):
    """Adjacency the cloth colouring must respect (M/GraphColoring/TriMeshVertexGraph.h:25-38 with -E): mesh edges plus the
    two vertices opposite each interior edge (they share a bending hinge). Returns a set-free CSR (offsets, neighbours)."""
    faces = np.asarray(faces, np.int64)
    e = np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]])
    opp = np.concatenate([faces[:, 2], faces[:, 0], faces[:, 1]])
    key = np.minimum(e[:, 0], e[:, 1]) * n_verts + np.maximum(e[:, 0], e[:, 1])
    order = np.argsort(key, kind="stable")
    ks, os_ = key[order], opp[order]
    same = np.nonzero(ks[1:] == ks[:-1])[0]
    hinge = np.stack([os_[same], os_[same + 1]], 1)
    pairs = np.concatenate([e, hinge])
    pairs = np.concatenate([pairs, pairs[:, ::-1]])
    pairs = np.unique(pairs, axis=0)
    offs = np.zeros(n_verts + 1, np.int64)
    np.add.at(offs, pairs[:, 0] + 1, 1)
    offs = np.cumsum(offs)
    return offs, pairs[:, 1].astype(np.int64)


def greedy_cloth_coloring(n_verts, faces):
    """Greedy colouring (ascending vertex id, smallest free colour) of cloth_vertex_graph: a restatement of the reference's
    offline colouring tool's greedy mode, good enough for synthetic meshes."""
    offs, nbr = cloth_vertex_graph(n_verts, faces)
    color = -np.ones(n_verts, np.int64)
    for v in range(n_verts):
        used = set(color[nbr[offs[v]:offs[v + 1]]].tolist())
        c = 0
        while c in used:
            c += 1
        color[v] = c
    return [np.nonzero(color == c)[0].astype(np.int32) for c in range(int(color.max()) + 1)]


def grid_cloth_coloring(n):
    """Closed-form colouring of grid_cloth: colour = (i mod 3) * 3 + (j mod 3) separates every pair of vertices within
    Chebyshev distance 2 on the grid, which covers mesh edges and hinge-opposite pairs (9 colours)."""
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    c = ((i % 3) * 3 + (j % 3)).ravel()
    return [np.nonzero(c == k)[0].astype(np.int32) for k in range(9)]


def check_cloth_coloring(n_verts, faces, colors):
    offs, nbr = cloth_vertex_graph(n_verts, faces)
    col = -np.ones(n_verts, np.int64)
    for k, c in enumerate(colors):
        col[np.asarray(c)] = k
    if (col < 0).any():
        return False
    src = np.repeat(np.arange(n_verts), np.diff(offs))
    return not (col[src] == col[nbr]).any()


def uv_sphere(radius=1.0, n_lat=16, n_lon=32, center=(0.0, 0.0, 0.0)):
    """Closed triangle sphere (stand-in for S/VBDCloth/ParameterGen/Data/SphereCollider.obj), outward orientation."""
    verts = [(0.0, radius, 0.0)]
    for a in range(1, n_lat):
        th = np.pi * a / n_lat
        for b in range(n_lon):
            ph = 2 * np.pi * b / n_lon
            verts.append((radius * np.sin(th) * np.cos(ph), radius * np.cos(th), radius * np.sin(th) * np.sin(ph)))
    verts.append((0.0, -radius, 0.0))
    verts = np.asarray(verts, np.float32) + np.asarray(center, np.float32)
    faces = []
    ring = lambda a, b: 1 + (a - 1) * n_lon + (b % n_lon)  # noqa: E731
    for b in range(n_lon):
        faces.append((0, ring(1, b + 1), ring(1, b)))
    for a in range(1, n_lat - 1):
        for b in range(n_lon):
            faces.append((ring(a, b), ring(a, b + 1), ring(a + 1, b + 1)))
            faces.append((ring(a, b), ring(a + 1, b + 1), ring(a + 1, b)))
    last = len(verts) - 1
    for b in range(n_lon):
        faces.append((last, ring(n_lat - 1, b), ring(n_lat - 1, b + 1)))
    return verts, np.asarray(faces, np.int32)


def icosphere(radius=1.0, level=3, center=(0.0, 0.0, 0.0)):
    """Subdivided icosahedron (level 3 = 642 vertices / 1280 triangles, the shape of S/VBDCloth/ParameterGen/Data/
    SphereCollider.obj): every vertex has valence 5 or 6, unlike uv_sphere whose poles collect n_lon edges."""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    verts = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
             (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    verts = [tuple(np.asarray(v, np.float64) / np.linalg.norm(v)) for v in verts]
    faces = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6), (7, 1, 8),
             (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    for _ in range(level):
        cache, new_faces = {}, []

        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = (np.asarray(verts[a]) + np.asarray(verts[b])) * 0.5
                verts.append(tuple(m / np.linalg.norm(m)))
                cache[key] = len(verts) - 1
            return cache[key]
        for a, b, c in faces:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            new_faces += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        faces = new_faces
    v = (np.asarray(verts, np.float64) * radius + np.asarray(center, np.float64)).astype(np.float32)
    return v, np.asarray(faces, np.int32)
