"""Host-side scene helpers (Python mirror of the reference's input layer, used by tests and bench).

  load_t                   MeshFrame2 `.t` reader  (3P/MeshFrame2/.../TetMesh/TMeshStatic.h:51-200)
  load_coloring            `*.vertexColoring.json` reader (M/TetMesh/TetMeshFEM.cpp:122-133)
  physics/collision/object params from the reference's JSON schema
                           (M/Parameters/PhysicsParameters.h:107-178, M/VBD/VBDPhysicsParameters.h:85-147,
                            M/CollisionDetector/CollisionDetertionParameters.h:82-109, M/Materials/Materials.h:55-75,
                            M/VBD/VBD_BaseMaterial.h:434-443, M/VBD/VBD_NeoHookean.cpp:1613-1624)
  kuhn_beam / greedy_coloring   synthetic meshes of BASELINE.json's configs 4-5
"""
import json
import os
import numpy as np

from . import capi


def load_t(path):
    """Vertices keep file order (remapped to 0..n-1 in read order); tets keep file order."""
    vid, verts, tets = {}, [], []
    with open(path) as f:
        for line in f:
            tok = line.split()
            if not tok:
                continue
            if tok[0] == "Vertex":
                vid[int(tok[1])] = len(verts)
                verts.append((np.float32(tok[2]), np.float32(tok[3]), np.float32(tok[4])))
            elif tok[0] == "Tet":
                tets.append([vid[int(t)] for t in tok[2:6]])
    return np.asarray(verts, np.float32).reshape(-1, 3), np.asarray(tets, np.int32).reshape(-1, 4)


def save_t(path, verts, tets):
    with open(path, "w") as f:
        for i, v in enumerate(verts):
            f.write(f"Vertex {i} {v[0]:.9g} {v[1]:.9g} {v[2]:.9g}\n")
        for i, t in enumerate(tets):
            f.write(f"Tet {i} {t[0]} {t[1]} {t[2]} {t[3]}\n")


def load_coloring(path):
    with open(path) as f:
        return [np.asarray(c, np.int32) for c in json.load(f)]


def resolve_path(p, repo_root):
    """${REPO_ROOT} substitution + '\\' normalisation (M/Parser/InputHandler.h:139-183)."""
    return p.replace("${REPO_ROOT}", repo_root).replace("\\", "/")


def _take(dst, src, names, conv=None):
    for n in names:
        if n in src:
            v = src[n]
            if isinstance(v, bool):
                v = int(v)
            setattr(dst, n, conv(v) if conv else v)


def physics_params_from_json(j):
    p = capi.default_physics_params()
    if "gravity" in j:
        p.gravity = capi.c_f3(*j["gravity"])
    if "worldBounds" in j:
        p.worldBoundsMin = capi.c_f3(*j["worldBounds"][0]); p.worldBoundsMax = capi.c_f3(*j["worldBounds"][1])
    _take(p, j, ["timeStep", "numSubsteps", "iterations", "usePlaneGround", "useBowlGround", "checkAndUpdateWorldBounds",
                 "boundaryFrictionDynamic", "dcdTetMeshSceneBVHRebuildSteps", "dcdSurfaceSceneBVHRebuildSteps",
                 "ccdBVHRebuildSteps", "stepSize", "stepSizeGD", "associateGravityWithInertia", "collisionStiffness",
                 "intermediateCollisionIterations", "boundaryCollisionStiffness", "boundaryFrictionEpsV",
                 "frictionStartIter", "useAccelerator", "acceleratorRho", "collisionOffHeight"])
    return p


def bowl_ground_from_json(j):
    """(centre, radius, cap) of PhysicsParams when useBowlGround is set (BasePhysicsParams::bowlCenter / bowlRadius / bowlCap, the
    reference's defaults where a key is missing), else None: pass it to Scene.set_bowl_ground after creating the scene"""
    if not j.get("useBowlGround"):
        return None
    return [float(x) for x in j.get("bowlCenter", [0.0, 0.0, 0.0])], float(j.get("bowlRadius", 1.0)), bool(j.get("bowlCap", False))


def collision_params_from_json(j):
    c = capi.default_collision_params()
    _take(c, j, [n for n, _ in capi.CollisionParams._fields_])
    return c


def object_params_from_json(j):
    o = capi.default_object_params()
    for n in ("scale", "rotation", "translation", "initialVelocity"):
        if n in j:
            setattr(o, n, capi.c_f3(*j[n]))
    _take(o, j, ["density", "hasNoGravZone", "noGravZoneThreshold", "maxVelocityMagnitude", "frameToAppear",
                 "frictionDynamic", "frictionEpsV", "exponentialVelDamping", "constantVelDamping", "initializationType",
                 "miu", "lmbd", "dampingHydrostatic", "dampingDeviatoric"])
    return o


def load_scene_json(models_json, params_json, repo_root):
    """-> (phys, col, [(verts, tets, objParams, colors, fixedPoints), ...]); meshes de-duplicated by path."""
    with open(models_json) as f:
        models = json.load(f)["Models"]
    with open(params_json) as f:
        params = json.load(f)
    phys = physics_params_from_json(params.get("PhysicsParams", {}))
    col = collision_params_from_json(params.get("CollisionParams", {}))
    cache, out = {}, []
    for m in models:
        path = resolve_path(m["path"], repo_root)
        if path not in cache:
            cache[path] = load_t(path)
        cpath = resolve_path(m["verticesColoringCategoriesPath"], repo_root)
        if cpath not in cache:
            cache[cpath] = load_coloring(cpath)
        v, t = cache[path]
        out.append((v, t, object_params_from_json(m), cache[cpath], np.asarray(m.get("fixedPoints", []), np.int32)))
    return phys, col, out


# ------------------------------------------------------------------------------------------ synthetic meshes
_KUHN = np.array([[0, 1, 3, 7], [0, 3, 2, 7], [0, 2, 6, 7], [0, 6, 4, 7], [0, 4, 5, 7], [0, 5, 1, 7]], np.int64)


def kuhn_beam(nx, ny, nz, h=1.0, origin=(0.0, 0.0, 0.0)):
    """Regular nx*ny*nz hex grid, 6 Kuhn tets per cell (all positively oriented), spacing h.
    Vertex id = (i*(ny+1) + j)*(nz+1) + k. Returns (verts f32 [nV,3], tets i32 [6*cells,4])."""
    gx, gy, gz = nx + 1, ny + 1, nz + 1
    i, j, k = np.meshgrid(np.arange(gx), np.arange(gy), np.arange(gz), indexing="ij")
    verts = np.stack([i, j, k], -1).reshape(-1, 3).astype(np.float32) * np.float32(h) + np.asarray(origin, np.float32)
    ci, cj, ck = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    base = ((ci * gy + cj) * gz + ck).reshape(-1).astype(np.int64)
    # corner c has offsets (c&1, (c>>1)&1, (c>>2)&1) in (i,j,k)
    corner = np.array([((c & 1) * gy + ((c >> 1) & 1)) * gz + ((c >> 2) & 1) for c in range(8)], np.int64)
    tets = (base[:, None, None] + corner[_KUHN][None, :, :]).reshape(-1, 4)
    # orientation: make det(Dm) > 0
    p = verts[tets].astype(np.float64)
    det = np.einsum("ij,ij->i", p[:, 1] - p[:, 0], np.cross(p[:, 2] - p[:, 0], p[:, 3] - p[:, 0]))
    flip = det < 0
    tets[flip] = tets[flip][:, [0, 2, 1, 3]]
    return verts, tets.astype(np.int32)


def kuhn_coloring(nx, ny, nz):
    """Proper 4-colouring of the Kuhn grid's vertex graph: (i+j+k) mod 4 (edges join v and v+d, d in {0,1}^3\\0)."""
    gx, gy, gz = nx + 1, ny + 1, nz + 1
    i, j, k = np.meshgrid(np.arange(gx), np.arange(gy), np.arange(gz), indexing="ij")
    c = ((i + j + k) % 4).reshape(-1)
    return [np.nonzero(c == q)[0].astype(np.int32) for q in range(4)]


def vertex_adjacency(n_verts, tets):
    """CSR of the tet-mesh vertex graph (M/GraphColoring/TetMeshVertexGraph.h:13)."""
    e = np.concatenate([tets[:, [a, b]] for a in range(4) for b in range(4) if a != b]).astype(np.int64)
    key = np.unique(e[:, 0] * n_verts + e[:, 1])
    src, dst = key // n_verts, key % n_verts
    offs = np.zeros(n_verts + 1, np.int64)
    np.add.at(offs, src + 1, 1)
    return np.cumsum(offs), dst


def greedy_coloring(n_verts, tets):
    """Ordered greedy colouring, vertices in index order, smallest free colour
    (restating OrderedGreedy::color, M/GraphColoring/gready.cpp:47)."""
    offs, dst = vertex_adjacency(n_verts, tets)
    color = -np.ones(n_verts, np.int32)
    for v in range(n_verts):
        used = set(color[dst[offs[v]:offs[v + 1]]].tolist())
        c = 0
        while c in used:
            c += 1
        color[v] = c
    return [np.nonzero(color == q)[0].astype(np.int32) for q in range(color.max() + 1)]


def check_coloring(n_verts, tets, colors):
    col = -np.ones(n_verts, np.int64)
    for c, vs in enumerate(colors):
        col[vs] = c
    if (col < 0).any():
        return False
    ct = col[tets]
    for a in range(4):
        for b in range(a + 1, 4):
            if (ct[:, a] == ct[:, b]).any():
                return False
    return True


def add_scene_meshes(scene, meshes):
    for v, t, obj, colors, fixed in meshes:
        scene.add_tet_mesh(v, t, obj, colors, fixed)
    scene.finalize()
    return scene


def write_ply_ascii(path, verts, faces):
    """ASCII PLY as M/IO/FileIO.cpp:8-50 (saveAsPly)."""
    with open(path, "w") as f:
        f.write("ply\nformat ascii 1.0\n")
        f.write(f"element vertex {len(verts)}\nproperty float x\nproperty float y\nproperty float z\n")
        f.write(f"element face {len(faces)}\nproperty list uchar int vertex_indices\nend_header\n")
        for v in verts:
            f.write("%f %f %f\n" % (v[0], v[1], v[2]))
        for t in faces:
            f.write("3 %d %d %d\n" % (t[0], t[1], t[2]))


REFERENCE_ROOT = "/root/reference"


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "Data"))


def kuhn_blob(n, radii=(1.0, 1.0, 1.0), h=1.0, power=2.0):
    """A blobby tet mesh: the cells of an n^3 Kuhn grid whose centre lies inside the super-ellipsoid
    sum(|x_i / r_i|^power) <= 1 (coordinates scaled to [-1,1]); vertices compacted, spacing h.
    Stands in for the reference's organic models (teapot, bunny, octopus, armadillo) in synthetic many-object scenes."""
    verts, tets = kuhn_beam(n, n, n, h=1.0)
    cells = tets.reshape(-1, 6, 4)
    cen = verts[cells[:, 0, :]].mean(1)                    # tet 0 of a cell spans corner 0 and corner 7: its centroid is near the cell centre
    u = (cen / n * 2.0 - 1.0) / np.asarray(radii, np.float32)
    keep = (np.abs(u) ** power).sum(1) <= 1.0
    t = cells[keep].reshape(-1, 4)
    used = np.unique(t)
    remap = -np.ones(len(verts), np.int64)
    remap[used] = np.arange(len(used))
    v = (verts[used] - n / 2.0) * np.float32(h)
    return v.astype(np.float32), remap[t].astype(np.int32)


def place(verts, obj):
    """World-space rest positions of an object: rotation (axis * angle) -> scale -> translation, as TetMeshFEM::initialize
    does (M/TetMesh/TetMeshFEM.cpp:369-384). float64 numpy; used for partitioning decisions only."""
    p = np.asarray(verts, np.float64)
    rot = np.array([obj.rotation[0], obj.rotation[1], obj.rotation[2]], np.float64)
    ang = np.linalg.norm(rot)
    if ang > 0:
        a = rot / ang
        K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
        R = np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * (K @ K)
        p = p @ R.T
    return p * np.array([obj.scale[0], obj.scale[1], obj.scale[2]]) + np.array([obj.translation[0], obj.translation[1], obj.translation[2]])
