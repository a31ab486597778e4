"""Host side of the cloth path for scenes in the reference's own file formats: the cloth `Models.json` / `Parameters.json`
(S/VBDCloth/ParameterGen, e.g. S03_MultiLayerClothOnCollider.py), `.obj` meshes and `*.vertexColoring*.json` colourings.

    python -m gaia_b200.cloth_host Models.json Parameters.json outFolder [-R repoRoot] [--frames N]

Mirrors what BaseClothPhsicsFramework does around the step (M/Framework/BaseClothSimPhsicsFramework.cpp: loadRunningparameters,
initialize, simulate, writeOutputs): parameters keep the reference's names, meshes are loaded in file order (simulated
models, a collider being a model whose vertices are all fixed, as S03 does it), frames are written as `A%08d.ply`
(all meshes together, ASCII) or `A%08d.bin` (raw float32 positions). `Deformers` (Rotator / Translator,
M/VBDCloth/VBDClothDeformer.h:202-242) and `ColliderParams.ColliderMeshes` of type "TriMeshSequence" (key-frame .obj files
of one topology, M/SpatialQuery/ColiiderTriMeshSequence.h) are handed to the library; other collider types are refused.
"""
import argparse
import json
import os

import numpy as np

from . import cloth, meshio


def load_obj(path):
    """Triangle .obj: `v x y z` and `f a b c` (with or without /vt/vn); returns (verts float32 [n,3], faces int32 [m,3])."""
    verts, faces = [], []
    with open(path) as f:
        for line in f:
            if line.startswith("v "):
                p = line.split()
                verts.append((float(p[1]), float(p[2]), float(p[3])))
            elif line.startswith("f "):
                ids = [int(tok.split("/")[0]) for tok in line.split()[1:]]
                if len(ids) != 3:
                    raise ValueError(f"{path}: only triangle faces are supported")
                faces.append([i - 1 if i > 0 else len(verts) + i for i in ids])
    return np.asarray(verts, np.float32), np.asarray(faces, np.int32)


def save_obj(path, verts, faces):
    with open(path, "w") as f:
        for p in verts:
            f.write("v %.6f %.6f %.6f\n" % tuple(p))
        for t in faces:
            f.write("f %d %d %d\n" % (t[0] + 1, t[1] + 1, t[2] + 1))


_PHYS = ("timeStep", "numSubsteps", "iterations", "boundaryFrictionDynamic", "stepSize", "associateGravityWithInertia",
         "applyAcceleration", "handleCollision", "contactStiffness", "contactRadius", "thickness", "conservativeStepRelaxation",
         "boundaryContactStiffness", "boundaryFrictionEpsValue", "contactEnergyTypes", "contactBVHReconstructionSteps",
         "degenerateTriangleThres")
_OBJ = ("density", "hasNoGravZone", "noGravZoneThreshold", "maxVelocityMagnitude", "frictionDynamic", "frictionEpsV",
        "bendingStiffness", "exponentialVelDamping", "constantVelDamping", "initializationType", "miu", "lambda", "dampingStVK",
        "dampingBending")


def cloth_physics_params_from_json(j):
    """PhysicsParams + ContactDetectorParams -> ClothPhysicsParams (absent keys keep the reference's defaults)."""
    pp = j.get("PhysicsParams", {})
    kw = {k: (int(pp[k]) if isinstance(pp[k], bool) else pp[k]) for k in _PHYS if k in pp}
    if "gravity" in pp:
        kw["gravity"] = pp["gravity"]
    if "worldBounds" in pp:
        kw["worldBoundsMin"], kw["worldBoundsMax"] = pp["worldBounds"][0], pp["worldBounds"][1]
    if "accelerationRho" in pp:            # the scripts write accelerationRho, the struct field is acceleratorRho
        kw["acceleratorRho"] = pp["accelerationRho"]
    if "acceleratorRho" in pp:
        kw["acceleratorRho"] = pp["acceleratorRho"]
    cd = j.get("ContactDetectorParams", {})
    if "maxQueryDis" in cd:
        kw["maxQueryDis"] = cd["maxQueryDis"]
    return cloth.default_cloth_physics_params(**kw)


def cloth_object_params_from_json(m):
    kw = {k: (int(m[k]) if isinstance(m[k], bool) else m[k]) for k in _OBJ if k in m}
    for k in ("scale", "rotation", "translation"):
        if k in m:
            kw[k] = m[k]
    return cloth.default_cloth_object_params(**kw)


def load_cloth_scene_json(models_json, params_json, repo_root, allow_deformers=True):
    """-> (ClothPhysicsParams, [(verts, faces, ClothObjectParams, colours, fixedPoints, initialVelocity)], raw parameter dict);
    the raw dict's "Deformers" list goes to build_scene(deformers=...)"""
    with open(params_json) as f:
        pj = json.load(f)
    with open(models_json) as f:
        mj = json.load(f)
    if pj.get("Deformers") and not allow_deformers:
        raise NotImplementedError("cloth Deformers refused by the caller")
    for c in pj.get("ColliderParams", {}).get("ColliderMeshes", []) or []:
        if c.get("colliderType") != "TriMeshSequence":
            raise NotImplementedError(f"Unrecognized collider type: {c.get('colliderType')}")
    phys = cloth_physics_params_from_json(pj)
    meshes = []
    for m in mj["Models"]:
        v, f = load_obj(meshio.resolve_path(m["path"], repo_root))
        colors = meshio.load_coloring(meshio.resolve_path(m["verticesColoringCategoriesPath"], repo_root))
        meshes.append((v, f, cloth_object_params_from_json(m), colors, np.asarray(m.get("fixedPoints", []), np.int32),
                       np.asarray(m.get("initialVelocity", [0, 0, 0]), np.float32)))
    return phys, meshes, pj


def add_deformers(scene, deformers):
    """loadClothDeformers: one library deformer per (deformer, selected mesh)"""
    for d in deformers or []:
        name = d.get("DeformerName")
        if len(d["selectedMeshes"]) != len(d["selectedVertices"]):
            raise ValueError("deformer: selectedMeshes / selectedVertices size mismatch")
        for mesh, verts in zip(d["selectedMeshes"], d["selectedVertices"]):
            if name == "Rotator":
                scene.add_rotator(mesh, verts, d["center"], d["axis"], d["angularVelocity"], d.get("rotationEndTime", -1))
            elif name == "Translator":
                scene.add_translator(mesh, verts, d["speed"], d.get("deformationStartTime", 0.0), d.get("deformationEndTime", -1))
            else:
                raise ValueError(f"Unrecognized deformer: {name}")


def load_collider_sequences(pj, repo_root):
    """ColliderParams.ColliderMeshes -> [(key positions [nKeys, nV, 3], faces, ClothObjectParams, keyFrames | None,
    interpolateSubstep, interpolateIter)] (ColliderTrimeshSequenceParams::fromJson)"""
    out = []
    for c in pj.get("ColliderParams", {}).get("ColliderMeshes", []) or []:
        files = c.get("meshFiles", [])
        kf = c.get("keyFrames") or None
        if kf is not None and len(kf) > len(files):
            kf = kf[:len(files)]
        n = len(kf) if kf is not None else len(files)
        if n == 0:
            raise ValueError("collider sequence without mesh files")
        keys, faces = [], None
        for path in files[:n]:
            v, f = load_obj(meshio.resolve_path(path, repo_root))
            if faces is not None and (f.shape != faces.shape or len(v) != len(keys[0])):
                raise ValueError("collider key frames must share one topology")
            faces = f if faces is None else faces
            keys.append(v)
        out.append((np.stack(keys), faces, cloth_object_params_from_json(c), kf, bool(c.get("interpolateSubstep", True)),
                    int(c.get("interpolateIter", 1))))
    return out


def load_cloth_constraints(models_json, params_json, repo_root):
    """-> (sewing, attachments): sewing = [(mesh, seams [n, 3], ratios [n], sewingStiffness, filtered vf pairs, filtered ee pairs)] from the models' sewingPath / filteredPairsPath files
    (TriMeshFEM::initialize, M/TriMesh/TriMesh.cpp:95-110), attachments = [(meshV, vertex, meshF, faceVerts, barys, stiffness)]
    from PhysicsParams.VertexFaceAttachmentConstraintParams (M/VBDCloth/VBDTriMeshConstraints.h:6-29)"""
    with open(params_json) as f:
        pj = json.load(f)
    with open(models_json) as f:
        mj = json.load(f)
    sewing = []
    for i, m in enumerate(mj["Models"]):
        if m.get("sewingPath"):
            path = meshio.resolve_path(m["sewingPath"], repo_root)
            if not path.endswith(".json"):
                raise NotImplementedError(f"Unsupported file format: {path}")
            with open(path) as f:
                sj = json.load(f)
            vf, ee = np.zeros((0, 2), np.int32), np.zeros((0, 2), np.int32)
            if m.get("filteredPairsPath"):        # generate_sewing_filters reads it only beside a sewing file (M/TriMesh/TriMesh.h:603-621)
                with open(meshio.resolve_path(m["filteredPairsPath"], repo_root)) as f:
                    fj = json.load(f)
                vf = np.asarray(fj.get("vf", []), np.int32).reshape(-1, 2); ee = np.asarray(fj.get("ee", []), np.int32).reshape(-1, 2)
            sewing.append((i, np.asarray(sj["vertices"], np.int32).reshape(-1, 3), np.asarray(sj["ratios"], np.float32), float(m.get("sewingStiffness", 0.0)), vf, ee))
    a = pj.get("PhysicsParams", {}).get("VertexFaceAttachmentConstraintParams")
    attachments = []
    if a:
        vi = np.asarray(a.get("vertexConstraintVertexInfos", []), np.int32).reshape(-1, 2)
        fi = np.asarray(a.get("vertexConstraintTargetFacesInfos", []), np.int32).reshape(-1, 4)
        ba = np.asarray(a.get("vertexConstraintBarys", []), np.float32).reshape(-1, 3)
        ks = a.get("vertexConstraintStiffness") or []
        if not (len(vi) == len(fi) == len(ba)) or (len(ks) and len(ks) != len(vi)):
            raise ValueError("VertexFaceAttachmentConstraintParams: one (mesh, vertex), one (mesh, v, v, v), one barycentric triple per constraint")
        for k in range(len(vi)):
            attachments.append((int(vi[k, 0]), int(vi[k, 1]), int(fi[k, 0]), fi[k, 1:].copy(), ba[k].copy(), float(ks[k]) if len(ks) else 1e4))
    return sewing, attachments


def build_scene(cls, phys, meshes, deformers=None, colliders=None, constraints=None, **kw):
    s = cls(phys, **kw)
    for v, f, obj, colors, fixed, v0 in meshes:
        s.add_tri_mesh(v, f, obj, colors, fixed=fixed, simulated=True)
    for keys, faces, obj, kf, sub, it in colliders or []:
        cid = s.add_tri_mesh(keys[0], faces, obj, simulated=False)
        s.set_collider_sequence(cid, keys, kf, sub, it)
    add_deformers(s, deformers)
    sewing, attachments = constraints if constraints is not None else ([], [])
    for mesh, seams, ratios, k, vf, ee in sewing:
        s.set_sewing(mesh, seams, ratios, k)
        if len(vf) or len(ee):
            s.set_filtered_pairs(mesh, vf, ee)
    for mv, v, mf, fv, b, k in attachments:
        s.add_attachment(mv, v, mf, fv, b, k)
    s.finalize()
    for i, (v, f, obj, colors, fixed, v0) in enumerate(meshes):
        if np.any(v0 != 0):
            s.set_state(i, None, np.tile(v0, (len(v), 1)))
    return s


def write_frame(scene, meshes, out_dir, frame, ext="ply"):
    pos = [scene.positions(i) for i in range(len(meshes))]
    name = os.path.join(out_dir, "A%08d.%s" % (frame, ext))
    if ext == "bin":
        np.concatenate(pos).astype(np.float32).tofile(name)
    else:
        base, faces = 0, []
        for p, m in zip(pos, meshes):
            faces.append(m[1] + base)
            base += len(p)
        meshio.write_ply_ascii(name, np.concatenate(pos), np.concatenate(faces))
    return name


def simulate(models_json, params_json, out_dir, repo_root=".", frames=None, device=0):
    phys, meshes, pj = load_cloth_scene_json(models_json, params_json, repo_root)
    pp = pj.get("PhysicsParams", {})
    n = int(pp.get("numFrames", 1)) if frames is None else frames
    ext = pp.get("outputExt", "ply")
    os.makedirs(out_dir, exist_ok=True)
    s = build_scene(cloth.ClothScene, phys, meshes, deformers=pj.get("Deformers"), colliders=load_collider_sequences(pj, repo_root),
                    constraints=load_cloth_constraints(models_json, params_json, repo_root), device=device)
    write_frame(s, meshes, out_dir, 0, ext)
    for fr in range(1, n + 1):
        s.run_frames(1)
        write_frame(s, meshes, out_dir, fr, ext)
    st = s.stats()
    s.close()
    return st


def main():
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("models"); ap.add_argument("parameters"); ap.add_argument("out")
    ap.add_argument("-R", "--repoRoot", default=".")
    ap.add_argument("--frames", type=int, default=None)
    ap.add_argument("--device", type=int, default=0)
    a = ap.parse_args()
    st = simulate(a.models, a.parameters, a.out, a.repoRoot, a.frames, a.device)
    print(json.dumps(st))


if __name__ == "__main__":
    main()
