"""The host side of the cloth path reads the reference's cloth scene files (Models.json / Parameters.json, .obj,
vertexColoring json) — CPU: schema + file formats, with the oracle as the solver; GPU: the same scene through the product."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import has_gpu
from gaia_b200 import cloth, cloth_host, meshio
from oracle.binding_cloth import OracleClothScene


def write_scene(tmp, n=10, frames=3):
    v, f = cloth.grid_cloth(n, 1.0)
    sv, sf = cloth.icosphere(0.25, 1)
    cloth_host.save_obj(os.path.join(tmp, "sheet.obj"), v, f)
    cloth_host.save_obj(os.path.join(tmp, "ball.obj"), sv, sf)
    with open(os.path.join(tmp, "sheet.obj.vertexColoring.json"), "w") as fh:
        json.dump([c.tolist() for c in cloth.grid_cloth_coloring(n)], fh)
    with open(os.path.join(tmp, "ball.obj.vertexColoring.json"), "w") as fh:
        json.dump([c.tolist() for c in cloth.greedy_cloth_coloring(len(sv), sf)], fh)
    base = {"density": 0.2, "miu": 2000.0, "lambda": 2000.0, "bendingStiffness": 0.5, "dampingStVK": 1e-6, "dampingBending": 1e-6,
            "initializationType": 2, "scale": [1, 1, 1], "rotation": [0, 0, 0], "hasNoGravZone": False, "materialName": "StVK_triMesh"}
    models = [dict(base, path="${REPO_ROOT}\\sheet.obj", verticesColoringCategoriesPath="${REPO_ROOT}\\sheet.obj.vertexColoring.json",
                   translation=[0.0, 0.88, 0.0], fixedPoints=[], initialVelocity=[0.0, -0.5, 0.0]),
              dict(base, path="${REPO_ROOT}\\ball.obj", verticesColoringCategoriesPath="${REPO_ROOT}\\ball.obj.vertexColoring.json",
                   translation=[0.0, 0.6, 0.0], fixedPoints=list(range(len(sv))))]
    params = {"PhysicsParams": {"timeStep": 0.01666666, "numSubsteps": 4, "iterations": 6, "gravity": [0.0, -10.0, 0.0], "numFrames": frames,
                                "contactStiffness": 1e5, "contactRadius": 0.02, "thickness": 1e-4, "conservativeStepRelaxation": 0.4,
                                "handleCollision": True, "applyAcceleration": False, "accelerationRho": 0.9, "outputExt": "ply",
                                "worldBounds": [[-10, -10, -10], [10, 10, 10]]},
              "ContactDetectorParams": {"maxQueryDis": 0.03}, "ViewerParams": {"enableViewer": False}, "Deformers": []}
    with open(os.path.join(tmp, "Models.json"), "w") as fh:
        json.dump({"Models": models}, fh)
    with open(os.path.join(tmp, "Parameters.json"), "w") as fh:
        json.dump(params, fh)
    return os.path.join(tmp, "Models.json"), os.path.join(tmp, "Parameters.json"), (v, f, sv, sf)


def test_scene_files_parse_and_drive_the_oracle(tmp_path):
    m, p, (v, f, sv, sf) = write_scene(str(tmp_path))
    phys, meshes, pj = cloth_host.load_cloth_scene_json(m, p, str(tmp_path))
    assert phys.numSubsteps == 4 and phys.iterations == 6 and phys.handleCollision == 1 and abs(phys.maxQueryDis - 0.03) < 1e-9
    assert abs(phys.acceleratorRho - 0.9) < 1e-6 and abs(phys.thickness - 1e-4) < 1e-10
    assert len(meshes) == 2 and np.allclose(meshes[0][0], v, atol=1e-6) and np.array_equal(meshes[0][1], f)
    assert meshes[0][2].initializationType == 2 and abs(meshes[0][2].lambda_ - 2000.0) < 1e-3
    assert len(meshes[1][4]) == len(sv) and np.allclose(meshes[0][5], [0, -0.5, 0])
    o = cloth_host.build_scene(OracleClothScene, phys, meshes, threads=2)
    x0 = o.positions(0)
    o.run_frames(3)
    x = o.positions(0)
    assert np.isfinite(x).all() and x[:, 1].mean() < x0[:, 1].mean() - 1e-3
    assert np.allclose(o.positions(1), sv + np.float32([0, 0.6, 0]), atol=2e-6)      # the collider stays (.obj keeps 6 decimals)
    out = cloth_host.write_frame(o, meshes, str(tmp_path), 3, "ply")
    head = open(out).read(400)
    assert "element vertex %d" % (len(v) + len(sv)) in head and "element face %d" % (len(f) + len(sf)) in head
    out = cloth_host.write_frame(o, meshes, str(tmp_path), 3, "bin")
    assert os.path.getsize(out) == 12 * (len(v) + len(sv))


def add_collider_sequence(tmp, p, sv, sf, collider_type="TriMeshSequence"):
    """ColliderParams as the cloth parameter scripts write them (S03_MultiLayerClothOnCollider.py:14-35, 143-145)"""
    files = []
    for k, dy in enumerate((0.0, 0.2, 0.25)):
        name = os.path.join(tmp, "mover%d.obj" % k)
        cloth_host.save_obj(name, sv + np.float32([0.05 * k, dy, 0]), sf)
        files.append("${REPO_ROOT}\\mover%d.obj" % k)
    pj = json.load(open(p))
    pj["ColliderParams"] = {"ColliderMeshes": [{"colliderType": collider_type, "path": files[0], "meshFiles": files, "keyFrames": [0, 2, 4],
                                                "interpolateSubstep": True, "interpolateIter": 2, "scale": [1, 1, 1], "rotation": [0, 0, 0],
                                                "translation": [0.0, 0.35, 0.0], "frictionDynamic": 0.3}]}
    json.dump(pj, open(p, "w"))


def test_collider_sequence_from_scene_files(tmp_path):
    m, p, (v, f, sv, sf) = write_scene(str(tmp_path))
    add_collider_sequence(str(tmp_path), p, sv, sf)
    phys, meshes, pj = cloth_host.load_cloth_scene_json(m, p, str(tmp_path))
    cols = cloth_host.load_collider_sequences(pj, str(tmp_path))
    assert len(cols) == 1 and cols[0][0].shape == (3, len(sv), 3) and cols[0][3] == [0, 2, 4] and cols[0][5] == 2
    o = cloth_host.build_scene(OracleClothScene, phys, meshes, colliders=cols, threads=2)
    assert o.num_meshes == 3
    y0 = o.positions(2)[:, 1].mean()
    o.run_frames(3)
    y3 = o.positions(2)[:, 1].mean()
    assert abs(y0 - (sv[:, 1].mean() + 0.35)) < 1e-5 and 0.2 < y3 - y0 < 0.25          # between key 1 (frame 2) and key 2 (frame 4)
    r = subprocess.run([CLOTH_HOST_BIN, m, p, str(tmp_path / "out"), "-R", str(tmp_path), "--dryRun"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    add_collider_sequence(str(tmp_path), p, sv, sf, collider_type="Implicit")
    with pytest.raises(NotImplementedError):
        cloth_host.load_cloth_scene_json(m, p, str(tmp_path))
    r = subprocess.run([CLOTH_HOST_BIN, m, p, str(tmp_path / "out"), "-R", str(tmp_path), "--dryRun"], capture_output=True, text=True)
    assert r.returncode == 3 and "Unrecognized collider type" in r.stderr


def add_constraints(tmp, m, p):
    """sewingPath / sewingStiffness of a model and PhysicsParams.VertexFaceAttachmentConstraintParams, in the reference's schema
    (M/TriMesh/TriMesh.cpp:95-110, 458; M/VBDCloth/VBDTriMeshConstraints.h:6-29); the seam keeps to the 10 x 10 sheet's colouring"""
    with open(os.path.join(tmp, "seams.json"), "w") as fh:
        json.dump({"vertices": [[44, 48, 59], [2, 22, 23]], "ratios": [0.4, 1.0]}, fh)
    mj = json.load(open(m))
    with open(os.path.join(tmp, "pairs.json"), "w") as fh:
        json.dump({"vf": [[44, 90], [45, 91]], "ee": [[10, 200]]}, fh)
    mj["Models"][0].update(sewingPath="${REPO_ROOT}\\seams.json", sewingStiffness=300.0, filteredPairsPath="${REPO_ROOT}\\pairs.json")
    json.dump(mj, open(m, "w"))
    pj = json.load(open(p))
    pj["PhysicsParams"]["VertexFaceAttachmentConstraintParams"] = {
        "vertexConstraintVertexInfos": [0, 0, 0, 9, 0, 55], "vertexConstraintTargetFacesInfos": [-1, 0, 0, 0, 1, 0, 1, 2, 0, 33, 34, 43],
        "vertexConstraintBarys": [[-0.5, 0.9, -0.5], [0.2, 0.3, 0.5], [0.5, 0.5, 0.0]], "vertexConstraintStiffness": [2e3, 1e3, 5e2]}
    json.dump(pj, open(p, "w"))


def test_sewing_and_attachments_from_scene_files(tmp_path):
    m, p, (v, f, sv, sf) = write_scene(str(tmp_path))
    add_constraints(str(tmp_path), m, p)
    phys, meshes, pj = cloth_host.load_cloth_scene_json(m, p, str(tmp_path))
    sewing, att = cloth_host.load_cloth_constraints(m, p, str(tmp_path))
    assert len(sewing) == 1 and sewing[0][0] == 0 and sewing[0][1].tolist() == [[44, 48, 59], [2, 22, 23]] and sewing[0][3] == 300.0
    assert sewing[0][4].tolist() == [[44, 90], [45, 91]] and sewing[0][5].tolist() == [[10, 200]]
    assert [a[:3] for a in att] == [(0, 0, -1), (0, 9, 1), (0, 55, 0)] and att[1][3].tolist() == [0, 1, 2] and att[2][5] == 500.0
    o = cloth_host.build_scene(OracleClothScene, phys, meshes, constraints=(sewing, att), threads=2)
    plain = cloth_host.build_scene(OracleClothScene, phys, meshes, threads=2)
    o.run_frames(2); plain.run_frames(2)
    x, xp = o.positions(0), plain.positions(0)
    assert np.isfinite(x).all()
    corner = np.float32([-0.5, 0.9, -0.5])
    assert np.linalg.norm(x[0] - corner) < 0.5 * np.linalg.norm(xp[0] - corner)       # tied to a fixed point above its rest place
    gap = lambda q: np.linalg.norm(q[44] - (0.6 * q[48] + 0.4 * q[59]))
    assert gap(x) < 0.8 * gap(xp)                                                      # the seam closes
    r = subprocess.run([CLOTH_HOST_BIN, m, p, str(tmp_path / "out"), "-R", str(tmp_path), "--dryRun"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    pj = json.load(open(p))
    pj["PhysicsParams"]["VertexFaceAttachmentConstraintParams"]["vertexConstraintBarys"].pop()
    json.dump(pj, open(p, "w"))
    r = subprocess.run([CLOTH_HOST_BIN, m, p, str(tmp_path / "out"), "-R", str(tmp_path), "--dryRun"], capture_output=True, text=True)
    assert r.returncode == 3 and "VertexFaceAttachmentConstraintParams" in r.stderr
    with pytest.raises(ValueError):
        cloth_host.load_cloth_constraints(m, p, str(tmp_path))


@pytest.mark.skipif(not meshio.reference_available(), reason="/root/reference only exists in the build container")
def test_reference_cloth_experiment_files_parse():
    d = ("/root/reference/Simulator/VBDCloth/ParameterGen/Parameters/S01_Twisting/"
         "C100_Twist_ContactR0.2_CStif100000.0_bd20_damping3e-05_relaxiation0.42_thick0.01_CDPerIter")
    phys, meshes, pj = cloth_host.load_cloth_scene_json(d + "/Models.json", d + "/Parameters.json", "/root/reference", allow_deformers=True)
    assert phys.numSubsteps == 5 and phys.iterations == 10 and abs(phys.contactRadius - 0.2) < 1e-7 and abs(phys.maxQueryDis - 0.3) < 1e-7
    v, f, obj, colors, fixed, v0 = meshes[0]
    assert v.shape == (10000, 3) and f.shape == (19602, 3) and abs(obj.bendingStiffness - 20) < 1e-6 and obj.initializationType == 2
    assert sum(len(c) for c in colors) == 10000 and cloth.check_cloth_coloring(10000, f, colors)
    assert len(pj["Deformers"]) == 2


@pytest.mark.gpu
def test_gpu_scene_with_sewing_and_attachments_from_files(tmp_path):
    """gaia_vbdcloth_b200 reads the seams and attachments from the scene files: its frames equal the Python binding's (same
    library underneath: bit-identical), and the binding's run matches the oracle"""
    m, p, _ = write_scene(str(tmp_path))
    add_constraints(str(tmp_path), m, p)
    pj = json.load(open(p))
    pj["PhysicsParams"].update(outputExt="bin")
    json.dump(pj, open(p, "w"))
    r = subprocess.run([CLOTH_HOST_BIN, m, p, str(tmp_path / "outc"), "-R", str(tmp_path), "--frames", "2"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    phys, meshes, _ = cloth_host.load_cloth_scene_json(m, p, str(tmp_path))
    con = cloth_host.load_cloth_constraints(m, p, str(tmp_path))
    g = cloth_host.build_scene(cloth.ClothScene, phys, meshes, constraints=con)
    o = cloth_host.build_scene(OracleClothScene, phys, meshes, constraints=con, threads=2)
    plain = cloth_host.build_scene(OracleClothScene, phys, meshes, threads=2)
    # (oracle against itself with every coordinate moved by 1 / 8 ulp: 1e-6 / 3e-5 after one and two frames, chaotic from the third)
    g.run_frames(1); o.run_frames(1); plain.run_frames(1)
    assert np.abs(g.positions(0) - o.positions(0)).max() < 5e-4 < 0.01 < np.abs(plain.positions(0) - o.positions(0)).max()
    g.run_frames(1); o.run_frames(1)
    assert np.abs(g.positions(0) - o.positions(0)).max() < 2e-3
    a = np.fromfile(tmp_path / "outc" / "A00000002.bin", np.float32).reshape(-1, 3)
    assert np.array_equal(a, np.concatenate([g.positions(0), g.positions(1)]))


@pytest.mark.gpu
def test_gpu_scene_from_files_matches_the_oracle(tmp_path):
    m, p, _ = write_scene(str(tmp_path))
    st = cloth_host.simulate(m, p, str(tmp_path / "out"), str(tmp_path), frames=3)
    assert st["frames"] == 3 and st["numberCollisionDetection"] >= 12
    assert sorted(os.listdir(tmp_path / "out")) == ["A%08d.ply" % i for i in range(4)]
    phys, meshes, _ = cloth_host.load_cloth_scene_json(m, p, str(tmp_path))
    g = cloth_host.build_scene(cloth.ClothScene, phys, meshes)
    o = cloth_host.build_scene(OracleClothScene, phys, meshes, threads=2)
    g.run_frames(3); o.run_frames(3)
    assert np.abs(g.positions(0) - o.positions(0)).max() < 1e-4


# ------------------------------------------------------------------------------------------------ C++ host (gaia_vbdcloth_b200)
CLOTH_HOST_BIN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gaia_b200", "host", "gaia_vbdcloth_b200")


def test_cpp_cloth_host_dry_run_reads_the_schema(tmp_path):
    m, p, (v, f, sv, sf) = write_scene(str(tmp_path))
    r = subprocess.run([CLOTH_HOST_BIN, m, p, str(tmp_path / "out"), "-R", str(tmp_path), "--dryRun"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert ("dryRun: 2 models, %d vertices, %d triangles, numFrames 3, numSubsteps 4, iterations 6, contactRadius 0.02, outputExt ply"
            % (len(v) + len(sv), len(f) + len(sf))) in r.stdout


def test_cpp_cloth_host_refuses_cpu_mode_and_missing_gpu(tmp_path):
    m, p, _ = write_scene(str(tmp_path))
    r = subprocess.run([CLOTH_HOST_BIN, m, p, str(tmp_path / "out"), "--CPU"], capture_output=True, text=True)
    assert r.returncode == 2 and "no CPU path" in r.stderr
    if not has_gpu():
        r = subprocess.run([CLOTH_HOST_BIN, m, p, str(tmp_path / "out"), "-R", str(tmp_path)], capture_output=True, text=True)
        assert r.returncode == 3 and "gaia_b200_cloth_create failed" in r.stderr        # no silent CPU fallback


@pytest.mark.skipif(not meshio.reference_available(), reason="/root/reference only exists in the build container")
def test_cpp_cloth_host_reads_a_reference_experiment(tmp_path):
    d = ("/root/reference/Simulator/VBDCloth/ParameterGen/Parameters/S01_Twisting/"
         "C100_Twist_ContactR0.2_CStif100000.0_bd20_damping3e-05_relaxiation0.42_thick0.01_CDPerIter")
    p = d + "/Parameters.json"                 # two Rotators twist the sheet
    r = subprocess.run([CLOTH_HOST_BIN, d + "/Models.json", p, "/tmp/unused", "-R", "/root/reference", "--dryRun"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "10000 vertices, 19602 triangles" in r.stdout and "numSubsteps 5, iterations 10, contactRadius 0.2" in r.stdout


@pytest.mark.gpu
def test_cpp_cloth_host_frames_match_the_python_binding(tmp_path):
    m, p, (v, f, sv, sf) = write_scene(str(tmp_path), frames=4)
    pj = json.load(open(p))
    pj["PhysicsParams"].update(outputExt="bin", binaryModeVisualizationSteps=2, outputRecoveryStateStep=2, outputStatistics=True)
    pj["Deformers"] = [{"DeformerName": "Translator", "speed": [0.0, 0.3, 0.0], "selectedMeshes": [0], "selectedVertices": [[0, 9]],
                        "deformationEndTime": -1}]
    json.dump(pj, open(p, "w"))
    add_collider_sequence(str(tmp_path), p, sv, sf)
    out = tmp_path / "out"
    r = subprocess.run([CLOTH_HOST_BIN, m, p, str(out), "-R", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    phys, meshes, pj = cloth_host.load_cloth_scene_json(m, p, str(tmp_path))
    g = cloth_host.build_scene(cloth.ClothScene, phys, meshes, deformers=pj["Deformers"],
                               colliders=cloth_host.load_collider_sequences(pj, str(tmp_path)))
    for fr in range(5):
        if fr:
            g.run_frames(1)
        a = np.fromfile(out / ("A%08d.bin" % fr), np.float32).reshape(-1, 3)
        assert np.array_equal(a, np.concatenate([g.positions(0), g.positions(1)])), fr       # same library underneath: bit-identical
    assert (out / "TemplateMesh" / "TemplateMesh_Model_000000_sheet.ply").exists()
    assert (out / "A00000002.ply").exists() and not (out / "A00000001.ply").exists()
    st = json.load(open(out / "RecoveryStates" / "A00000002.json"))
    assert st["frameId"] == 2 and st["format"] == "binary" and st["meshesState"]["nVerts"] == [len(v), len(sv)]
    stats = json.load(open(out / "Statistics" / "A00000003.json"))
    assert stats["substeps"] == 4 and stats["numberCollisionDetection"] >= 4
    head = open(out / "A00000002.ply").read(300)
    assert "element vertex %d" % (len(v) + len(sv)) in head and "element face %d" % (len(f) + len(sf)) in head
    # a run recovered from frame 2 carries on (positions and velocities are the whole state the reference saves too)
    part = tmp_path / "part"
    r = subprocess.run([CLOTH_HOST_BIN, m, p, str(part), "-R", str(tmp_path), "-r", str(out / "RecoveryStates" / "A00000002.json")],
                       capture_output=True, text=True)
    assert r.returncode == 0 and "Recovering state from" in r.stdout, r.stderr
    a = np.fromfile(out / "A00000004.bin", np.float32)
    b = np.fromfile(part / "A00000004.bin", np.float32)
    assert a.shape == b.shape and np.abs(a - b).max() < 1e-3
    g.close()


def test_oracle_deformers_follow_the_reference_schedule():
    """ClothDeformerRotator / ClothDeformerTranslater (M/VBDCloth/VBDClothDeformer.h:43-200) in the oracle: the selected
    vertices become fixed, the rotator places them at root + R(curTime * w) RootP, the translator adds speed * dt per substep
    while start <= curTime < end."""
    n = 10
    v, f = cloth.grid_cloth(n, 1.0, origin=(0, 1.0, 0))
    phys = cloth.default_cloth_physics_params(numSubsteps=4, iterations=4, contactRadius=0.02, maxQueryDis=0.03, gravity=[0, 0, 0])
    obj = cloth.default_cloth_object_params(density=0.2, miu=2e3, bendingStiffness=0.5, initializationType=2)
    obj.lambda_ = 2e3
    o = OracleClothScene(phys, threads=2)
    o.add_tri_mesh(v, f, obj, cloth.grid_cloth_coloring(n))
    left = np.nonzero(v[:, 0] < v[:, 0].min() + 1e-6)[0]
    right = np.nonzero(v[:, 0] > v[:, 0].max() - 1e-6)[0]
    o.add_rotator(0, left, [0, 1.0, 0], [1, 0, 0], 2.0, -1.0)
    o.add_translator(0, right, [0.0, 0.6, 0.0], 0.0, 0.05)
    o.finalize()
    o.run_frames(6)
    x = o.positions(0)
    th = 2.0 * o.cur_time
    assert abs(o.cur_time - 6 * phys.timeStep) < 1e-5
    # reference quirk kept: root = ((P - o).axis) axis is a vector from o but is subtracted from the absolute P, so the
    # turn is about the axis through the ORIGIN; `center` only enters through its component along the axis
    ry, rz = v[left, 1], v[left, 2]
    assert np.allclose(x[left, 0], v[left, 0], atol=1e-6)
    assert np.allclose(x[left, 1], np.cos(th) * ry - np.sin(th) * rz, atol=2e-6)
    assert np.allclose(x[left, 2], np.sin(th) * ry + np.cos(th) * rz, atol=2e-6)
    dt = phys.timeStep / 4
    steps = sum(1 for k in range(1, 25) if np.float32(k * dt) < 0.05)
    assert np.allclose(x[right, 1] - v[right, 1], 0.6 * dt * steps, atol=1e-5)
    assert np.isfinite(x).all() and np.abs(x - v).max() > 0.05
    assert (o.velocities(0)[left] == 0).all()


def test_scene_files_with_deformers_reach_the_solver(tmp_path):
    m, p, (v, f, sv, sf) = write_scene(str(tmp_path))
    pj = json.load(open(p))
    corner = [0, 9]
    pj["Deformers"] = [{"DeformerName": "Translator", "speed": [0.0, 0.5, 0.0], "selectedMeshes": [0], "selectedVertices": [corner],
                        "deformationEndTime": -1},
                       {"DeformerName": "Rotator", "center": [0, 0.9, 0], "axis": [0, 1, 0], "angularVelocity": 1.0, "selectedMeshes": [0],
                        "selectedVertices": [[90, 99]], "rotationEndTime": 60}]
    json.dump(pj, open(p, "w"))
    phys, meshes, pj2 = cloth_host.load_cloth_scene_json(m, p, str(tmp_path))
    o = cloth_host.build_scene(OracleClothScene, phys, meshes, deformers=pj2["Deformers"], threads=2)
    y0 = o.positions(0)[corner, 1].copy()
    o.run_frames(2)
    assert np.allclose(o.positions(0)[corner, 1] - y0, 0.5 * 2 * phys.timeStep, atol=1e-5)
    r = subprocess.run([CLOTH_HOST_BIN, m, p, str(tmp_path / "out"), "-R", str(tmp_path), "--dryRun"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    pj["Deformers"].append({"DeformerName": "Bender", "selectedMeshes": [], "selectedVertices": []})
    json.dump(pj, open(p, "w"))
    r = subprocess.run([CLOTH_HOST_BIN, m, p, str(tmp_path / "out"), "-R", str(tmp_path), "--dryRun"], capture_output=True, text=True)
    assert r.returncode == 3 and "Unrecognized deformer" in r.stderr


@pytest.mark.parametrize("interp_substep,interp_iter", [(True, 1), (True, 2), (False, 1)])
def test_oracle_collider_sequence_follows_the_reference_interpolation(interp_substep, interp_iter):
    """ColliderTrimeshSequence::update (M/SpatialQuery/ColiiderTriMeshSequence.h:59-104) restated in Python: after every frame
    the collider sits where the LAST update of that frame put it: t = (f - f0)/n + substep/(S n) + iter/(I S n) for the
    first I iterations of a substep (interpolateSubstep), or (f - f0)/n once per frame; past the last key nothing moves."""
    S, IT = 3, 4
    n = 8
    v, f = cloth.grid_cloth(n, 1.0, origin=(0, 2.0, 0))
    sv, sf = cloth.icosphere(0.2, 1)
    keys = np.stack([sv, sv + np.float32([0.3, 0.0, 0.0]), sv + np.float32([0.3, 0.6, 0.0])])
    kf = [1, 3, 6]
    phys = cloth.default_cloth_physics_params(numSubsteps=S, iterations=IT, contactRadius=0.02, maxQueryDis=0.03, gravity=[0, 0, 0])
    o = OracleClothScene(phys, threads=2)
    o.add_tri_mesh(v, f, cloth.default_cloth_object_params(), cloth.grid_cloth_coloring(n))
    c = o.add_tri_mesh(sv, sf, cloth.default_cloth_object_params(), simulated=False)
    o.set_collider_sequence(c, keys, kf, interp_substep, interp_iter)
    o.finalize()
    interp_substep = interp_substep or interp_iter > 1

    def lerp(k, t):
        t = np.float32(t)
        return keys[k] * (np.float32(1) - t) + keys[k + 1] * t

    expect = None
    first = True
    for frame in range(8):
        for sub in range(S):
            for it in range(IT):
                if first:                                   # set_state(frameId)
                    first = False
                    expect = keys[0] if frame < kf[0] else keys[-1]
                    continue
                if frame < kf[0] or frame >= kf[-1]:
                    continue
                k = max(i for i in range(len(kf)) if kf[i] <= frame)
                nfr = kf[k + 1] - kf[k]
                if interp_substep:
                    if it < interp_iter:
                        t = np.float32(frame - kf[k]) / np.float32(nfr)
                        t = t + np.float32(sub) / np.float32(S * nfr)
                        t = t + np.float32(it) / np.float32(interp_iter * S * nfr)
                        expect = lerp(k, t)
                elif sub == 0 and it == 0:
                    expect = lerp(k, np.float32(frame - kf[k]) / np.float32(nfr))
        o.run_frames(1)
        assert np.allclose(o.positions(c), expect, atol=1e-6), frame
    assert np.array_equal(o.positions(0), v)                # far away, no gravity: the sheet never moved
