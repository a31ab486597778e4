"""-m gpu: the C++ host adapter (gaia_vbd_b200, the VBDDynamics-shaped driver) end to end: Models.json / Parameters.json in,
per-frame output meshes out. Its frames are bit-identical to the same scene driven through the Python binding (same
library underneath), the file formats follow M/IO/FileIO.cpp, and a run recovered from a RecoveryStates file continues
bit-identically."""
import json
import os
import subprocess

import numpy as np
import pytest

from test_host_and_abi import write_scene, HOST_BIN
from gaia_b200 import capi, meshio

pytestmark = pytest.mark.gpu


def _python_run(m, p, root, frames):
    phys, col, meshes = meshio.load_scene_json(m, p, root)
    s = capi.Scene(phys, col)
    meshio.add_scene_meshes(s, meshes)
    if len(meshes) > 1:
        s.set_solver_lanes(8)          # the host driver's rule for small many-object scenes (VBDPhysicsB200::initialize)
    out = [np.concatenate([s.surface_positions(i) for i in range(len(meshes))])]
    for _ in range(frames):
        s.run_frames(1)
        out.append(np.concatenate([s.surface_positions(i) for i in range(len(meshes))]))
    return out, s


def test_binary_outputs_match_python_binding(tmp_path):
    m, p = write_scene(str(tmp_path), n_frames=4, ext="bin")
    out = tmp_path / "out"
    r = subprocess.run([HOST_BIN, m, p, str(out), "-R", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    frames, scene = _python_run(m, p, str(tmp_path), 4)
    for f in range(5):
        a = np.fromfile(out / ("A%08d.bin" % f), np.float32).reshape(-1, 3)
        assert np.array_equal(a, frames[f]), f
    # template meshes + periodic visualisation PLY + recovery states + statistics
    assert (out / "TemplateMesh" / "TemplateMesh_Model_000000_block0.ply").exists()
    assert (out / "TemplateMesh" / "TemplateMesh_Model_000001_block1.ply").exists()
    assert (out / "A00000002.ply").exists() and not (out / "A00000001.ply").exists()
    st = json.load(open(out / "RecoveryStates" / "A00000002.json"))
    assert st["frameId"] == 2 and st["format"] == "binary" and st["meshesState"]["nVerts"] == [36, 27]
    pos = np.fromfile(out / "RecoveryStates" / "A00000002.pos", np.float32)
    assert pos.size == 63 * 3
    stats = json.load(open(out / "Statistics" / "A00000003.json"))
    assert stats["substeps"] == 2 and stats["vertexSolves"] == 63 * 2 * 5
    # ASCII PLY layout (saveAsPly): header, %f vertices, "3 i j k" faces indexing the concatenated surface vertices
    lines = open(out / "A00000002.ply").read().split("\n")
    assert lines[0] == "ply" and lines[2] == "comment Mocap generated" and lines[3] == "element vertex %d" % len(frames[2])
    hdr = lines.index("end_header")
    v0 = np.array([float(x) for x in lines[hdr + 1].split()])
    assert np.allclose(v0, frames[2][0], atol=1e-6)
    nF = int(lines[7].split()[2])
    faces = np.array([[int(x) for x in ln.split()[1:]] for ln in lines[hdr + 1 + len(frames[2]): hdr + 1 + len(frames[2]) + nF]])
    assert faces.min() == 0 and faces.max() == len(frames[2]) - 1
    scene.close()


def test_recovery_state_continues_bit_identically(tmp_path):
    m, p = write_scene(str(tmp_path), n_frames=6, ext="bin", recovery_step=3)
    full = tmp_path / "full"
    r = subprocess.run([HOST_BIN, m, p, str(full), "-R", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    part = tmp_path / "part"
    r = subprocess.run([HOST_BIN, m, p, str(part), "-R", str(tmp_path), "-r", str(full / "RecoveryStates" / "A00000003.json")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Recovering state from" in r.stdout
    for f in (4, 5, 6):
        a = np.fromfile(full / ("A%08d.bin" % f), np.float32)
        b = np.fromfile(part / ("A%08d.bin" % f), np.float32)
        assert np.array_equal(a, b), f
    assert not (part / "A00000001.bin").exists()


def test_ply_output_and_text_recovery_state(tmp_path):
    m, p = write_scene(str(tmp_path), n_frames=2, ext="ply", recovery_step=1)
    par = json.load(open(p)); par["PhysicsParams"]["outputRecoveryStateBinary"] = False; par["PhysicsParams"]["saveAllModelsTogether"] = False
    json.dump(par, open(p, "w"))
    out = tmp_path / "out"
    r = subprocess.run([HOST_BIN, m, p, str(out), "-R", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert (out / "A00000002_Model_000001.ply").exists()
    st = json.load(open(out / "RecoveryStates" / "A00000002.json"))
    assert st["format"] == "text" and len(st["meshesState"]) == 2 and len(st["meshesState"][0]["position"]) == 36
    frames, scene = _python_run(m, p, str(tmp_path), 2)
    allpos = np.concatenate([scene.positions(i) for i in range(2)])
    txt = np.array(st["meshesState"][0]["position"] + st["meshesState"][1]["position"], np.float32)
    assert np.allclose(txt, allpos, rtol=0, atol=1e-6)
    scene.close()


def test_line_search_key_and_short_recovery_state(tmp_path):
    """PhysicsParams.applyLocalLineSearch (+ tau, tau_GD, backtracingLineSearchMaxIters) switches the host's scene to the
    line-search sweep: frames identical to the Python binding with set_line_search, different from the plain run. And a
    text recovery state whose arrays are shorter than the mesh is refused instead of being read past its end."""
    m, p = write_scene(str(tmp_path), n_frames=3, ext="bin")
    par = json.load(open(p))
    par["PhysicsParams"].update({"applyLocalLineSearch": True, "tau": 0.5, "tau_GD": 0.1, "backtracingLineSearchMaxIters": 8, "stepSize": 2.5})
    json.dump(par, open(p, "w"))
    out = tmp_path / "out"
    r = subprocess.run([HOST_BIN, m, p, str(out), "-R", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    phys, col, meshes = meshio.load_scene_json(m, p, str(tmp_path))
    assert abs(phys.stepSize - 2.5) < 1e-7
    runs = {}
    for ls in (True, False):
        s = capi.Scene(phys, col)
        meshio.add_scene_meshes(s, meshes)
        s.set_solver_lanes(8)
        if ls:
            s.set_line_search(True, 0.5, 0.1, 8)
        s.run_frames(3)
        runs[ls] = np.concatenate([s.surface_positions(i) for i in range(len(meshes))])
        if ls:
            assert s.line_search_shrinks() > 0
        s.close()
    a = np.fromfile(out / "A00000003.bin", np.float32).reshape(-1, 3)
    assert np.array_equal(a, runs[True]) and not np.array_equal(a, runs[False])
    # short text recovery state
    st = {"frameId": 1, "curTime": 0.0166666, "format": "text",
          "meshesState": [{"position": [[0, 0, 0]] * 5, "velocities": [[0, 0, 0]] * 5}, {"position": [], "velocities": []}]}
    json.dump(st, open(tmp_path / "short.json", "w"))
    r = subprocess.run([HOST_BIN, m, p, str(tmp_path / "out2"), "-R", str(tmp_path), "-r", str(tmp_path / "short.json")], capture_output=True, text=True)
    assert r.returncode == 3 and "vertex count mismatch" in r.stderr


@pytest.mark.gpu
def test_surface_snapshots_are_ordered_behind_the_step_and_match_the_per_mesh_getter():
    """row f4: gaia_b200_surface_snapshot_begin / _wait - the device-side surface gather into the page-locked two-slot ring.
    A snapshot taken before a frame keeps showing that state after the next frame was enqueued and snapshotted into the other
    slot; its contents are the per-mesh surface positions concatenated (writeAllToBinary order)."""
    from gaia_b200 import capi, meshio
    from conftest import make_contact_scene
    g = make_contact_scene(capi.Scene, n_objects=3)
    g.run_frames(2)
    before = np.concatenate([g.surface_positions(m) for m in range(3)])
    g.surface_snapshot_begin(0)
    g.run_frames(1)                       # enqueued behind the snapshot
    g.surface_snapshot_begin(1)
    s0, s1 = g.surface_snapshot_wait(0), g.surface_snapshot_wait(1)
    after = np.concatenate([g.surface_positions(m) for m in range(3)])
    assert np.array_equal(s0, before) and np.array_equal(s1, after) and not np.array_equal(s0, s1)
    for m in range(3):                    # the per-mesh getter picks the surface out of all positions the same way
        assert np.array_equal(g.surface_positions(m), g.positions(m)[g.surface_vertex_ids(m)])
