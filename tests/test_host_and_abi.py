"""CPU: the drop-in boundary. The C-ABI library loads and exports every symbol include/gaia_b200.h declares; without a
GPU it fails loudly (no CPU fallback); the C++ host adapter reads the reference's Models.json / Parameters.json schema
(including the reference's own checked-in S01 experiment when /root/reference is mounted); the Python scene readers
agree with it."""
import ctypes as C
import json
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, has_gpu
from gaia_b200 import capi, meshio

HOST_BIN = os.path.join(ROOT, "gaia_b200", "host", "gaia_vbd_b200")


def declared_symbols(header="gaia_b200.h"):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gaia_b200_\w+)\s*\(", text)))


def test_library_exports_every_declared_cloth_symbol_and_layouts_match():
    from gaia_b200 import cloth
    lib = capi.load_library()
    names = declared_symbols("gaia_b200_cloth.h")
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/gaia_b200_cloth.h but not exported"
    text = open(os.path.join(ROOT, "include", "gaia_b200_cloth.h")).read()
    for cname, cls in (("gaia_cloth_physics_params", cloth.ClothPhysicsParams), ("gaia_cloth_object_params", cloth.ClothObjectParams),
                       ("gaia_cloth_vf_contact", cloth.ClothVFContact), ("gaia_cloth_ee_contact", cloth.ClothEEContact),
                       ("gaia_cloth_stats", cloth.ClothStats)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (cname, cname), text, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        fields = []
        for ln in body.split(";"):
            m = re.match(r"\s*\w+\s+(.*)", ln, re.S)
            if m:
                fields += [re.match(r"(\w+)", x.strip()).group(1) for x in m.group(1).split(",")]
        assert fields == [n.rstrip("_") for n, _ in cls._fields_], cname
    p = cloth.ClothPhysicsParams(); lib.gaia_b200_cloth_default_physics_params(C.byref(p))
    assert bytes(p) == bytes(cloth.default_cloth_physics_params())
    o = cloth.ClothObjectParams(); lib.gaia_b200_cloth_default_object_params(C.byref(o))
    assert bytes(o) == bytes(cloth.default_cloth_object_params())


@pytest.mark.skipif(has_gpu(), reason="checks the no-GPU failure mode")
def test_cloth_create_fails_loudly_without_gpu():
    from gaia_b200 import cloth
    with pytest.raises(capi.GaiaError) as e:
        cloth.ClothScene()
    assert "no CPU path" in str(e.value)


def test_library_exports_every_declared_symbol():
    lib = capi.load_library()
    names = declared_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/gaia_b200.h but not exported"
    assert lib.gaia_b200_abi_version() == 1


def test_struct_layouts_match_the_header():
    # the ctypes mirrors must have one field per header field, in order
    text = open(os.path.join(ROOT, "include", "gaia_b200.h")).read()
    for cname, cls in (("gaia_physics_params", capi.PhysicsParams), ("gaia_collision_params", capi.CollisionParams),
                       ("gaia_object_params", capi.ObjectParams), ("gaia_contact", capi.Contact), ("gaia_stats", capi.Stats)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (cname, cname), text, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        fields = [re.match(r"\s*\w+\s+(\w+)", ln).group(1) for ln in body.split(";") if re.match(r"\s*\w+\s+\w+", ln)]
        assert fields == [n for n, _ in cls._fields_], cname
    lib = capi.load_library()
    p = capi.PhysicsParams(); lib.gaia_b200_default_physics_params(C.byref(p))
    q = capi.default_physics_params()
    assert bytes(p) == bytes(q)
    c = capi.CollisionParams(); lib.gaia_b200_default_collision_params(C.byref(c))
    assert bytes(c) == bytes(capi.default_collision_params())
    o = capi.ObjectParams(); lib.gaia_b200_default_object_params(C.byref(o))
    assert bytes(o) == bytes(capi.default_object_params())


@pytest.mark.skipif(has_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_create_fails_loudly():
    with pytest.raises(capi.GaiaError) as e:
        capi.Scene()
    assert "no CPU path" in str(e.value)


def write_scene(tmp, n_frames=4, ext="bin", recovery_step=2, collisions=True):
    """Two Kuhn blocks + the reference's JSON schema, with ${REPO_ROOT} and Windows separators on purpose."""
    os.makedirs(os.path.join(tmp, "meshes"), exist_ok=True)
    models = []
    for i, (dims, h, org) in enumerate((((3, 2, 2), 0.25, (0, 0.3, 0)), ((2, 2, 2), 0.2, (0.1, 0.72, 0.05)))):
        v, t = meshio.kuhn_beam(*dims, h=h, origin=(0, 0, 0))
        path = os.path.join(tmp, "meshes", f"block{i}.t")
        meshio.save_t(path, v, t)
        with open(path + ".vertexColoring.json", "w") as f:
            json.dump([c.tolist() for c in meshio.kuhn_coloring(*dims)], f)
        models.append({"materialName": "NeoHookean", "path": "${REPO_ROOT}\\meshes\\block%d.t" % i,
                       "verticesColoringCategoriesPath": "${REPO_ROOT}/meshes/block%d.t.vertexColoring.json" % i,
                       "translation": list(org), "scale": [1, 1, 1], "rotation": [0, 0.1 * i, 0], "initialVelocity": [0, -2.0 * i, 0],
                       "density": 50, "miu": 1e5, "lmbd": 1e6, "frictionDynamic": 0.2, "frictionEpsV": 0.01, "dampingShear": 1e-3,
                       "fixedPoints": [], "frameToAppear": 0})
    params = {"PhysicsParams": {"gravity": [0, -10, 0], "worldBounds": [[-10, 0, -10], [10, 20, 10]], "numFrames": n_frames, "numSubsteps": 2,
                                "iterations": 5, "timeStep": 0.0166666, "outputExt": ext, "outputRecoveryStateStep": recovery_step,
                                "binaryModeVisualizationSteps": 2, "collisionStiffness": 1e5, "useGPU": True, "debugVerboseLvl": 0,
                                "outputStatistics": True},
              "CollisionParams": {"allowDCD": collisions, "allowCCD": collisions}, "ViewerParams": {"enableViewer": False}}
    with open(os.path.join(tmp, "Models.json"), "w") as f:
        json.dump({"Models": models}, f)
    with open(os.path.join(tmp, "Parameters.json"), "w") as f:
        json.dump(params, f)
    return os.path.join(tmp, "Models.json"), os.path.join(tmp, "Parameters.json")


def test_host_adapter_dry_run_reads_the_schema(tmp_path):
    m, p = write_scene(str(tmp_path))
    r = subprocess.run([HOST_BIN, m, p, str(tmp_path / "out"), "-R", str(tmp_path), "--dryRun"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "2 models, 63 vertices, 120 tets, numFrames 4, numSubsteps 2, iterations 5, outputExt bin" in r.stdout
    # the Python reader sees the same scene
    phys, col, meshes = meshio.load_scene_json(m, p, str(tmp_path))
    assert phys.numSubsteps == 2 and phys.iterations == 5 and col.allowDCD == 1 and len(meshes) == 2
    assert sum(len(x[0]) for x in meshes) == 63 and sum(len(x[1]) for x in meshes) == 120
    assert abs(meshes[1][2].initialVelocity[1] + 2.0) < 1e-7 and meshes[1][2].frameToAppear == 0


def test_host_adapter_reads_the_bowl_ground_and_initial_guess_5(tmp_path):
    """useBowlGround with bowlCenter / bowlRadius / bowlCap (M/Parameters/PhysicsParameters.h:110, 130-131, 148) and
    initializationType 5 are accepted by both readers (they were refused until round 2)"""
    m, p = write_scene(str(tmp_path))
    pj = json.load(open(p))
    pj["PhysicsParams"].update(useBowlGround=True, bowlCenter=[0.5, 40.0, -0.25], bowlRadius=45.5, bowlCap=True)     # a wide bowl: the blocks start inside
    json.dump(pj, open(p, "w"))
    mj = json.load(open(m))
    mj["Models"][0]["initializationType"] = 5
    json.dump(mj, open(m, "w"))
    r = subprocess.run([HOST_BIN, m, p, str(tmp_path / "out"), "-R", str(tmp_path), "--dryRun"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "bowl ground, centre (0.5, 40, -0.25), radius 45.5, cap 1" in r.stdout
    phys, col, meshes = meshio.load_scene_json(m, p, str(tmp_path))
    assert phys.useBowlGround == 1 and meshes[0][2].initializationType == 5
    assert meshio.bowl_ground_from_json(pj["PhysicsParams"]) == ([0.5, 40.0, -0.25], 45.5, True)
    assert meshio.bowl_ground_from_json({"useBowlGround": True}) == ([0.0, 0.0, 0.0], 1.0, False) and meshio.bowl_ground_from_json({}) is None
    # the oracle runs such a scene
    from oracle.binding import OracleScene
    o = OracleScene(phys, col, threads=1)
    for v, t, obj, colors, fixed in meshes:
        o.add_tet_mesh(v, t, obj, colors, fixed)
    o.finalize()
    o.set_bowl_ground(*meshio.bowl_ground_from_json(pj["PhysicsParams"]))
    o.run_frames(2)
    assert np.isfinite(o.positions(0)).all()


@pytest.mark.skipif(has_gpu(), reason="checks the no-GPU failure mode")
def test_host_adapter_fails_loudly_without_gpu(tmp_path):
    m, p = write_scene(str(tmp_path))
    r = subprocess.run([HOST_BIN, m, p, str(tmp_path / "out"), "-R", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU path" in r.stderr
    r = subprocess.run([HOST_BIN, m, p, str(tmp_path / "out"), "--CPU"], capture_output=True, text=True)
    assert r.returncode == 2 and "no CPU path" in r.stderr


@pytest.mark.skipif(not meshio.reference_available(), reason="/root/reference only exists in the build container")
def test_reference_s01_inputs_parse_unchanged():
    d = ("/root/reference/Simulator/VBDDynamics/ParameterGen/Parameters/S01_Experiment_HybridModelsDrop_sequentiallyAppear/"
         "Test_64HybridModels_2steps_40Iters_stiff_Friction")
    r = subprocess.run([HOST_BIN, d + "/Models.json", d + "/Parameters.json", "/tmp/unused", "-R", "/root/reference", "--dryRun"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "32 models, 87339 vertices, 330335 tets, numFrames 720, numSubsteps 4, iterations 20, outputExt ply" in r.stdout
    phys, col, meshes = meshio.load_scene_json(d + "/Models.json", d + "/Parameters.json", "/root/reference")
    assert len(meshes) == 32 and phys.numSubsteps == 4 and phys.iterations == 20
    assert [m[2].frameToAppear for m in meshes[:3]] == [120, 126, 132]
    for v, t, obj, colors, fixed in meshes[:4]:
        assert meshio.check_coloring(len(v), t, colors)


def test_t_file_roundtrip_and_colourings(tmp_path):
    v, t = meshio.kuhn_beam(3, 2, 1, h=0.5)
    path = str(tmp_path / "a.t")
    meshio.save_t(path, v, t)
    v2, t2 = meshio.load_t(path)
    assert np.array_equal(v, v2) and np.array_equal(t, t2)
    for colors in (meshio.kuhn_coloring(3, 2, 1), meshio.greedy_coloring(len(v), t)):
        assert meshio.check_coloring(len(v), t, colors)
    bad = [np.arange(len(v), dtype=np.int32)]
    assert not meshio.check_coloring(len(v), t, bad)
