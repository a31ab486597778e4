#!/usr/bin/env python
"""gaia_vbd_b200 --gpus N against gaia_vbd_b200 on one GPU: a twisting beam (two Rotator deformers, the reference's JSON schema)
must come out byte-identical in every frame file. Usage (on a box with >= N GPUs): python scripts/check_cpp_multigpu.py [N] [frames]"""
import filecmp
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gaia_b200 import meshio  # noqa: E402

HOST = os.path.join(ROOT, "gaia_b200", "host", "gaia_vbd_b200")


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    frames = int(sys.argv[2]) if len(sys.argv) > 2 else 6
    nx, ny, nz, h = 48, 8, 8, 0.05
    tmp = tempfile.mkdtemp(prefix="gaia_cpp_mgpu_")
    v, t = meshio.kuhn_beam(nx, ny, nz, h=h, origin=(0.0, 1.0, 0.0))
    meshio.save_t(os.path.join(tmp, "beam.t"), v, t)
    with open(os.path.join(tmp, "beam.t.vertexColoring.json"), "w") as f:
        json.dump([c.tolist() for c in meshio.kuhn_coloring(nx, ny, nz)], f)
    gx = np.arange(len(v)) // ((ny + 1) * (nz + 1))
    left, right = np.nonzero(gx == 0)[0].tolist(), np.nonzero(gx == nx)[0].tolist()
    cy, cz = 1.0 + 0.5 * ny * h, 0.5 * nz * h
    models = [{"materialName": "NeoHookean", "path": "${REPO_ROOT}/beam.t", "verticesColoringCategoriesPath": "${REPO_ROOT}/beam.t.vertexColoring.json",
               "translation": [0, 0, 0], "scale": [1, 1, 1], "rotation": [0, 0, 0], "initialVelocity": [0, 0, 0], "density": 100, "miu": 5e3,
               "lmbd": 1e5, "dampingHydrostatic": 1e-6, "dampingDeviatoric": 1e-6, "fixedPoints": [], "frameToAppear": 0}]
    deformers = [{"DeformerName": "Rotator", "center": [0, cy, cz], "axis": [1, 0, 0], "angularVelocity": 0.785398, "rotationEndTime": -1,
                  "selectedMeshes": [0], "selectedVertices": [left]},
                 {"DeformerName": "Rotator", "center": [0, cy, cz], "axis": [1, 0, 0], "angularVelocity": -0.785398, "rotationEndTime": -1,
                  "selectedMeshes": [0], "selectedVertices": [right]}]
    params = {"PhysicsParams": {"gravity": [0, 0, 0], "worldBounds": [[-100, -100, -100], [100, 100, 100]], "numFrames": frames, "numSubsteps": int(os.environ.get("GAIA_CHECK_SUBSTEPS", "5")),
                                "iterations": int(os.environ.get("GAIA_CHECK_ITERS", "20")), "timeStep": 0.0166666, "outputExt": "bin", "outputRecoveryState": False, "binaryModeVisualizationSteps": 1000,
                                "useGPU": True, "debugVerboseLvl": 1},
              "CollisionParams": {"allowDCD": bool(int(os.environ.get("GAIA_CHECK_DCD", "0"))), "allowCCD": False}, "ViewerParams": {"enableViewer": False}, "Deformers": deformers}
    layout = os.environ.get("GAIA_CHECK_LAYOUT", "ring")
    if os.environ.get("GAIA_CHECK_FOLDED", "0") == "1":
        # a beam folded onto itself, the upper arm's end pressed onto the lower arm: self-contact between far-apart parts of the mesh
        from gaia_b200 import workloads
        wl = workloads.workload_folded_beam(96, 8, 8, 0.05, 0.03, 0.6, 2, 10)
        fv, ft, fobj, fcol, ffixed = wl["meshes"][0]
        meshio.save_t(os.path.join(tmp, "beam.t"), fv, ft)
        with open(os.path.join(tmp, "beam.t.vertexColoring.json"), "w") as f:
            json.dump([c.tolist() for c in fcol], f)
        models[0].update({"density": 100, "miu": 5e4, "lmbd": 5e5, "dampingHydrostatic": 1e-5, "dampingDeviatoric": 1e-5, "frictionDynamic": 0.2,
                          "fixedPoints": [int(x) for x in ffixed]})
        tip = wl["deformers"][0][1]
        params["Deformers"] = [{"DeformerName": "Translator", "speed": [0.0, -0.6, 0.0], "deformationEndTime": -1, "selectedMeshes": [0],
                                "selectedVertices": [[int(x) for x in tip]]}]
        params["PhysicsParams"].update({"gravity": [0, -10, 0], "worldBounds": [[-10, 0, -10], [10, 20, 10]], "numSubsteps": 2, "iterations": 10,
                                        "collisionStiffness": 1e5})
        params["CollisionParams"] = {"allowDCD": True, "allowCCD": True}
    with open(os.path.join(tmp, "Models.json"), "w") as f:
        json.dump({"Models": models}, f)
    with open(os.path.join(tmp, "Parameters.json"), "w") as f:
        json.dump(params, f)
    outs = []
    for g in (1, n):
        out = os.path.join(tmp, f"out{g}")
        r = subprocess.run([HOST, os.path.join(tmp, "Models.json"), os.path.join(tmp, "Parameters.json"), out, "-R", tmp, "--gpus", str(g), "--layout", layout],
                           capture_output=True, text=True, timeout=600)
        print(f"--gpus {g}: rc={r.returncode}", r.stdout.strip().splitlines()[-1:] , r.stderr.strip()[-300:])
        if r.returncode:
            sys.exit(1)
        outs.append(out)
    same, moved = True, 0.0
    first = np.fromfile(os.path.join(outs[0], "A%08d.bin" % 0), np.float32)
    for fr in range(frames + 1):
        name = "A%08d.bin" % fr
        a, b = os.path.join(outs[0], name), os.path.join(outs[1], name)
        ok = os.path.exists(a) and os.path.exists(b) and filecmp.cmp(a, b, shallow=False)
        same &= ok
        xa = np.fromfile(a, np.float32)
        moved = max(moved, float(np.abs(xa - first).max()))
        if not ok and os.path.exists(b):
            xb = np.fromfile(b, np.float32)
            d = np.abs(xa - xb).reshape(-1, 3).max(1) if len(xa) == len(xb) else None
            print(f"  frame {fr}: sizes {len(xa)} / {len(xb)}", "" if d is None else f"max diff {d.max():.3e} at surface vertex {int(d.argmax())}, {int((d > 0).sum())} vertices differ")
            if d is not None and fr == 1:
                x0 = first.reshape(-1, 3)[:, 0]
                for lo in np.arange(0.0, x0.max() + 1e-6, x0.max() / 8):
                    sel = (x0 >= lo) & (x0 < lo + x0.max() / 8 + 1e-9)
                    print(f"    x in [{lo:.2f}, {lo + x0.max() / 8:.2f}): {int((d[sel] > 0).sum())} of {int(sel.sum())} differ, max {d[sel].max():.2e}")
    print(f"[check_cpp_multigpu] gpus={n} layout={layout} folded={os.environ.get('GAIA_CHECK_FOLDED', '0')} frames={frames} byte_identical={same} max_displacement={moved:.4f}")
    sys.exit(0 if same and moved > 1e-3 else 1)


if __name__ == "__main__":
    main()
