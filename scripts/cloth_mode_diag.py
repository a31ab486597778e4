"""Diagnostic: deviation of the cloth kernel variants from the oracle on the key-framed collider scene.
GAIA_B200_CLOTH_EXACT_ORDER: unset = fast kernels, 1 = exact both, 2 = exact contact terms only, 3 = exact colour solve only."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_gpu_cloth_parity as T  # noqa: E402

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 6
res = {}
for mode in ("", "1", "2", "3"):
    if mode:
        os.environ["GAIA_B200_CLOTH_EXACT_ORDER"] = mode
    else:
        os.environ.pop("GAIA_B200_CLOTH_EXACT_ORDER", None)
    g, o, _ = T.moving_collider_scene(True, 1)
    out = []
    for fr in range(frames):
        g.run_frames(1); o.run_frames(1)
        d = np.abs(g.positions(0) - o.positions(0))
        out.append((float(d.max()), float(d.mean()), g.stats()["numberCollisionDetection"], o.stats()["numberCollisionDetection"]))
    res[mode or "fast"] = out
for k, v in res.items():
    print(k, " ".join("%.2e/%.1e(%d,%d)" % x for x in v))
