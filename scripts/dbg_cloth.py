import os, sys
import numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import test_gpu_cloth_parity as T
energy = 0
scenes = {}
for mode in ("0", "1", "2", "3"):
    os.environ["GAIA_B200_CLOTH_EXACT_ORDER"] = mode
    g, o = T.two_layer_scene(phys_kw=dict(contactEnergyTypes=energy))
    scenes[mode] = g
for it in range(1, 4):
    o.run_iterations(1)
    for mode, g in scenes.items():
        g.run_iterations(1)
        print(f"call {it} mode {mode}: " + " ".join(f"{np.abs(g.positions(m) - o.positions(m)).max():.3e}" for m in range(2)))
