"""Run the same many-object scene twice on one GPU: are the trajectories bit-identical?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from gaia_b200 import capi

name = sys.argv[1] if len(sys.argv) > 1 else "drop_pile_8"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 40
for dcd, ccd in ((1, 1), (1, 0), (0, 1), (0, 0)):
    res = []
    for rep in range(2):
        wl = bench.WORKLOADS[name]()
        wl["col"].allowDCD = dcd; wl["col"].allowCCD = ccd
        s = bench.build_scene(capi.Scene, wl, device=0)
        hist = []
        for f in range(frames):
            s.run_frames(1)
            hist.append(s.all_positions().copy())
        res.append(hist)
        st = s.stats(); s.close()
    first = next((f for f in range(frames) if not np.array_equal(res[0][f], res[1][f])), None)
    print(f"DCD {dcd} CCD {ccd}: first differing frame = {first}, final max diff = {np.abs(res[0][-1] - res[1][-1]).max():.3e}, contacts {st['numContactsDCD']}/{st['numContactsCCD']}", flush=True)
