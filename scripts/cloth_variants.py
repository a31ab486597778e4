"""Time the cloth colour sweep variants (GAIA_B200_CLOTH_VARIANT) on the cloth_1m workload: us per launch from CUDA events."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from gaia_b200 import cloth

wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "cloth_1m"]()
ref = None
for variant in (sys.argv[2].split(",") if len(sys.argv) > 2 else ("1", "0", "2", "3", "4")):
    os.environ["GAIA_B200_CLOTH_VARIANT"] = variant
    s = bench.build_scene(cloth.ClothScene, wl, device=0)
    s.run_frames(1)
    s.synchronize(); s.reset_stats(); s.set_profiling(True)
    t0 = time.time(); s.run_frames(1); s.synchronize(); wall = time.time() - t0
    st = s.stats()
    x = s.positions(0)
    if ref is None:
        ref = x
    print(f"variant {variant}: {1e3 * st['solveKernelMs'] / st['colourSweeps']:.2f} us per colour launch, frame {1e3 * wall:.1f} ms (profiling on), "
          f"detect {st['timeCsmpUpdatingCollisionInfoDCD']:.1f} ms / {st['numberCollisionDetection']} detections, max |x - x_variant1| = {np.abs(x - ref).max():.2e}", flush=True)
    s.close()
