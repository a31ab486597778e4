#!/usr/bin/env python
"""bench.py — VBD vertex-solves/s on B200 (BASELINE.json metric), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one frame of the hot path (VBDPhysics::runStep: numSubsteps x iterations colour sweeps + initial step,
deformers, velocity update [+ collision detection when the workload enables it]) over one synthetic scene.
Default workload at N=1: BASELINE.json configs[3], the single ~10M-tet Neo-Hookean beam twisted by two Rotator deformers
(SURVEY.md §8(d) cfg 4) — the configuration north_star's roofline and scaling targets are stated on; it fits one GPU.
  value        whole-job vertex-solves/s with the scene resident in HBM (CUDA events on the library's stream)
  e2e          the same metric through the C ABI with HOST buffers: per step H2D of positions+velocities, runStep, D2H of
               positions, all inside the timed region
  roofline     dominant kernel (k_solve_colour): algorithmic bytes (SURVEY.md §8(d) formula) / CUDA-event kernel time
  cpu_baseline the CPU oracle (a port of the reference's CPU/TBB path; the reference itself needs Eigen/oneTBB/Embree,
               absent here) on a bounded sample of the same workload, all host threads
--impl reference times that CPU oracle as the reference arm.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from gaia_b200 import capi, cloth, meshio  # noqa: E402


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# stdout carries exactly ONE JSON line: everything libraries print on fd 1 (NCCL's version banner ...) goes to stderr
_REAL_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(obj):
    _REAL_STDOUT.write(json.dumps(obj) + "\n")
    _REAL_STDOUT.flush()


from gaia_b200.workloads import WORKLOADS, build_scene, is_cloth, workload_cloth_drape, workload_drop_pile, workload_twist_beam  # noqa: E402


# --------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.samples, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            t0 = time.time()
            while not self.samples and time.time() - t0 < 3.0:      # nvidia-smi needs a moment before its first line
                time.sleep(0.02)
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            f = [x.strip() for x in line.split(",")]
            if len(f) >= 8:
                self.samples.append(f)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        mx = [float(s[2]) for s in self.samples if s[2].replace(".", "").isdigit()]
        reasons = set()
        for s in self.samples:
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.samples)}


def physical_cores():
    """distinct (physical id, core id) pairs of /proc/cpuinfo; falls back to the logical count"""
    try:
        seen, phys, core = set(), None, None
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("physical id"):
                    phys = line.split(":")[1].strip()
                elif line.startswith("core id"):
                    core = line.split(":")[1].strip()
                elif not line.strip():
                    if phys is not None and core is not None:
                        seen.add((phys, core))
                    phys = core = None
        return len(seen) or os.cpu_count()
    except OSError:
        return os.cpu_count()


# --------------------------------------------------------------------------------------------- CPU arm
def cpu_baseline(wl_name, max_seconds=25.0, threads=0):
    """Time the CPU oracle on a bounded sample of the workload: a beam of the same cross-section and material, 1/8 of the
    length, one frame (5 substeps x 20 iterations); iterations are reduced if one substep already exceeds the budget."""
    from oracle import binding, binding_cloth
    from oracle.binding import OracleScene
    from oracle.binding_cloth import OracleClothScene
    speed = binding.use_speed_build()          # -O3 -march=x86-64-v3, FMA contraction allowed (oracle/Makefile: liboracle_fast.so)
    binding_cloth.use_speed_build()
    if wl_name.startswith("cloth"):
        # same cell size, material and contact parameters on a 160 x 160 sheet (sphere scaled with it)
        sample_wl = WORKLOADS[wl_name]() if wl_name == "cloth_small" else workload_cloth_drape(160, width=16.0, sphere_level=3, preroll=0)
        OracleScene = OracleClothScene
    elif wl_name.startswith("twist_beam"):
        sample_wl = workload_twist_beam(59, 60, 60, collisions="nodcd" not in wl_name) if "10m" in wl_name or "1m" in wl_name else WORKLOADS[wl_name]()
    elif wl_name.startswith("drop_pile"):
        # same shapes, material, contact parameters and step settings on at most 32 objects, timed from the start of the drop
        full = WORKLOADS[wl_name]()
        sample_wl = workload_drop_pile(min(32, len(full["meshes"])), substeps=full["phys"].numSubsteps, iterations=full["phys"].iterations, preroll=0)
    elif wl_name == "s01":
        from gaia_b200.workloads import workload_s01
        sample_wl = workload_s01(first_frame=1, n_models=4)      # the scene's first four models (appearing at frames 1, 7, 13, 19), timed from frame 0
        sample_wl["preroll"] = 0
    else:
        sample_wl = WORKLOADS[wl_name]()
    t0 = time.time()
    o = build_scene(OracleScene, sample_wl, threads=threads)
    setup = time.time() - t0
    # calibrate with one iteration-only call, then run as many whole frames as fit
    o.reset_stats()
    t0 = time.time()
    o.run_iterations(2)
    t_it = (time.time() - t0) / 2
    per_frame = t_it * sample_wl["phys"].numSubsteps * sample_wl["phys"].iterations
    frames = max(1, int(max_seconds / max(per_frame, 1e-9)))
    frames = min(frames, 3)
    o.reset_stats()
    t0 = time.time()
    if per_frame > 2 * max_seconds:
        n_it = max(2, int(max_seconds / t_it))
        o.run_iterations(n_it)
        what = f"{n_it} iterations"
    else:
        o.run_frames(frames)
        what = f"{frames} frame(s)"
    dt = time.time() - t0
    st = o.stats()
    return {"value": st["vertexSolves"] / dt, "unit": "vertex-solves/s", "cores": o.num_threads, "physical_cores": physical_cores(), "kind": "port",
            "build": "liboracle_fast.so: -O3 -march=x86-64-v3 -ffp-contract=fast" if speed else "liboracle.so (parity build: -O3 -march=x86-64-v2 -ffp-contract=off)",
            "sample_workload": sample_wl["name"],
            "sample": f"{sample_wl['name']}: {what} in {dt:.1f}s (setup {setup:.1f}s), oracle = CPU restatement of "
                      + ("VBDClothSimulationFramework::runStep_CDEveryIter" if wl_name.startswith("cloth") else
                         "VBDPhysics::runStep_hybridCollisionHandling") + " with std::thread pool"}, st["vertexSolves"], dt


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]()
    vals, tot_solves, tot_t = [], 0, 0.0
    base = None
    for i in range(args.warmup + args.steps):
        # each step is a bounded sample (the CPU cannot run the full step in minutes); keep them short
        base, solves, dt = cpu_baseline(args.workload, max_seconds=max(4.0, 40.0 / max(1, args.steps + args.warmup)))
        if i >= args.warmup:
            tot_solves += solves; tot_t += dt
    value = tot_solves / tot_t
    base["value"] = value
    out = {"impl": "reference", "metric": "VBD vertex-solves/sec", "value": value, "unit": "vertex-solves/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(1, args.steps), "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": base["sample_workload"] + " [bounded sample of: " + wl["name"] + "]",
                      "parallelism": f"cpu: {base['cores']} threads on {base['physical_cores']} physical cores"},
           "cpu_baseline": base, "e2e": {"value": value, "unit": "vertex-solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    emit(out)


# --------------------------------------------------------------------------------------------- our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="twist_beam_10m", choices=sorted(WORKLOADS))
    ap.add_argument("--layout", default="ring", choices=["ring", "replicated"],
                    help="N > 1, one mesh: 'ring' = slabs with a one-ring ghost layer (contacts seen inside a slab and its ring), "
                         "'replicated' = solve partitioned, whole mesh on every rank as collision geometry (contacts between any two ranks)")
    ap.add_argument("--multi-colour-launch", type=int, default=None, choices=[0, 1, 2],
                    help="gaia_b200_set_multi_colour_launch mode (library default 1)")
    ap.add_argument("--refit-margin", type=float, default=None, help="lazy-refit margin of the DCD trees in typical tet edges (library default 0.2)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile-steps", type=int, default=1, help="extra untimed steps with per-kernel CUDA events for the roofline")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)

    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"

    t0 = time.time()
    wl = WORKLOADS[args.workload]()
    if args.refit_margin is not None:
        wl["refit_margin"] = args.refit_margin
        wl["name"] += f", refit margin {args.refit_margin} edges"
    if world > 1:
        if is_cloth(wl):
            raise SystemExit("the cloth workloads run on one GPU (north_star partitions configs 4 and 5 only)")
        from gaia_b200 import multigpu
        if len(wl["meshes"]) > 1:      # many objects: shard by object with ghost objects (config 5)
            scene = multigpu.build_object_sharded_scene(wl, rank, world, local_rank)
        else:                          # one big mesh: x-slab partition with a one-ring ghost layer (config 4)
            scene = (multigpu.build_replicated_scene if args.layout == "replicated" else multigpu.build_partitioned_scene)(wl, rank, world, local_rank)
    else:
        scene = build_scene(cloth.ClothScene if is_cloth(wl) else capi.Scene, wl, device=local_rank)
    if args.multi_colour_launch is not None and not is_cloth(wl):
        scene.set_multi_colour_launch(args.multi_colour_launch)
        wl["name"] += f", multi-colour launch mode {args.multi_colour_launch}"
    log(f"[bench] rank {rank}: scene '{wl['name']}' built in {time.time() - t0:.1f}s")
    stream = torch.cuda.ExternalStream(scene.cuda_stream, device=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(n):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if world > 1:
            import torch.distributed as dist
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        barrier()
        return ms

    step = lambda: scene.run_frames(1)  # noqa: E731
    if wl.get("preroll"):
        scene.run_frames(wl["preroll"])       # part of the set-up: bring the objects into contact
        scene.synchronize()
    for _ in range(args.warmup):
        step()
    scene.synchronize()
    scene.reset_stats()
    clocks = ClockSampler(local_rank)
    clocks.start()
    ms = timed(step, args.steps)
    clk = clocks.stop()
    st = scene.stats()
    solves = st["vertexSolves"]
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([solves], device="cuda", dtype=torch.float64)
        dist.all_reduce(t)
        solves = float(t.item())
    value = solves / (ms * 1e-3)

    # ---- end to end through the C ABI with host buffers
    nV = scene.total_vertices
    pos_h = scene.all_positions()
    vel_h = np.zeros_like(pos_h)
    if not is_cloth(wl):                      # the step's host buffers are page-locked once, as the reference does with its Eigen storage
        scene.register_host_buffer(pos_h); scene.register_host_buffer(vel_h)

    def e2e_step():
        scene.set_all_state(pos_h, vel_h)     # H2D of this step's inputs (pinned staging inside the library)
        scene.run_frames(1)
        scene.all_positions(out=pos_h)        # D2H of this step's result

    e2e_step()
    scene.reset_stats()
    t_wall0 = time.time()
    ms_e2e = timed(e2e_step, args.steps)
    wall_e2e = time.time() - t_wall0
    st2 = scene.stats()
    solves2 = st2["vertexSolves"]
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([solves2], device="cuda", dtype=torch.float64)
        dist.all_reduce(t)
        solves2 = float(t.item())
    nV_all = nV
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([nV], device="cuda", dtype=torch.float64)
        dist.all_reduce(t)
        nV_all = int(t.item())         # owned + ghost vertices of every partition
    e2e = {"value": solves2 / (ms_e2e * 1e-3), "unit": "vertex-solves/s", "h2d_bytes_per_step": int(2 * nV_all * 12),
           "d2h_bytes_per_step": int(nV_all * 12), "wall_ms_per_step": 1e3 * wall_e2e / args.steps}

    # ---- roofline of the dominant kernel: per-launch CUDA events on the launching stream (untimed extra steps)
    # (every rank runs the same steps: partitions advance in lock-step through the halo protocol)
    roof = None
    loop_ms = None
    if not is_cloth(wl):
        # pass A: phase timers only - the colour loop runs exactly as in the timed region (CUDA graph where it is one)
        scene.reset_stats()
        scene.set_profiling(2)
        for _ in range(max(1, args.profile_steps)):
            scene.run_frames(1)
        scene.synchronize()
        loop_ms = scene.stats()["timeCsmpMaterialSolve"] / max(1, args.profile_steps)
    # pass B: an event pair around every colour-solve launch (no host sync between launches)
    scene.reset_stats()
    scene.set_profiling(True)
    for _ in range(max(1, args.profile_steps)):
        scene.run_frames(1)
    scene.synchronize()
    sp = scene.stats()
    scene.set_profiling(False)
    if rank == 0:
        k_ms = sp["solveKernelMs"]
        launches = sp["colourSweeps"]
        alg_bytes = sp["algorithmicBytes"]
        achieved = alg_bytes / (k_ms * 1e-3) / 1e9
        phase_keys = (("timeCsmpUpdatingCollisionInfoDCD",) if is_cloth(wl) else
                      ("timeCsmpColDetectDCD", "timeCsmpInitialStep", "timeCsmpColDetectCCD", "timeCsmpMaterialSolve", "timeCsmpUpdateVelocity"))
        contacts = ({"vf": sp["numContactsVF"], "ee": sp["numContactsEE"], "detections": sp["numberCollisionDetection"]} if is_cloth(wl) else
                    {"dcd": sp["numContactsDCD"], "ccd": sp["numContactsCCD"]})
        roof = {"kernel": "gbc::k_solve_colour (cloth)" if is_cloth(wl) else "gb::k_sweep_tiles", "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "traffic": None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes / max(1, launches), "avg_launch_us": 1e3 * k_ms / max(1, launches),
                "launches_profiled": launches, "kernel_share_of_step": (k_ms / max(1, args.profile_steps)) / (ms / args.steps),
                "avg_launch_us_isolated": 1e3 * k_ms / max(1, launches),
                "phases_ms_per_step": {k: sp[k] / max(1, args.profile_steps) for k in phase_keys},
                "contacts": contacts, "detection": "DCD %s, CCD %s" % (("on" if wl["col"].allowDCD else "off"), ("on" if wl["col"].allowCCD else "off")) if not is_cloth(wl) else "VF + EE every iteration that needs it"}
        sweep_only = (not is_cloth(wl)) and sp["numContactsDCD"] + sp["numContactsCCD"] == 0 and not wl["col"].allowDCD and not wl["col"].allowCCD and world == 1
        if sweep_only and loop_ms:
            # the loop holds nothing but sweep launches: their time inside the replayed loop is the figure that adds up to the
            # step (event pairs between launches cost ~7 %, which is why the isolated figure alone gave a share above 1)
            per = loop_ms / max(1, launches / max(1, args.profile_steps))
            roof["avg_launch_us"] = 1e3 * per
            roof["achieved"] = roof["algorithmic_bytes_per_launch"] / (per * 1e-3) / 1e9
            roof["frac"] = roof["achieved"] / hbm_peak
            roof["kernel_share_of_step"] = loop_ms / (ms / args.steps)
            roof["timing"] = "colour loop timed as replayed in the step (CUDA events around the loop) / launches"
        else:
            roof["timing"] = ("CUDA event pair around every sweep launch; a multi-colour launch (one per contact-free stretch of the loop) counts as "
                              "the colour steps it covers, so avg_launch_us, algorithmic bytes and traffic are all per COLOUR STEP")
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath) and world == 1:
            try:
                with open(tpath) as f:
                    tj = json.load(f)
                roof["traffic"] = tj.get(args.workload, {}).get("dram_bytes_per_launch")
                if roof["traffic"]:
                    # this layout moves fewer bytes than the reference-layout formula counts: also report the real DRAM rate
                    roof["dram_achieved"] = roof["traffic"] / (roof["avg_launch_us"] * 1e-6) / 1e9
                    roof["dram_frac"] = roof["dram_achieved"] / hbm_peak
            except (OSError, ValueError):
                pass

    if is_cloth(wl):
        per_iter = 656.0 * nV
    else:
        lb = scene.layout_bytes()
        per_iter = lb["records"] + lb["neighbours"] + lb["vertex_state"] + lb["tiles"]
    l2_note = ("inputs larger than L2 (one iteration streams %.2f GB of records, neighbour entries and vertex state vs 126 MB of L2), no flush needed"
               % (per_iter / 1e9)) if per_iter > 2 * 126e6 else \
              ("inputs SMALLER than L2 (%.1f MB per iteration vs 126 MB): the step is latency-bound, L2-resident by nature of the scene; no flush"
               % (per_iter / 1e6))
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu, _, _ = cpu_baseline(args.workload)

    if rank == 0:
        out = {"metric": "VBD vertex-solves/sec", "value": value, "unit": "vertex-solves/s", "n_gpus": world, "steps": args.steps,
               "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
               "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": wl["name"], "parallelism": (f"{'objects sharded' if len(wl['meshes']) > 1 else ('x-slab partition, ' + ('whole mesh replicated as collision geometry' if args.layout == 'replicated' else 'one-ring ghosts'))} over {world} GPU(s)") if world > 1 else "1 GPU",
                          "l2": l2_note,
                          "sim_steps_per_s": st["substeps"] / (ms * 1e-3), "frames_per_s": args.steps / (ms * 1e-3)},
               "clocks": clk, "e2e": e2e, "gpu_launches": int(st["kernelLaunches"]), "roofline": roof, "cpu_baseline": cpu}
        emit(out)
    scene.close()
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
