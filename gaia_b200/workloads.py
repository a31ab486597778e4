"""Synthetic and reference-fixture workloads of BASELINE.json's configs, shared by bench.py, the
parity tests and the scripts. Each builder returns a dict {name, meshes, phys, col, deformers[, preroll, kind]}; `build_scene`
adds it to any scene class that has the C ABI's calls (the product's `capi.Scene`, the oracle's `OracleScene`, ...)."""
import gzip
import os

import numpy as np

from . import capi, cloth, meshio

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_ROOT = os.path.join(REPO, "tests", "golden", "ref_root")     # reference-held input files (see its README.md)


def workload_twist_beam(nx=472, ny=60, nz=60, h=0.02, collisions=False):
    """SURVEY.md §8(d) cfg 4: regular hex grid split into 6 Kuhn tets per cell, both end caps driven by Rotator deformers
    (+-pi/4 rad/s about x), NeoHookean mu=5e3 lambda=1e5 density 100 damping 1e-6, gravity 0, 5 substeps x 20 iterations."""
    verts, tets = meshio.kuhn_beam(nx, ny, nz, h=h, origin=(0.0, 1.0, 0.0))
    colors = meshio.kuhn_coloring(nx, ny, nz)
    phys = capi.default_physics_params(gravity=[0, 0, 0], numSubsteps=5, iterations=20, collisionStiffness=2e4,
                                       intermediateCollisionIterations=5 if collisions else -1,
                                       worldBoundsMin=[-100, -100, -100], worldBoundsMax=[100, 100, 100])
    col = capi.default_collision_params(allowDCD=1 if collisions else 0, allowCCD=0)
    obj = capi.default_object_params(miu=5e3, lmbd=1e5, density=100, dampingHydrostatic=1e-6, dampingDeviatoric=1e-6)
    gx = np.arange(len(verts)) // ((ny + 1) * (nz + 1))
    left = np.nonzero(gx == 0)[0].astype(np.int32)
    right = np.nonzero(gx == nx)[0].astype(np.int32)
    cy, cz = 1.0 + 0.5 * ny * h, 0.5 * nz * h
    deformers = [("rotator", left, (0.0, cy, cz), (1.0, 0.0, 0.0), np.pi / 4, -1.0),
                 ("rotator", right, (0.0, cy, cz), (1.0, 0.0, 0.0), -np.pi / 4, -1.0)]
    name = (f"twist_beam {nx}x{ny}x{nz} cells ({len(tets)} tets, {len(verts)} verts), 5 substeps x 20 iters, 4 colours, "
            + ("DCD on (every substep + every 5 iterations), CCD off" if collisions else "collision detection off"))
    return dict(name=name, meshes=[(verts, tets, obj, colors, np.zeros(0, np.int32))], phys=phys, col=col, deformers=deformers)


_BLOB_CACHE = {}


def workload_drop_pile(n_objects=32, grid=(4, 4), substeps=4, iterations=20, preroll=40, seed=123):
    """SURVEY.md §8(d) cfg 2 (S01_Experiment_HybridModelsDrop), synthetic stand-in: n_objects tet models cycling four
    blob shapes (~2-3.5k vertices, 8 colours each, like teapot / bunny / octopus / armadillo) dropping into a box with
    self / inter-object collision: S01's step settings (4 substeps x 20 iterations, DCD + CCD, collisionStiffness 1e5,
    friction 0.2 / epsV 0.01, material mu 1e5 lambda 1e6 density 50). All objects are present from frame 0 (stacked
    layers); `preroll` untimed frames bring them into contact before anything is measured."""
    rng = np.random.RandomState(seed)
    shapes = [(16, (1, 1, 1), 2.0), (18, (1, 0.6, 0.8), 2.0), (16, (1, 1, 0.55), 4.0), (17, (0.75, 1, 0.75), 1.5)]
    meshes = []
    h = 0.1
    per_layer = grid[0] * grid[1]
    for i in range(n_objects):
        key = i % len(shapes)
        if key not in _BLOB_CACHE:
            n, r, pw = shapes[key]
            v, t = meshio.kuhn_blob(n, r, h, pw)
            # coloured by the library's restatement of the reference tool (`-a mcs`, no balancing: what the reference ships)
            _BLOB_CACHE[key] = (v, t, capi.color_tet_mesh(len(v), t, balance=False, algorithm="mcs"))
        v, t, colors = _BLOB_CACHE[key]
        layer, j = divmod(i, per_layer)
        gx, gz = divmod(j, grid[1])
        pos = [(gx - (grid[0] - 1) / 2) * 2.1 + rng.uniform(-0.1, 0.1), 1.2 + 2.0 * layer + rng.uniform(0, 0.2), (gz - (grid[1] - 1) / 2) * 2.1 + rng.uniform(-0.1, 0.1)]
        obj = capi.default_object_params(miu=1e5, lmbd=1e6, density=50, frictionDynamic=0.2, frictionEpsV=0.01, translation=pos,
                                         rotation=[0, float(rng.uniform(0, 3.0)), 0], initialVelocity=[0, -2.0, 0], maxVelocityMagnitude=25.0,
                                         exponentialVelDamping=0.95, constantVelDamping=0.02)
        meshes.append((v, t, obj, colors, np.zeros(0, np.int32)))
    half = grid[0] * 2.1 / 2 + 0.3
    phys = capi.default_physics_params(numSubsteps=substeps, iterations=iterations, collisionStiffness=1e5, boundaryCollisionStiffness=1e5,
                                       boundaryFrictionDynamic=0.1, worldBoundsMin=[-half, 0, -half], worldBoundsMax=[half, 50, half])
    col = capi.default_collision_params(allowDCD=1, allowCCD=1)
    nV = sum(len(m[0]) for m in meshes); nT = sum(len(m[1]) for m in meshes)
    name = f"drop_pile {n_objects} tet models ({nT} tets, {nV} verts) into a box, DCD+CCD, {substeps} substeps x {iterations} iters, 8 colours"
    return dict(name=name, meshes=meshes, phys=phys, col=col, deformers=[], preroll=preroll)


def workload_cloth_drape(n=1000, width=100.0, substeps=20, iterations=15, preroll=2, sphere_level=4):
    """SURVEY.md §8(d) cfg 3: n x n vertex sheet (the reference's C{N}_horizontal.obj family, 100 units wide) centred above a
    sphere collider of radius 0.3 x width (all collider vertices fixed), S03_MultiLayerClothOnCollider.py values: StVK
    mu = lambda = 1e4, density 0.02, bendingStiffness 10, damping 2e-6, gravity (0,-1000,0), contactStiffness 1e5,
    contactRadius 0.3 and maxQueryDis 0.45 scaled by cell / cell(C70), thickness 1e-4, conservativeStepRelaxation 0.4,
    friction 0.1, 20 substeps x 15 iterations; VF + EE contact with the collider and with itself. `preroll` untimed frames
    let the sheet land on the sphere before anything is measured."""
    scale = (width / (n - 1)) / (100.0 / 69.0)
    R = 0.3 * width
    v, f = cloth.grid_cloth_fast(n, width, origin=(0.0, R + 0.6 * 0.45 * scale, 0.0))
    sv, sf = cloth.icosphere(R, sphere_level, center=(0.0, 0.0, 0.0))
    phys = cloth.default_cloth_physics_params(gravity=[0, -1000, 0], numSubsteps=substeps, iterations=iterations, contactStiffness=1e5,
                                              contactRadius=0.3 * scale, maxQueryDis=0.45 * scale, thickness=1e-4 * scale,
                                              conservativeStepRelaxation=0.4, contactEnergyTypes=1)
    obj = cloth.default_cloth_object_params(miu=1e4, density=0.02, bendingStiffness=10.0, dampingStVK=2e-6, dampingBending=2e-6,
                                            frictionDynamic=0.1, frictionEpsV=0.1, initializationType=2)
    obj.lambda_ = 1e4
    name = (f"cloth_drape {n}x{n} sheet ({len(v)} verts, {len(f)} tris) over a sphere collider ({len(sv)} verts), VF+EE contact, "
            f"{substeps} substeps x {iterations} iters, 9 colours")
    meshes = [(v, f, obj, cloth.grid_cloth_coloring(n), np.zeros(0, np.int32), True),
              (sv, sf, cloth.default_cloth_object_params(frictionDynamic=0.1), [], np.zeros(0, np.int32), False)]
    return dict(kind="cloth", name=name, meshes=meshes, phys=phys, col=None, deformers=[], preroll=preroll)


def workload_folded_beam(nx=16, ny=2, nz=2, h=0.2, gap=0.03, press=0.6, substeps=2, iterations=5):
    """One beam bent into a U (the bent shape IS the rest shape), lower arm on the ground and clamped at its free end, the
    upper arm's end pressed down onto the lower arm by a translator: self-contact between the two ENDS of one mesh. With
    `owner` (slabs along the unbent beam) every one of those contacts joins vertices of the first and the last rank - the
    case a one-ring ghost layer cannot see (cross-partition contacts)."""
    from . import meshio
    flat, tets = meshio.kuhn_beam(nx, ny, nz, h=h)
    thick, L = ny * h, nx * h
    R = 0.5 * (thick + gap)                         # radius of the centre line in the bend
    arc = np.pi * R
    arm = 0.5 * (L - arc)
    s_, off, z = flat[:, 0].astype(np.float64), flat[:, 1].astype(np.float64) - 0.5 * thick, flat[:, 2].astype(np.float64)
    x = np.empty_like(s_); y = np.empty_like(s_)
    lo, hi = s_ <= arm, s_ >= arm + arc
    mid = ~(lo | hi)
    y0 = 0.5 * thick + 0.02
    x[lo], y[lo] = s_[lo], y0 - off[lo]             # lower arm runs +x, its "up" is -off so that the upper arm keeps orientation
    th = (s_[mid] - arm) / R                        # 0..pi around the centre (arm, y0 + R)
    x[mid] = arm + (R + off[mid]) * np.sin(th)
    y[mid] = y0 + R - (R + off[mid]) * np.cos(th)
    x[hi], y[hi] = arm - (s_[hi] - arm - arc), y0 + 2 * R + off[hi]
    verts = np.stack([x, y, z], 1).astype(np.float32)
    p = verts[tets].astype(np.float64)
    det = np.einsum("ij,ij->i", p[:, 1] - p[:, 0], np.cross(p[:, 2] - p[:, 0], p[:, 3] - p[:, 0]))
    tets = tets.copy()
    tets[det < 0] = tets[det < 0][:, [0, 2, 1, 3]]
    colors = meshio.kuhn_coloring(nx, ny, nz)
    fixed = np.nonzero(flat[:, 0] < 1e-6)[0].astype(np.int32)
    tip = np.nonzero(flat[:, 0] > L - 1e-6)[0].astype(np.int32)
    from . import capi
    phys = capi.default_physics_params(numSubsteps=substeps, iterations=iterations, worldBoundsMin=[-10, 0, -10], worldBoundsMax=[10, 20, 10],
                                       collisionStiffness=1e5)
    col = capi.default_collision_params(allowDCD=1, allowCCD=1)
    obj = capi.default_object_params(miu=5e4, lmbd=5e5, density=100, dampingHydrostatic=1e-5, dampingDeviatoric=1e-5, frictionDynamic=0.2)
    return dict(name=f"folded_beam {nx}x{ny}x{nz} cells ({len(tets)} tets, {len(verts)} verts), DCD+CCD, self-contact between the two ends",
                meshes=[(verts, tets, obj, colors, fixed)], phys=phys, col=col, deformers=[("translator", tip, [0.0, -press, 0.0], -1.0)],
                flat=flat)


def is_cloth(wl):
    return wl.get("kind") == "cloth"


def build_scene(cls, wl, **kw):
    if is_cloth(wl):
        s = cls(wl["phys"], **kw)
        for v, f, obj, colors, fixed, simulated in wl["meshes"]:
            s.add_tri_mesh(v, f, obj, colors, fixed=fixed, simulated=simulated)
        s.finalize()
        return s
    s = cls(wl["phys"], wl["col"], **kw)
    if cls is capi.Scene and wl.get("refit_margin") is not None:
        s.set_lazy_refit_margin(wl["refit_margin"])
    if cls is capi.Scene and len(wl["meshes"]) > 1:      # many-object scenes: the rule the sharded builder applies too
        from gaia_b200.multigpu import scene_solver_lanes
        s.set_solver_lanes(scene_solver_lanes([m[3] for m in wl["meshes"]]))
    for v, t, obj, colors, fixed in wl["meshes"]:
        s.add_tet_mesh(v, t, obj, colors, fixed)
    for d in wl["deformers"]:
        if d[0] == "rotator":
            s.add_rotator(0, d[1], d[2], d[3], d[4], d[5])
        else:
            s.add_translator(0, d[1], d[2], d[3])
    s.finalize()
    return s


WORKLOADS = {
    # SURVEY.md §8(d) cfg 4 as written: DCD on, CCD off, intermediateCollisionIterations 5 (the *_nodcd variants: detection off)
    "twist_beam_10m": lambda: workload_twist_beam(472, 60, 60, collisions=True),
    "twist_beam_10m_nodcd": lambda: workload_twist_beam(472, 60, 60),
    "twist_beam_1m": lambda: workload_twist_beam(59, 60, 60, collisions=True),
    "twist_beam_1m_nodcd": lambda: workload_twist_beam(59, 60, 60),
    "twist_beam_small": lambda: workload_twist_beam(24, 12, 12),
    "drop_pile_32": lambda: workload_drop_pile(32),
    "drop_pile_8": lambda: workload_drop_pile(8, grid=(2, 2), preroll=30),
    # SURVEY.md §8(d) cfg 5 shape: many replicated models on a jittered lattice, 2 substeps x 20 iterations, DCD + CCD
    "drop_pile_512": lambda: workload_drop_pile(512, grid=(8, 8), substeps=2, iterations=20, preroll=20),
    "drop_pile_4096": lambda: workload_drop_pile(4096, grid=(16, 16), substeps=2, iterations=20, preroll=60),
    "cloth_1m": lambda: workload_cloth_drape(1000, preroll=12),
    "cloth_250k": lambda: workload_cloth_drape(500, width=50.0, sphere_level=4),
    "cloth_small": lambda: workload_cloth_drape(120, width=12.0, substeps=4, iterations=6, sphere_level=2),
}




# --------------------------------------------------------------------------------------------- reference-held meshes
REF_MESHES = {
    "teapot": "Data/mesh_models/UnitTest/Teapot/teapot.t",
    "bunny_small": "Data/mesh_models/t/bunny_small.t",
    "beam_1k": "Data/mesh_models/UnitTest/Beam/Beam_1k.t",
    "beam_lowres": "Data/mesh_models/UnitTest/Beam/Beam_lowRes.t",
    "octopus_5k": "Data/mesh_models/ModelsDifferentRes/Octopus3_Inflated/Octopus_Inflated_5k_noIntersection.t",
    "armadillo_lowres": "Data/mesh_models/t/Armadilo_lowres.t",
}
_REF_CACHE = {}


def ref_mesh(name):
    """(verts, tets, colours) of a reference mesh with the vertex colouring the reference ships for it."""
    if name not in _REF_CACHE:
        p = os.path.join(REF_ROOT, REF_MESHES[name])
        v, t = meshio.load_t(p)
        _REF_CACHE[name] = (v, t, meshio.load_coloring(p + ".vertexColoring.json"))
    return _REF_CACHE[name]


def workload_model_drop(model="teapot", frames_hint=100, scale=(1.5, 1.5, 1.5), rotation=(-1.5707963, 0.0, 0.0), height=10.0):
    """SURVEY.md §8(d) cfg 1 (BASELINE configs[0]): ONE tet model of Data/mesh_models with its shipped colouring dropping
    onto the ground plane; NeoHookean mu 1e5 lambda 1e6 density 50 (S01 teapot values, S01 scale / rotation), drop height 10,
    initialVelocity (0,-10,0), gravity (0,-10,0), worldBounds [[-10,0,-10],[10,20,10]], timeStep 0.01666666, 4 substeps x 20
    iterations, DCD / CCD off, boundaryCollisionStiffness 1e5, boundaryFrictionDynamic 0.1."""
    v, t, colors = ref_mesh(model)
    phys = capi.default_physics_params(gravity=[0, -10, 0], numSubsteps=4, iterations=20, timeStep=0.01666666,
                                       boundaryCollisionStiffness=1e5, boundaryFrictionDynamic=0.1,
                                       worldBoundsMin=[-10, 0, -10], worldBoundsMax=[10, 20, 10])
    col = capi.default_collision_params(allowDCD=0, allowCCD=0)
    obj = capi.default_object_params(miu=1e5, lmbd=1e6, density=50, initialVelocity=[0, -10, 0], scale=list(scale), rotation=list(rotation),
                                     translation=[0.0, height, 0.0])
    name = (f"model_drop {model} ({len(t)} tets, {len(v)} verts, shipped colouring, {len(colors)} colours) onto the ground plane, "
            f"4 substeps x 20 iters, ground contact only")
    return dict(name=name, meshes=[(v, t, obj, colors, np.zeros(0, np.int32))], phys=phys, col=col, deformers=[])


def workload_s01(first_frame=None, n_models=None):
    """SURVEY.md §8(d) cfg 2 (BASELINE configs[1]): exactly the reference's checked-in S01_Experiment Models.json /
    Parameters.json (32 models = 10 teapots, 9 bunnies, 8 octopuses, 5 armadillos: 87 339 verts / 330 335 tets, 4 substeps x
    20 iterations, DCD + CCD, frameToAppear = 120 + 6 i). `first_frame`: frameToAppear is shifted so that model 0 appears
    at that frame (the reference scene idles for 120 frames first); `n_models` keeps the first n models only."""
    d = os.path.join(REF_ROOT, "S01")
    phys, col, meshes = meshio.load_scene_json(os.path.join(d, "Models.json"), os.path.join(d, "Parameters.json"), REF_ROOT)
    if n_models:
        meshes = meshes[:n_models]
    if first_frame is not None:
        shift = min(m[2].frameToAppear for m in meshes) - first_frame
        for m in meshes:
            m[2].frameToAppear -= shift
    last = max(m[2].frameToAppear for m in meshes)
    nV = sum(len(m[0]) for m in meshes); nT = sum(len(m[1]) for m in meshes)
    name = (f"S01_Experiment_HybridModelsDrop_sequentiallyAppear: {len(meshes)} tet models ({nT} tets, {nV} verts) dropping into a box, "
            f"DCD+CCD, {phys.numSubsteps} substeps x {phys.iterations} iters, 8 colour groups, timed after frame {last + 60} (all models in the pile)")
    return dict(name=name, meshes=meshes, phys=phys, col=col, deformers=[], preroll=last + 60)


def workload_ref_stack(models=("teapot", "bunny_small", "octopus_5k", "armadillo_lowres"), substeps=4, iterations=20, gap=-0.05):
    """Config 2 in miniature: the four S01 meshes (S01 scales, shipped colourings, S01 material / contact parameters) stacked
    with a small initial overlap above the ground plane, so DCD + CCD contacts exist from the first frames."""
    s01 = {"teapot": ((1.5, 1.5, 1.5), (-1.5707963, 0.0, 0.0)), "bunny_small": ((0.2, 0.2, 0.2), (0.0, 0.0, 0.0)),
           "octopus_5k": ((0.8, 0.8, 0.8), (0.0, 0.0, 0.0)), "armadillo_lowres": ((3.0, 3.0, 3.0), (0.0, 0.0, 0.0)),
           "beam_1k": ((0.5, 0.5, 0.5), (0.0, 0.0, 0.3))}
    meshes, y = [], 0.05
    for i, mname in enumerate(models):
        v, t, colors = ref_mesh(mname)
        sc, rot = s01[mname]
        o = capi.default_object_params(scale=list(sc), rotation=list(rot))
        w = meshio.place(v, o)
        lo, hi = w.min(0), w.max(0)
        obj = capi.default_object_params(miu=1e5, lmbd=1e6, density=50, frictionDynamic=0.2, frictionEpsV=0.01, scale=list(sc), rotation=list(rot),
                                         translation=[float(-0.5 * (lo[0] + hi[0]) + 0.07 * i), float(y - lo[1]), float(-0.5 * (lo[2] + hi[2]) - 0.05 * i)],
                                         initialVelocity=[0.0, -1.0 * i, 0.0], maxVelocityMagnitude=25.0, exponentialVelDamping=0.9, constantVelDamping=0.1)
        meshes.append((v, t, obj, colors, np.zeros(0, np.int32)))
        y += float(hi[1] - lo[1]) + gap
    phys = capi.default_physics_params(gravity=[0, -10, 0], numSubsteps=substeps, iterations=iterations, timeStep=0.01666666, collisionStiffness=1e5,
                                       boundaryCollisionStiffness=1e5, boundaryFrictionDynamic=0.1, worldBoundsMin=[-10, 0, -10], worldBoundsMax=[10, 20, 10])
    col = capi.default_collision_params(allowDCD=1, allowCCD=1)
    nV = sum(len(m[0]) for m in meshes); nT = sum(len(m[1]) for m in meshes)
    return dict(name=f"ref_stack {'+'.join(models)} ({nT} tets, {nV} verts), DCD+CCD, {substeps} substeps x {iterations} iters",
                meshes=meshes, phys=phys, col=col, deformers=[])


WORKLOADS.update({
    "model_drop_teapot": lambda: workload_model_drop("teapot"),
    "s01": lambda: workload_s01(first_frame=1),
    "ref_stack4": lambda: workload_ref_stack(),
})
