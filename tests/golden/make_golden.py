#!/usr/bin/env python
"""Generate tests/golden/ref_vectors.npz from the REFERENCE's own code.

The vectors are outputs of oracle/_ref/libgaia_ref.so, i.e. of the reference's __host__ __device__ VBD math
(Simulator/Modules/VBD/VBDPhysicsCompute.cu, VBD_GeneralCompute.h, 3rdParty/CuMatrix) and of its own CCD solver
(Simulator/Modules/CollisionDetector/CCDSolver.h + 3rdParty/CyCodeBase), compiled from /root/reference by
oracle/Makefile. The reference ships no golden vectors of its own; these pin the CPU oracle (and through
it the CUDA path) to the reference's arithmetic on seeded inputs. Run in the build container only:

    python tests/golden/make_golden.py

Inputs are regenerated from the seeds by the tests, so only outputs + inputs needed for replay are stored.
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from golden_inputs import (cloth_group_merge_case, group_merge_case, collider_sequence_cases, cloth_deformer_cases, run_cloth_oracle_deformer, deformer_cases, run_oracle_deformer, inertia_cases, vertex_step_tail_cases, prepared_collision_cases, prepared_collision_inputs, step_schedule_case, cloth_bounds_case, bowl_cases, cloth_constraint_case, boundary_cases_cpu, cloth_step_sequence_cases, run_cloth_oracle_step_sequence, step_sequence_cases, run_oracle_step_sequence, cloth_feasible_region_cases, feasible_region_cases, traverse_cases, cloth_contact_case, cloth_material_case, neohookean_case, single_tet_cases, solve3x3_cases, ccd_cases, cubic_cases, contact_cases, boundary_cases,
                           point_in_tet_cases, closest_point_triangle_cases, segment_cases, cloth_repulsion_cases)  # noqa: E402
from oracle.binding import load_ref, ref_available  # noqa: E402

FP = C.POINTER(C.c_float)
IP = C.POINTER(C.c_int32)


def fp(a):
    return a.ctypes.data_as(FP)


def ip(a):
    return a.ctypes.data_as(IP)


def main():
    if not ref_available():
        raise SystemExit("oracle/_ref/libgaia_ref.so missing: run `make -C oracle` where /root/reference exists")
    ref = load_ref()
    out = {}
    # ---- Neo-Hookean per-vertex force / Hessian (GAIA::accumulateMaterialForceAndHessianForVertex_NeoHookean)
    cases = [("nh_plain", neohookean_case(seed=11, dampH=0.0, dampD=0.0))] + \
            [(f"nh_damped{i}", c) for i, c in enumerate(single_tet_cases())]
    for tag, c in cases:
        nV, nT = len(c["pos"]), len(c["tets"])
        F = np.zeros((nV, 3), np.float32); H = np.zeros((nV, 9), np.float32)
        for v in range(nV):
            f = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
            ref.ref_neohookean_vertex(nV, nT, fp(c["pos"]), fp(c["prev"]), ip(c["tets"]), fp(c["DmInv"]), fp(c["restVol"]),
                                      ip(c["infos"]), ip(c["nbrTets"]), ip(c["nbrOrder"]), c["miu"], c["lmbd"], c["dampH"], c["dampD"],
                                      c["dt"], v, fp(f), fp(h))
            F[v] = f; H[v] = h
        out[tag + "_force"] = F; out[tag + "_hessian"] = H
    # ---- per-tet Neo-Hookean energy (GAIA::evaluateNeoHookeanEnergy, minus its rest value): what the line search's merit energy sums
    c = neohookean_case(seed=11, dampH=0.0, dampD=0.0)
    nT = len(c["tets"])
    E = np.zeros(nT, np.float32)
    pos = np.ascontiguousarray(c["pos"]); dm = np.ascontiguousarray(c["DmInv"]); tets = np.ascontiguousarray(c["tets"], np.int32)
    for t in range(nT):
        E[t] = ref.ref_neohookean_tet_energy(c["miu"], c["lmbd"], float(c["restVol"][t]), t, ip(tets[t]), fp(pos), fp(dm))
    out["nh_tet_energy"] = E
    # ---- cloth material: VBDTriMeshStVk::intializeEdgeQuadraticForms + ::accumlateMaterialForceAndHessian (membrane + bending)
    c = cloth_material_case()
    nV, nF, nE = len(c["pos"]), len(c["faces"]), len(c["edges"])
    Q = np.zeros((nE, 16), np.float32)
    ref.ref_cloth_bending_forms(nV, fp(c["rest"]), nF, ip(c["faces"]), nE, ip(c["edges"]), fp(Q))
    out["cloth_Q"] = Q
    for which, tag in ((0, "cloth_material"), (1, "cloth_stvk"), (2, "cloth_bending")):
        F = np.zeros((nV, 3), np.float32); H = np.zeros((nV, 9), np.float32)
        for v in range(nV):
            f = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
            ref.ref_cloth_material_vertex(nV, fp(c["pos"]), fp(c["prev"]), nF, ip(c["faces"]), fp(c["dminv"]), fp(c["area"]),
                                          ip(c["vf"][0]), ip(c["vf"][1]), ip(c["vf"][2]), nE, ip(c["edges"]), fp(Q),
                                          ip(c["vb"][0]), ip(c["vb"][1]), ip(c["vb"][2]), c["miu"], c["lmbd"], c["dampS"], c["bend"], c["dampB"],
                                          c["dt"], c["thres"], v, which, fp(f), fp(h))
            F[v] = f; H[v] = h
        out[tag + "_force"] = F; out[tag + "_hessian"] = H
    # ---- cloth contacts, boundary and friction: VBDClothSimulationFramework::accumulateVF/EEContactForceAndHessian,
    #      ::accumlateBoundaryForceAndHessian (with ::computeFriction, ::computeContactRepulsiveForce), three contact energies
    for et in (0, 1, 2):
        c = cloth_contact_case(et)
        M = c["meshes"]
        VF = np.zeros((len(c["vf"]), 4, 13), np.float32)
        for i, (a, v, b, f) in enumerate(c["vf"]):
            tri = M[b]["pos"][M[b]["faces"][f]]
            cp = np.zeros(3, np.float32); bary = np.zeros(3, np.float32)
            ref.ref_closest_point_triangle(fp(M[a]["pos"][v]), fp(tri[0]), fp(tri[1]), fp(tri[2]), fp(cp), fp(bary))
            fr4 = np.array([M[a]["mu"], M[a]["eps"], M[b]["mu"], M[b]["eps"]], np.float32)
            for order in range(4):
                fo = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
                d = ref.ref_cloth_vf_contact(fp(c["params8"]), fp(fr4), fp(M[a]["pos"]), fp(M[a]["prev"]), v, fp(M[b]["pos"]), fp(M[b]["prev"]),
                                             ip(M[b]["faces"]), f, fp(cp), fp(bary), order, 1, fp(fo), fp(h))
                VF[i, order] = np.concatenate([[d], fo, h])
        out[f"cloth_vf_e{et}"] = VF
        EE = np.zeros((len(c["ee"]), 4, 13), np.float32)
        for i, (a, e1, b, e2) in enumerate(c["ee"]):
            E1, E2 = M[a]["edges"][e1], M[b]["edges"][e2]
            o8 = np.zeros(8, np.float32)
            ref.ref_closest_points_segments(fp(M[a]["pos"][E1[0]]), fp(M[a]["pos"][E1[1]]), fp(M[b]["pos"][E2[0]]), fp(M[b]["pos"][E2[1]]), fp(o8))
            fr4 = np.array([M[a]["mu"], M[a]["eps"], M[b]["mu"], M[b]["eps"]], np.float32)
            c1 = np.ascontiguousarray(o8[0:3]); c2 = np.ascontiguousarray(o8[3:6])
            for order in range(4):
                fo = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
                d = ref.ref_cloth_ee_contact(fp(c["params8"]), fp(fr4), fp(M[a]["pos"]), fp(M[a]["prev"]), len(M[a]["edges"]), ip(M[a]["edges"]), e1,
                                             fp(M[b]["pos"]), fp(M[b]["prev"]), len(M[b]["edges"]), ip(M[b]["edges"]), e2, fp(c1), fp(c2),
                                             float(o8[6]), float(o8[7]), order, 1, fp(fo), fp(h))
                EE[i, order] = np.concatenate([[d], fo, h])
        out[f"cloth_ee_e{et}"] = EE
        if et == 1:
            B = []
            for mi, m in enumerate(M):
                for v in range(len(m["pos"])):
                    fo = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
                    ref.ref_cloth_boundary(fp(c["params8"]), fp(c["bmin"]), fp(c["bmax"]), fp(m["pos"]), fp(m["prev"]), v, fp(fo), fp(h))
                    B.append(np.concatenate([fo, h]))
            out["cloth_boundary"] = np.asarray(B, np.float32)
    # ---- TetMeshFEM::tetrahedralTraverseTo (static-stack variant): reached?, tets traversed, stop reason
    for c in traverse_cases():
        R = np.zeros((len(c["rays"]), 3), np.int32)
        for i, (org, d, md, st, sf, tg) in enumerate(c["rays"]):
            n = np.zeros(1, np.int32); why = np.zeros(1, np.int32)
            ok = ref.ref_tet_traverse(len(c["pos"]), fp(c["pos"]), len(c["tets"]), ip(c["tets"]), ip(c["nbr"]), ip(c["xor"]), fp(org), fp(d), float(md),
                                      st, sf, tg, c["eps"], ip(n), ip(why))
            R[i] = (ok, n[0], why[0])
        out["traverse_" + c["name"]] = R
    # ---- DiscreteCollisionDetector::checkFeasibleRegion (+ checkEdge / checkVertexFeasibleRegion): closest-point type, verdict
    for c in feasible_region_cases():
        R = np.zeros((len(c["pts"]), 2), np.int32)
        for i, (f, q) in enumerate(c["pts"]):
            tri = c["pos"][c["faces"][f]]
            cp = np.zeros(3, np.float32); bary = np.zeros(3, np.float32)
            typ = ref.ref_closest_point_triangle(fp(q), fp(tri[0]), fp(tri[1]), fp(tri[2]), fp(cp), fp(bary))
            ok = ref.ref_check_feasible_region(len(c["pos"]), fp(c["pos"]), len(c["faces"]), ip(c["faces"]), ip(c["f3"]), ip(c["t2s"]), len(c["off"]) - 1,
                                               ip(c["off"]), ip(c["nb"]), fp(q), f, typ, c["eps"])
            R[i] = (typ, ok)
        out["feasible_" + c["name"]] = R
    # ---- cloth: checkFeasibleRegion (TriMeshFEM) + checkEdge / checkVertexFeasibleRegion
    c = cloth_feasible_region_cases()
    R = np.zeros((len(c["pts"]), 2), np.int32)
    for i, (f, q) in enumerate(c["pts"]):
        tri = c["pos"][c["faces"][f]]
        cp = np.zeros(3, np.float32); bary = np.zeros(3, np.float32)
        typ = ref.ref_closest_point_triangle(fp(q), fp(tri[0]), fp(tri[1]), fp(tri[2]), fp(cp), fp(bary))
        ok = ref.ref_cloth_check_feasible_region(len(c["pos"]), fp(c["pos"]), len(c["faces"]), ip(c["faces"]), ip(c["f3"]), len(c["edges"]), ip(c["edges"]),
                                                 ip(c["off"]), ip(c["nb"]), fp(q), f, typ, c["eps"])
        R[i] = (typ, ok)
    out["cloth_feasible"] = R
    # ---- VBDBaseTetMesh::applyInitialStep (computeInertia + forwardStep) and VBDPhysics::updateVelocities (+ dampVelocities)
    c = step_sequence_cases()
    for init_type, kw in c["cases"]:
        o, params12 = c["make"](init_type, kw)
        forces = []
        solved = np.ascontiguousarray(run_oracle_step_sequence(o, c, forces)[3])          # the "solved" positions build on the oracle's initial guesses
        n = len(c["pos0"])
        a = np.zeros((c["n_steps"], n, 3), np.float32); b = np.zeros_like(a); v = np.zeros_like(a)
        # type 5 (forwardStepSimplecticEuler) reads the elastic forces (the oracle's: its force routine is pinned separately) and 1 / mass
        fo = np.ascontiguousarray(np.array(forces, np.float32)); inv = np.ascontiguousarray((np.float32(1.0) / o.vertex_masses(0)).astype(np.float32))
        ref.ref_step_sequence(n, fp(c["pos0"]), fp(c["vel0"]), fp(params12), len(c["fixed"]), ip(c["fixed"]), c["n_steps"], fp(solved), fp(a), fp(b), fp(v),
                              fp(fo) if init_type == 5 else None, fp(inv) if init_type == 5 else None)
        out[f"step_init{init_type}"] = np.stack([a, b, v]); out[f"step_solved{init_type}"] = solved
        if init_type == 5:
            out["step_forces5"] = fo
    # ---- cloth: VBDBaseTriMesh::applyInitialStep (forwardStep_intertia), VBDClothSimulationFramework::updateVelocity, ::applyConservativeBoundTruncation
    c = cloth_step_sequence_cases()
    for init_type, kw in c["cases"]:
        o, params12 = c["make"](init_type, kw)
        solved = np.ascontiguousarray(run_cloth_oracle_step_sequence(o, c)[3])
        n = len(c["pos0"])
        a = np.zeros((c["n_steps"], n, 3), np.float32); b = np.zeros_like(a); v = np.zeros_like(a)
        ref.ref_cloth_step_sequence(n, fp(c["pos0"]), fp(c["vel0"]), fp(params12), len(c["fixed"]), ip(c["fixed"]), fp(c["bounds"]), c["n_steps"], fp(solved),
                                    fp(a), fp(b), fp(v))
        out[f"cloth_step_init{init_type}"] = np.stack([a, b, v]); out[f"cloth_step_solved{init_type}"] = solved
    tp = np.ascontiguousarray(c["trunc_pos"].copy()); tf = np.zeros(len(tp), np.int32)
    ref.ref_cloth_truncate(len(c["pos0"]), fp(c["pos0"]), fp(c["bounds"]), len(tp), ip(c["trunc_verts"]), fp(tp), ip(tf))
    out["cloth_truncate_pos"] = tp; out["cloth_truncate_flag"] = tf
    # ---- VBDPhysics::accumlateBoundaryForceAndHessian, CPU semantics (friction per violated axis), with and without friction
    bc = boundary_cases_cpu()
    B = np.zeros((2, len(bc["p"]), 12), np.float32)
    for fr in (0, 1):
        for i in range(len(bc["p"])):
            f = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
            ref.ref_boundary_cpu(fp(bc["p"][i]), fp(bc["pPrev"][i]), fp(bc["mn"]), fp(bc["mx"]), bc["kB"], bc["muB"], bc["epsVB"], bc["dt"], fr, fp(f), fp(h))
            B[fr, i] = np.concatenate([f, h])
    out["boundary_cpu"] = B
    # ---- the same function with useBowlGround: [cap][friction][case][f3, h9]
    bw = bowl_cases()
    BW = np.zeros((2, 2, len(bw["p"]), 12), np.float32)
    for cap in (0, 1):
        for fr in (0, 1):
            for i in range(len(bw["p"])):
                f = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
                ref.ref_boundary_bowl_cpu(fp(bw["p"][i]), fp(bw["pPrev"][i]), fp(bw["center"]), bw["radius"], cap, bw["kB"], bw["muB"], bw["epsVB"], bw["dt"], fr, fp(f), fp(h))
                BW[cap, fr, i] = np.concatenate([f, h])
    out["boundary_bowl_cpu"] = BW
    # ---- VBDPhysics::updateCollisionInfo + ::accumlateCollisionForceAndHessianPerCollision (the CPU contact term): [case][order][depth, f3, h9]
    cc = contact_cases()
    CC = np.zeros((len(cc["p"]), 4, 13), np.float32)
    for i in range(len(cc["p"])):
        for order in range(4):
            f = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
            d = ref.ref_contact_cpu(fp(cc["p"][i]), fp(cc["pPrev"][i]), fp(cc["tri"][i]), fp(cc["triPrev"][i]), order, cc["k"], cc["mu0"], cc["mu1"],
                                    cc["eps0"], cc["eps1"], cc["dt"], fp(f), fp(h))
            CC[i, order] = np.concatenate([[d], f, h])
    out["contact_cpu"] = CC
    # ---- cloth sewing and attachments: VBDBaseTriMesh::accumlateSewingForceAndHessian, the attachment set-up loop of
    #      VBDClothSimulationFramework::initialize and ::accumulateConstraintForceAndHessian
    c = cloth_constraint_case()
    pos0 = np.ascontiguousarray(c["meshes"][0][0])
    SW = np.zeros((len(pos0), 12), np.float32)
    for v in range(len(pos0)):
        f = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
        ref.ref_cloth_sewing(len(pos0), fp(pos0), len(c["ratios"]), ip(c["seams"]), fp(c["ratios"]), c["sew_k"], v, fp(f), fp(h))
        SW[v] = np.concatenate([f, h])
    out["cloth_sewing"] = SW
    allpos = np.ascontiguousarray(np.concatenate([m[0] for m in c["meshes"]]))
    voff = np.zeros(len(c["meshes"]) + 1, np.int32); voff[1:] = np.cumsum([len(m[0]) for m in c["meshes"]])
    n_sim_v = int(voff[c["n_sim"]])
    for tag, st in (("", c["stiff"]), ("_default_k", None)):
        A = np.zeros((n_sim_v, 12), np.float32); cnt = np.zeros(n_sim_v, np.int32)
        ref.ref_cloth_attachments(len(c["meshes"]), c["n_sim"], ip(voff), fp(allpos), len(c["v_infos"]), ip(c["v_infos"]), ip(c["f_infos"]), fp(c["barys"]),
                                  fp(st) if st is not None else None, fp(A), ip(cnt))
        out["cloth_attach" + tag] = A; out["cloth_attach_count" + tag] = cnt
    # ---- the tail of VBDPhysics::VBDStepWithCollision (threshold, CuMatrix::solve3x3_psd_stable, gradient fall-back, update)
    c = vertex_step_tail_cases()
    xs = c["x"].copy(); order = np.zeros(4, np.int32)
    ref.ref_vertex_step_tail(len(xs), fp(c["f"]), fp(c["h"]), fp(xs), c["step"], c["step_gd"], 0, ip(order))
    out["vertex_step_tail_x"] = xs; out["vertex_step_order"] = order
    # ---- the Chebyshev weights: VBDPhysics::getAcceleratorOmega (tet step) and ChebysevAccelerator (cloth step), 40 iterations per rho
    rhos = np.array([0.0, 0.3, 0.8, 0.93, 0.95, 0.99], np.float32)
    W = np.zeros((2, len(rhos), 40), np.float32)
    for which in (0, 1):
        for i, rho in enumerate(rhos):
            ref.ref_accelerator_omegas(which, float(rho), 40, fp(W[which, i]))
    out["accel_rhos"] = rhos; out["accel_omegas"] = W
    rng = np.random.default_rng(141)
    yv = rng.uniform(-3, 3, (300, 3)).astype(np.float32); ypp = (yv + rng.normal(0, 0.01, yv.shape)).astype(np.float32)
    acc = yv.copy()
    ref.ref_accelerate(len(acc), fp(acc), fp(ypp), 1.7321)
    out["accel_y"] = yv; out["accel_ypp"] = ypp; out["accel_out"] = acc
    # ---- DeformerRotator / DeformerTranslater at the oracle's own curTime after each frame
    c = deformer_cases()
    for i, case in enumerate(c["cases"]):
        times, _, o = run_oracle_deformer(case, c)
        kind, sel, centre, a, w, end = case
        res = np.zeros((len(times), len(sel), 3), np.float32)
        ref.ref_deformer(kind, len(c["verts"]), fp(np.ascontiguousarray(c["verts"])), len(sel), ip(sel), fp(np.array(centre, np.float32)), fp(np.array(a, np.float32)),
                         w, end, len(times), times.ctypes.data_as(C.POINTER(C.c_double)), fp(res))
        out[f"deformer{i}_times"] = times; out[f"deformer{i}_pos"] = res
    # ---- ClothDeformerRotator / ClothDeformerTranslater, applied once per substep at the cloth oracle's own clock
    c = cloth_deformer_cases()
    for i, case in enumerate(c["cases"]):
        times, dt, _ = run_cloth_oracle_deformer(case, c)
        kind, sel, centre, a, w, t0, t1 = case
        res = np.zeros((len(times), len(sel), 3), np.float32)
        ref.ref_cloth_deformer(kind, len(c["verts"]), fp(np.ascontiguousarray(c["verts"])), len(sel), ip(sel), fp(np.array(centre, np.float32)), fp(np.array(a, np.float32)),
                               w, t0, t1, len(times), fp(times), dt, fp(res))
        out[f"cloth_deformer{i}_pos"] = res[3::4].copy()          # after the last substep of each frame
    # ---- ColliderTrimeshSequence::update / ::set_state
    c = collider_sequence_cases()
    nV = len(c["verts"])
    for i, (isub, iit) in enumerate(c["settings"]):
        res = np.zeros((len(c["calls"]), nV, 3), np.float32); upd = np.zeros(len(c["calls"]), np.int32)
        ref.ref_collider_sequence(nV, len(c["key_frames"]), ip(c["key_frames"]), fp(c["keys"]), isub, iit, len(c["calls"]), ip(c["calls"]), c["substeps"], c["iters"], fp(res), ip(upd))
        out[f"collider_seq{i}_pos"] = res; out[f"collider_seq{i}_upd"] = upd
    # ---- findSmallestParallelGroup + the merge loop of VBDPhysics::initialize, on colour classes sorted large first (ties in file order)
    meshes = group_merge_case()
    srt = [sorted(cols, key=lambda c: -len(c)) for _, _, cols in meshes]
    n_colors = np.array([len(c) for c in srt], np.int32); n_verts = np.array([len(v) for v, _, _ in meshes], np.int32)
    flat = [c for cols in srt for c in cols]
    class_off = np.zeros(len(flat) + 1, np.int32); class_off[1:] = np.cumsum([len(c) for c in flat])
    class_verts = np.ascontiguousarray(np.concatenate(flat), np.int32)
    g_off = np.zeros(int(n_colors.max()) + 1, np.int32); g_pairs = np.zeros(2 * int(n_verts.sum()), np.int32)
    nG = ref.ref_merge_groups(len(meshes), ip(n_verts), ip(n_colors), ip(class_off), ip(class_verts), ip(g_off), ip(g_pairs))
    out["group_merge_off"] = g_off[: nG + 1].copy(); out["group_merge_pairs"] = g_pairs[: g_off[nG]].reshape(-1, 2).copy()
    # ---- VBDClothSimulationFramework::initializeParallelGroups + ::findSmallestParallelGroup (classes in file order)
    meshes = cloth_group_merge_case()
    n_colors = np.array([len(c) for _, _, c in meshes], np.int32); n_verts = np.array([len(v) for v, _, _ in meshes], np.int32)
    # the mesh loader sorts the classes large to small before the merge sees them (TriMesh.cpp:747-750, needs the mesh class: done
    # here, stable on ties as the oracle does)
    flat = [c for _, _, cols in meshes for c in sorted(cols, key=lambda c: -len(c))]
    class_off = np.zeros(len(flat) + 1, np.int32); class_off[1:] = np.cumsum([len(c) for c in flat])
    class_verts = np.ascontiguousarray(np.concatenate(flat), np.int32)
    g_off = np.zeros(int(n_colors.max()) + 1, np.int32); g_pairs = np.zeros(2 * int(n_verts.sum()), np.int32); g_col = np.zeros(int(n_verts.sum()), np.int32)
    nG = ref.ref_cloth_merge_groups(len(meshes), ip(n_verts), ip(n_colors), ip(class_off), ip(class_verts), ip(g_off), ip(g_pairs), ip(g_col))
    out["cloth_group_merge_off"] = g_off[: nG + 1].copy(); out["cloth_group_merge_pairs"] = g_pairs[: g_off[nG]].reshape(-1, 2).copy()
    # ---- VBDBaseTetMesh::accumlateInertiaForceAndHessian
    c = inertia_cases()
    f = c["f0"].copy(); h = c["h0"].copy()
    ref.ref_inertia_terms(len(f), fp(c["mass"]), fp(c["inertia"]), fp(c["x"]), c["dtsq"], fp(f), fp(h))
    out["inertia_f"] = f; out["inertia_h"] = h
    # ---- the schedule of VBDPhysics::runStep_hybridCollisionHandling with its own dcd / ccd / intermediate passes (stand-in members and
    #      detectors that record their calls): three frames with both detectors, one frame with DCD alone, one with CCD alone
    up8 = lambda a: a.ctypes.data_as(C.POINTER(C.c_uint8))
    for tag, dcd, ccd, frames in (("", 1, 1, 3), ("_dcd", 1, 0, 1), ("_ccd", 0, 1, 1)):
        c = step_schedule_case(dcd, ccd)
        pen = np.zeros(c["n_verts"], np.uint8)
        for frame in range(frames):
            c["scene"].trace_begin(); c["scene"].run_frames(1)
            rec = np.ascontiguousarray(c["scene"].trace_read())
            args = lambda o_, cap: (ip(c["params"](frame)), 3, ip(c["active"][frame]), ip(c["fixed_off"]), up8(c["mask"]), c["n_groups"], ip(c["group_off"]), ip(c["pairs"]),
                                    ip(rec), len(rec), o_, cap, ip(c["sv_off"]), ip(c["svids"]))
            n = ref.ref_step_trace(*args(None, 0), up8(pen.copy()))
            tr = np.zeros(n, np.int32)
            ref.ref_step_trace(*args(ip(tr), n), up8(pen))          # (this call carries the penetrated mask on to the next frame)
            out[f"step_trace{tag}_f{frame}"] = tr.reshape(-1, 4)
    # ---- VBDPhysics::prepareCollisionDataCPU + ActiveCollisionList::addToActiveCollisionList on the oracle's detection results
    for tag, o in prepared_collision_cases():
        q = prepared_collision_inputs(o)
        nG = len(q["group_sizes"]); nVt = int(q["v_off"][-1])
        a_off = np.zeros(nG + 1, np.int32); a_pairs = np.zeros((4096, 2), np.int32); r_off = np.zeros(nVt + 1, np.int32); rows = np.zeros((8192, 5), np.int32)
        mask = np.zeros(nVt, np.uint8)
        nr = ref.ref_prepare_collision_data(q["nM"], ip(q["v_off"]), ip(q["sv_off"]), ip(q["f_off"]), ip(q["vertex_group"]), ip(q["svids"]), ip(q["faces"]),
                                            up8(q["has"]), up8(q["ccd"]), up8(q["path"]), ip(q["hit_mesh"]), ip(q["hit_face"]), nG, ip(q["group_sizes"]), 8.0,
                                            ip(a_off), ip(a_pairs), len(a_pairs), ip(r_off), ip(rows), len(rows), up8(mask))
        # activeCollisionListPreAllocationRatio 8 (the reference's default 1 overflows on these small groups: it then drops contacts
        # and, having counted before it checks - ActiveCollisionList.cpp:63-65 -, reads a slot it never wrote)
        assert nr >= 0 and (np.diff(a_off) <= 8 * q["group_sizes"]).all()
        out[f"prep_{tag}_aoff"] = a_off; out[f"prep_{tag}_apairs"] = a_pairs[: a_off[-1]].copy()
        out[f"prep_{tag}_roff"] = r_off; out[f"prep_{tag}_rows"] = rows[:nr].copy(); out[f"prep_{tag}_mask"] = mask
    # ---- the conservative-bounds loop of VBDClothSimulationFramework::applyCollisionDetection
    c = cloth_bounds_case()
    nV, nF, nE = c["sizes"]
    f_off, f_adj, e_off, e_adj = c["scene"].incidence(0)
    b = np.zeros(nV, np.float32); mk = np.zeros(nV, np.uint8)
    up = lambda a: a.ctypes.data_as(C.POINTER(C.c_uint8))
    ref.ref_cloth_bounds(nV, nF, nE, ip(f_off), ip(f_adj), ip(e_off), ip(e_adj), c["R"], c["thick"], c["relax"], fp(c["vf_min"]), ip(c["vf_cnt"]), fp(c["face_min"]),
                         ip(c["face_cnt"]), fp(c["ee_min"]), ip(c["ee_cnt"]), up(c["mask"]), fp(b), up(mk))
    out["cloth_bounds"] = b; out["cloth_bounds_mask"] = mk
    # ---- CuMatrix::solve3x3_psd_stable
    Ms, Bs = solve3x3_cases()
    X = np.zeros_like(Bs); ok = np.zeros(len(Ms), np.int32)
    for i in range(len(Ms)):
        x = np.zeros(3, np.float32)
        ok[i] = ref.ref_solve3x3(fp(Ms[i]), fp(Bs[i]), fp(x)); X[i] = x
    out["solve_x"] = X; out["solve_ok"] = ok
    # ---- cy::IntersectContinuousTriPoint<float>
    x0, x1, p0, p1 = ccd_cases()
    hit = np.zeros(len(x0), np.int32); tt = np.zeros(len(x0), np.float32); bary = np.zeros((len(x0), 3), np.float32)
    for i in range(len(x0)):
        t = C.c_float(-1); b = np.zeros(3, np.float32)
        hit[i] = ref.ref_ccd_tri_point(fp(x0[i]), fp(x1[i]), fp(p0[i]), fp(p1[i]), C.byref(t), fp(b))
        tt[i] = t.value; bary[i] = b
    out["ccd_hit"] = hit; out["ccd_t"] = tt; out["ccd_bary"] = bary
    # ---- cy::CubicRoots<float>
    coefs = cubic_cases()
    nr = np.zeros(len(coefs), np.int32); roots = np.zeros((len(coefs), 3), np.float32)
    for i in range(len(coefs)):
        r = np.zeros(3, np.float32)
        nr[i] = ref.ref_cubic_roots(fp(coefs[i]), 0.0, 1.0, fp(r)); roots[i] = r
    out["cubic_n"] = nr; out["cubic_roots"] = roots
    # ---- vertex-face contact force / Hessian (GAIA::accumlateCollisionForceAndHessianPerCollision_v2)
    cc = contact_cases()
    Fc = np.zeros((len(cc["p"]), 4, 3), np.float32); Hc = np.zeros((len(cc["p"]), 4, 9), np.float32)
    for i in range(len(cc["p"])):
        for order in range(4):
            f = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
            ref.ref_collision_per_collision(fp(cc["p"][i]), fp(cc["pPrev"][i]), fp(cc["tri"][i]), fp(cc["triPrev"][i]), order,
                                            cc["k"], cc["mu0"], cc["mu1"], cc["eps0"], cc["eps1"], cc["dt"], fp(f), fp(h))
            Fc[i, order] = f; Hc[i, order] = h
    out["contact_force"] = Fc; out["contact_hessian"] = Hc
    # ---- world-box penalty + friction (GAIA::accumulateBoundaryForceAndHessianGPU; one violated axis per case, where the
    #      reference's CPU "per axis" and GPU "once" friction rules coincide)
    bc = boundary_cases()
    Fb = np.zeros((len(bc["p"]), 3), np.float32); Hb = np.zeros((len(bc["p"]), 9), np.float32)
    bounds6 = np.ascontiguousarray(np.concatenate([bc["mn"], bc["mx"]]), np.float32)   # [min xyz, max xyz] (M/VBD/VBDPhysics.cpp:637-642)
    for i in range(len(bc["p"])):
        f = np.zeros(3, np.float32); h = np.zeros(9, np.float32)
        ref.ref_boundary(fp(bc["p"][i]), fp(bc["pPrev"][i]), fp(bounds6), bc["kB"], bc["muB"], bc["epsVB"], bc["dt"], fp(f), fp(h))
        Fb[i] = f; Hb[i] = h
    out["boundary_force"] = Fb; out["boundary_hessian"] = Hb
    # ---- collision predicates (GAIA::pointInTet, closestPointTriangle, get_closest_points_between_segments,
    #      M/CollisionDetector/CollisionGeometry.h:56-267, compiled with the vendored M/common Embree math)
    p, tets = point_in_tet_cases()
    out["pit_inside"] = np.array([ref.ref_point_in_tet(fp(p[i]), fp(tets[i, 0]), fp(tets[i, 1]), fp(tets[i, 2]), fp(tets[i, 3])) for i in range(len(p))], np.int32)
    p, tri = closest_point_triangle_cases()
    cpt = np.zeros((len(p), 3), np.float32); cpb = np.zeros((len(p), 3), np.float32); cpk = np.zeros(len(p), np.int32)
    for i in range(len(p)):
        c3 = np.zeros(3, np.float32); b3 = np.zeros(3, np.float32)
        cpk[i] = ref.ref_closest_point_triangle(fp(p[i]), fp(tri[i, 0]), fp(tri[i, 1]), fp(tri[i, 2]), fp(c3), fp(b3))
        cpt[i] = c3; cpb[i] = b3
    out["cpt_point"] = cpt; out["cpt_bary"] = cpb; out["cpt_type"] = cpk
    seg = segment_cases()
    so = np.zeros((len(seg), 8), np.float32)
    for i in range(len(seg)):
        o8 = np.zeros(8, np.float32)
        ref.ref_closest_points_segments(fp(seg[i, 0]), fp(seg[i, 1]), fp(seg[i, 2]), fp(seg[i, 3]), fp(o8))
        so[i] = o8
    out["seg_out"] = so
    # ---- cloth contact energy derivatives (VBDClothSimulationFramework::computeContactRepulsiveForce, VBDClothPhysics.cpp:1533-1579)
    dis, th, r, k = cloth_repulsion_cases()
    rep = np.zeros((3, len(dis), 2), np.float32)
    for et in range(3):
        for i in range(len(dis)):
            a, b = C.c_float(0), C.c_float(0)
            ref.ref_cloth_contact_repulsion(float(dis[i]), float(th[i]), float(r[i]), float(k[i]), et, C.byref(a), C.byref(b))
            rep[et, i] = (a.value, b.value)
    out["cloth_repulsion"] = rep
    path = os.path.join(ROOT, "tests", "golden", "ref_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
