// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// C entry points of the CPU cloth oracle; they shadow include/gaia_b200_cloth.h one to one with the prefix
// `oracle_cloth_` (handle = void*), plus a few probes for the unit tests.
#include "oracle_cloth.h"

using namespace orc;

extern "C" {

void* oracle_cloth_create(const gaia_cloth_physics_params* phys, int nThreads) {
    ClothScene* s = new ClothScene();
    s->phys = *phys;
    s->dt = s->phys.timeStep / s->phys.numSubsteps;          // BasePhysicFramework::timeStepEqualize
    s->dtSqrReciprocal = 1.f / (s->dt * s->dt);
    std::memset(&s->stats, 0, sizeof(s->stats));
    if (nThreads <= 0) nThreads = (int)std::thread::hardware_concurrency();
    s->pool = new Pool(nThreads);
    return s;
}
void oracle_cloth_destroy(void* h) { delete (ClothScene*)h; }
int oracle_cloth_num_threads(void* h) { return ((ClothScene*)h)->pool->size(); }
void oracle_cloth_set_brute_force(void* h, int on) { ((ClothScene*)h)->bruteForce = on != 0; }

int oracle_cloth_add_tri_mesh(void* h, int32_t simulated, int32_t nV, int32_t nF, const float* xyz, const int32_t* faces,
                              const gaia_cloth_object_params* obj, int32_t nColors, const int32_t* colorOffsets, const int32_t* colorVerts,
                              int32_t nFixed, const int32_t* fixed) {
    ClothScene* s = (ClothScene*)h;
    if (s->finalized) return GAIA_ERR_STATE;
    if (simulated && !s->meshes.empty() && !s->meshes.back().simulated) return GAIA_ERR_STATE;   // colliders come last
    s->meshes.emplace_back();
    TriMesh& m = s->meshes.back();
    m.id = (int)s->meshes.size() - 1;
    cloth_mesh_initialize(m, simulated != 0, nV, nF, xyz, faces, *obj, nColors, colorOffsets, colorVerts, nFixed, fixed);
    return m.id;
}
int oracle_cloth_add_rotator(void* h, int mesh, const int32_t* verts, int32_t n, const float* center, const float* axis,
                             double angularVelocity, double rotationEndTime) {
    ClothScene* s = (ClothScene*)h;
    if (s->finalized || mesh < 0 || mesh >= (int)s->meshes.size()) return GAIA_ERR_STATE;
    cloth_add_rotator(*s, mesh, verts, n, center, axis, angularVelocity, rotationEndTime);
    return (int)s->deformers.size() - 1;
}
int oracle_cloth_set_collider_sequence(void* h, int mesh, int32_t nKeys, const int32_t* keyFrames, const float* xyz, int32_t interpolateSubstep,
                                       int32_t interpolateIter) {
    ClothScene* s = (ClothScene*)h;
    if (s->finalized || mesh < 0 || mesh >= (int)s->meshes.size() || s->meshes[mesh].simulated) return GAIA_ERR_STATE;
    cloth_set_collider_sequence(*s, mesh, nKeys, keyFrames, xyz, interpolateSubstep != 0, interpolateIter);
    return GAIA_OK;
}
int oracle_cloth_add_translator(void* h, int mesh, const int32_t* verts, int32_t n, const float* speed, float tStart, float tEnd) {
    ClothScene* s = (ClothScene*)h;
    if (s->finalized || mesh < 0 || mesh >= (int)s->meshes.size()) return GAIA_ERR_STATE;
    cloth_add_translator(*s, mesh, verts, n, speed, tStart, tEnd);
    return (int)s->deformers.size() - 1;
}
int oracle_cloth_set_sewing(void* h, int mesh, int32_t nSeams, const int32_t* verts3, const float* ratios, float stiffness) {
    ClothScene* s = (ClothScene*)h;
    if (s->finalized) return GAIA_ERR_STATE;
    if (mesh < 0 || mesh >= (int)s->meshes.size() || !s->meshes[mesh].simulated || nSeams < 0 || (nSeams && (!verts3 || !ratios))) return GAIA_ERR_INVALID_ARG;
    for (int i = 0; i < 3 * nSeams; i++) if (verts3[i] < 0 || verts3[i] >= s->meshes[mesh].nV) return GAIA_ERR_INVALID_ARG;
    cloth_set_sewing(*s, mesh, nSeams, verts3, ratios, stiffness);
    return GAIA_OK;
}
int oracle_cloth_set_filtered_pairs(void* h, int mesh, int32_t nVF, const int32_t* vfPairs, int32_t nEE, const int32_t* eePairs) {
    ClothScene* s = (ClothScene*)h;
    if (s->finalized) return GAIA_ERR_STATE;
    if (mesh < 0 || mesh >= (int)s->meshes.size() || !s->meshes[mesh].simulated || nVF < 0 || nEE < 0 || (nVF && !vfPairs) || (nEE && !eePairs)) return GAIA_ERR_INVALID_ARG;
    const TriMesh& m = s->meshes[mesh];
    for (int i = 0; i < nVF; i++) if (vfPairs[2 * i] < 0 || vfPairs[2 * i] >= m.nV || vfPairs[2 * i + 1] < 0 || vfPairs[2 * i + 1] >= m.nF) return GAIA_ERR_INVALID_ARG;
    for (int i = 0; i < 2 * nEE; i++) if (eePairs[i] < 0 || eePairs[i] >= m.nE) return GAIA_ERR_INVALID_ARG;
    cloth_set_filtered_pairs(*s, mesh, nVF, vfPairs, nEE, eePairs);
    return GAIA_OK;
}
int oracle_cloth_add_attachment(void* h, int meshV, int32_t vertex, int meshF, const int32_t* faceVerts, const float* baryOrPos, float stiffness) {
    ClothScene* s = (ClothScene*)h;
    if (s->finalized) return GAIA_ERR_STATE;
    const int nM = (int)s->meshes.size();
    if (meshV < 0 || meshV >= nM || !s->meshes[meshV].simulated || vertex < 0 || vertex >= s->meshes[meshV].nV || meshF < -1 || meshF >= nM || !baryOrPos)
        return GAIA_ERR_INVALID_ARG;
    if (meshF >= 0) {
        if (!faceVerts) return GAIA_ERR_INVALID_ARG;
        for (int k = 0; k < 3; k++) {
            if (faceVerts[k] < 0 || faceVerts[k] >= s->meshes[meshF].nV) return GAIA_ERR_INVALID_ARG;
            if (s->meshes[meshF].simulated && baryOrPos[k] != 0.f && faceVerts[k] >= s->meshes[meshV].nV) return GAIA_ERR_INVALID_ARG;   // read from meshV (:1472)
        }
    }
    return cloth_add_attachment(*s, meshV, vertex, meshF, faceVerts, baryOrPos, stiffness);
}
// probes: the two terms alone, at the scene's current positions
void oracle_cloth_sewing_force_hessian(void* h, int mesh, int v, float* f3, float* h9) {
    ClothScene* s = (ClothScene*)h;
    V3 f; M3 H;
    cloth_sewing_force_hessian(s->meshes[mesh], v, f, H);
    f3[0] = f.x; f3[1] = f.y; f3[2] = f.z; std::memcpy(h9, H.m, 36);
}
void oracle_cloth_constraint_force_hessian(void* h, int mesh, int v, float* f3, float* h9) {
    ClothScene* s = (ClothScene*)h;
    V3 f; M3 H;
    cloth_constraint_force_hessian(*s, mesh, v, f, H);
    f3[0] = f.x; f3[1] = f.y; f3[2] = f.z; std::memcpy(h9, H.m, 36);
}
// probe: the conservative-bounds fold alone, from given query leftovers (smallest distance and number of contacts per vertex / face /
// edge of simulated `mesh`) and a given mask; also hands out the vertex -> faces / edges incidence it folds over (CSR; NULL to skip)
int oracle_cloth_probe_bounds(void* h, int mesh, const float* vfMinDis, const int32_t* vfCount, const float* faceMinDis, const int32_t* faceCount,
                              const float* eeMinDis, const int32_t* eeCount, const uint8_t* maskBefore, float* bounds, uint8_t* maskAfter) {
    ClothScene* s = (ClothScene*)h;
    if (mesh < 0 || mesh >= s->numSim) return GAIA_ERR_INVALID_ARG;
    TriMesh& m = s->meshes[mesh];
    for (int v = 0; v < m.nV; v++) { m.vfMinDis[v] = vfMinDis[v]; m.vf[v].assign((size_t)vfCount[v], VFContact()); m.activeCollisionMask[v] = maskBefore[v]; }
    for (int f = 0; f < m.nF; f++) { m.faceMinDis[f] = faceMinDis[f]; m.faceLinks[f].assign((size_t)faceCount[f], FVLink()); }
    for (int e = 0; e < m.nE; e++) { m.eeMinDis[e] = eeMinDis[e]; m.ee[e].assign((size_t)eeCount[e], EEContact()); }
    cloth_conservative_bounds(*s, m);
    for (int v = 0; v < m.nV; v++) { bounds[v] = m.bounds[v]; maskAfter[v] = m.activeCollisionMask[v]; }
    return GAIA_OK;
}
int oracle_cloth_get_incidence(void* h, int mesh, int32_t* fOff, int32_t* fAdj, int32_t* eOff, int32_t* eAdj) {
    const TriMesh& m = ((ClothScene*)h)->meshes[mesh];
    for (int v = 0; v <= m.nV; v++) { fOff[v] = m.vNbrFOff[v]; eOff[v] = m.vNbrEOff[v]; }
    for (size_t i = 0; i < m.vNbrF.size(); i++) fAdj[i] = m.vNbrF[i];
    for (size_t i = 0; i < m.vNbrE.size(); i++) eAdj[i] = m.vNbrE[i];
    return GAIA_OK;
}
void oracle_cloth_accelerator_omegas(float rho, int n, float* out) { float w = 1.f; for (int k = 1; k <= n; k++) { w = cloth_accelerator_omega(rho, k, w); out[k - 1] = w; } }
int oracle_cloth_probe_update_colliders(void* h, int frameId, int substep, int iIter) { return cloth_probe_update_colliders(*(ClothScene*)h, frameId, substep, iIter) ? 1 : 0; }
int oracle_cloth_finalize(void* h) { cloth_finalize(*(ClothScene*)h); return GAIA_OK; }
int oracle_cloth_run_frames(void* h, int32_t n) {
    ClothScene* s = (ClothScene*)h;
    for (int i = 0; i < n; i++) cloth_run_substeps(*s, s->phys.numSubsteps);
    return GAIA_OK;
}
int oracle_cloth_run_iterations(void* h, int32_t n) { cloth_run_iterations(*(ClothScene*)h, n); return GAIA_OK; }
int oracle_cloth_synchronize(void*) { return GAIA_OK; }
int oracle_cloth_frame_id(void* h) { return ((ClothScene*)h)->frameId; }
double oracle_cloth_cur_time(void* h) { return (double)((ClothScene*)h)->curTime; }
int oracle_cloth_set_time(void* h, int32_t frameId, double curTime) { ((ClothScene*)h)->frameId = frameId; ((ClothScene*)h)->curTime = (float)curTime; return GAIA_OK; }
int oracle_cloth_num_meshes(void* h) { return (int)((ClothScene*)h)->meshes.size(); }
int oracle_cloth_mesh_sizes(void* h, int mesh, int32_t* nV, int32_t* nF, int32_t* nE) {
    const TriMesh& m = ((ClothScene*)h)->meshes[mesh];
    if (nV) *nV = m.nV;
    if (nF) *nF = m.nF;
    if (nE) *nE = m.nE;
    return GAIA_OK;
}
static void copy_out(const std::vector<V3>& v, float* out) { std::memcpy(out, v.data(), v.size() * sizeof(V3)); }
int oracle_cloth_get_positions(void* h, int mesh, float* xyz) { copy_out(((ClothScene*)h)->meshes[mesh].pos, xyz); return GAIA_OK; }
int oracle_cloth_get_velocities(void* h, int mesh, float* xyz) { copy_out(((ClothScene*)h)->meshes[mesh].vel, xyz); return GAIA_OK; }
int oracle_cloth_set_state(void* h, int mesh, const float* pos, const float* vel) {
    TriMesh& m = ((ClothScene*)h)->meshes[mesh];
    if (pos) std::memcpy((void*)m.pos.data(), pos, m.pos.size() * sizeof(V3));
    if (vel) std::memcpy((void*)m.vel.data(), vel, m.vel.size() * sizeof(V3));
    return GAIA_OK;
}
int oracle_cloth_set_prev_positions(void* h, int mesh, const float* prev) {
    TriMesh& m = ((ClothScene*)h)->meshes[mesh];
    std::memcpy((void*)m.prev.data(), prev, m.prev.size() * sizeof(V3));
    return GAIA_OK;
}
int64_t oracle_cloth_total_vertices(void* h) { int64_t n = 0; for (auto& m : ((ClothScene*)h)->meshes) n += m.nV; return n; }
int oracle_cloth_get_all_positions(void* h, float* xyz) {
    for (auto& m : ((ClothScene*)h)->meshes) { copy_out(m.pos, xyz); xyz += 3 * (size_t)m.nV; }
    return GAIA_OK;
}
int oracle_cloth_set_all_state(void* h, const float* pos, const float* vel) {
    for (auto& m : ((ClothScene*)h)->meshes) {
        if (pos) { std::memcpy((void*)m.pos.data(), pos, m.pos.size() * sizeof(V3)); pos += 3 * (size_t)m.nV; }
        if (vel) { std::memcpy((void*)m.vel.data(), vel, m.vel.size() * sizeof(V3)); vel += 3 * (size_t)m.nV; }
    }
    return GAIA_OK;
}
int oracle_cloth_get_edges(void* h, int mesh, int32_t* out) {
    const TriMesh& m = ((ClothScene*)h)->meshes[mesh];
    for (int e = 0; e < m.nE; e++) {
        const ClothEdge& E = m.edges[e];
        const int32_t v[6] = {E.eV1, E.eV2, E.eV12Next, E.eV21Next, E.fId1, E.fId2};
        std::memcpy(out + 6 * (size_t)e, v, sizeof(v));
    }
    return GAIA_OK;
}
int oracle_cloth_num_parallel_groups(void* h) { return (int)((ClothScene*)h)->groups.size(); }
int oracle_cloth_get_parallel_group(void* h, int g, int32_t* out, int32_t* nPairs) {
    const auto& grp = ((ClothScene*)h)->groups[g];
    if (nPairs) *nPairs = (int32_t)(grp.size() / 2);
    if (out) std::memcpy(out, grp.data(), grp.size() * sizeof(int32_t));
    return GAIA_OK;
}
int oracle_cloth_get_vertex_masses(void* h, int mesh, float* out) {
    const TriMesh& m = ((ClothScene*)h)->meshes[mesh];
    std::memcpy(out, m.mass.data(), m.mass.size() * sizeof(float));
    return GAIA_OK;
}
int oracle_cloth_detect_contacts(void* h) { cloth_detect(*(ClothScene*)h); return GAIA_OK; }
int oracle_cloth_get_vf_contacts(void* h, gaia_cloth_vf_contact* out, int64_t cap, int64_t* n) {
    int64_t k = 0;
    for (auto& m : ((ClothScene*)h)->meshes)
        for (auto& l : m.vf)
            for (auto& c : l) {
                if (out && k < cap) out[k] = {c.vMesh, c.v, c.fMesh, c.f, c.primType, c.primId};
                k++;
            }
    if (n) *n = k;
    return GAIA_OK;
}
int oracle_cloth_get_ee_contacts(void* h, gaia_cloth_ee_contact* out, int64_t cap, int64_t* n) {
    int64_t k = 0;
    for (auto& m : ((ClothScene*)h)->meshes)
        for (auto& l : m.ee)
            for (auto& c : l) {
                if (out && k < cap) out[k] = {c.m1, c.e1, c.m2, c.e2};
                k++;
            }
    if (n) *n = k;
    return GAIA_OK;
}
int oracle_cloth_get_conservative_bounds(void* h, int mesh, float* out) {
    const TriMesh& m = ((ClothScene*)h)->meshes[mesh];
    std::memcpy(out, m.bounds.data(), m.bounds.size() * sizeof(float));
    return GAIA_OK;
}
int oracle_cloth_get_stats(void* h, gaia_cloth_stats* out) { *out = ((ClothScene*)h)->stats; return GAIA_OK; }
int oracle_cloth_reset_stats(void* h) {
    ClothScene* s = (ClothScene*)h;
    const int64_t a = s->stats.numContactsVF, b = s->stats.numContactsEE;
    std::memset(&s->stats, 0, sizeof(s->stats));
    s->stats.numContactsVF = a; s->stats.numContactsEE = b;
    return GAIA_OK;
}
int oracle_cloth_set_profiling(void*, int) { return GAIA_OK; }

// ---- probes
int oracle_cloth_material_force_hessian(void* h, int mesh, int v, float* f3, float* h9) {
    ClothScene* s = (ClothScene*)h;
    V3 f; M3 H;
    cloth_material_force_hessian(*s, s->meshes[mesh], v, f, H);
    f3[0] = f.x; f3[1] = f.y; f3[2] = f.z;
    std::memcpy(h9, H.m, sizeof(H.m));
    return GAIA_OK;
}
int oracle_cloth_vertex_force_hessian(void* h, int mesh, int v, float* f3, float* h9) {
    ClothScene* s = (ClothScene*)h;
    V3 f; M3 H;
    cloth_vertex_force_hessian(*s, mesh, v, f, H);
    f3[0] = f.x; f3[1] = f.y; f3[2] = f.z;
    std::memcpy(h9, H.m, sizeof(H.m));
    return GAIA_OK;
}
static void out_fh(const V3& f, const M3& H, float* f3, float* h9) { f3[0] = f.x; f3[1] = f.y; f3[2] = f.z; std::memcpy(h9, H.m, sizeof(H.m)); }
float oracle_cloth_probe_vf(void* h, int vMesh, int v, int fMesh, int f, int order, float* f3, float* h9) {
    V3 F; M3 H; const float d = cloth_probe_vf(*(ClothScene*)h, vMesh, v, fMesh, f, order, F, H); out_fh(F, H, f3, h9); return d;
}
float oracle_cloth_probe_ee(void* h, int m1, int e1, int m2, int e2, int order, float* f3, float* h9) {
    V3 F; M3 H; const float d = cloth_probe_ee(*(ClothScene*)h, m1, e1, m2, e2, order, F, H); out_fh(F, H, f3, h9); return d;
}
void oracle_cloth_probe_boundary(void* h, int mesh, int v, float* f3, float* h9) {
    V3 F; M3 H; cloth_probe_boundary(*(ClothScene*)h, mesh, v, F, H); out_fh(F, H, f3, h9);
}
// topology the cloth feasible-region test reads: edge id of each face edge [3 nF], CSR of each vertex's neighbour vertices
int oracle_cloth_get_topology(void* h, int mesh, int32_t* face3Edges, int32_t* vNbrOff, int32_t* vNbrV, int64_t* nNbr) {
    const TriMesh& m = ((ClothScene*)h)->meshes[mesh];
    if (face3Edges) std::memcpy(face3Edges, m.face3Edges.data(), m.face3Edges.size() * 4);
    if (vNbrOff) std::memcpy(vNbrOff, m.vNbrVOff.data(), m.vNbrVOff.size() * 4);
    if (vNbrV) std::memcpy(vNbrV, m.vNbrV.data(), m.vNbrV.size() * 4);
    if (nNbr) *nNbr = (int64_t)m.vNbrV.size();
    return GAIA_OK;
}
int oracle_cloth_check_feasible_region(void* h, int mesh, const float* p3, int faceId, int type, float eps) {
    return cloth_probe_feasible_region(((ClothScene*)h)->meshes[mesh], V3(p3[0], p3[1], p3[2]), faceId, type, eps) ? 1 : 0;
}
int oracle_cloth_probe_initial_step(void* h) { cloth_probe_initial_step(*(ClothScene*)h); return GAIA_OK; }
int oracle_cloth_probe_update_velocity(void* h) { cloth_probe_update_velocity(*(ClothScene*)h); return GAIA_OK; }
int oracle_cloth_probe_truncate(void* h, int mesh, int v, float* newPos3) {
    V3 p(newPos3[0], newPos3[1], newPos3[2]);
    const bool t = cloth_probe_truncate(*(ClothScene*)h, mesh, v, p);
    newPos3[0] = p.x; newPos3[1] = p.y; newPos3[2] = p.z;
    return t ? 1 : 0;
}
int oracle_cloth_get_inertia(void* h, int mesh, float* xyz) { copy_out(((ClothScene*)h)->meshes[mesh].inertia, xyz); return GAIA_OK; }
// (tests) conservative bounds and the positions of the last detection, as a detection would have left them
int oracle_cloth_set_bounds(void* h, int mesh, const float* bounds, const float* posAtCD) {
    TriMesh& m = ((ClothScene*)h)->meshes[mesh];
    if (bounds) { m.bounds.resize(m.nV); std::memcpy(m.bounds.data(), bounds, m.nV * sizeof(float)); }
    if (posAtCD) { m.posAtCD.resize(m.nV); std::memcpy(m.posAtCD.data(), posAtCD, m.nV * sizeof(V3)); }
    return GAIA_OK;
}
float oracle_cloth_elastic_energy(void* h, int mesh) { ClothScene* s = (ClothScene*)h; return cloth_elastic_energy(*s, s->meshes[mesh]); }
void oracle_cloth_contact_repulsion(void* h, float dis, float* dEdD, float* d2E) { cloth_contact_repulsion(*(ClothScene*)h, dis, *dEdD, *d2E); }
int oracle_cloth_get_dminv(void* h, int mesh, float* out4, float* area) {
    const TriMesh& m = ((ClothScene*)h)->meshes[mesh];
    if (out4) std::memcpy(out4, m.DmInv.data(), m.DmInv.size() * sizeof(float));
    if (area) std::memcpy(area, m.restArea.data(), m.restArea.size() * sizeof(float));
    return GAIA_OK;
}
int oracle_cloth_get_bending_forms(void* h, int mesh, float* out16) {
    const TriMesh& m = ((ClothScene*)h)->meshes[mesh];
    for (int e = 0; e < m.nE; e++) std::memcpy(out16 + 16 * (size_t)e, m.Q[e].data(), 16 * sizeof(float));
    return GAIA_OK;
}
void oracle_cloth_closest_points_segments(const float* p0, const float* p1, const float* q0, const float* q1, float* out8) {
    V3 ps, qt; float s, t;
    orc::closest_points_segments(V3(p0[0], p0[1], p0[2]), V3(p1[0], p1[1], p1[2]), V3(q0[0], q0[1], q0[2]), V3(q1[0], q1[1], q1[2]), ps, qt, s, t);
    out8[0] = ps.x; out8[1] = ps.y; out8[2] = ps.z; out8[3] = qt.x; out8[4] = qt.y; out8[5] = qt.z; out8[6] = s; out8[7] = t;
}

}  // extern "C"
