// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// C entry points of the CPU oracle; they shadow the product's C ABI (include/gaia_b200.h) one to one, with the
// prefix `oracle_`, so that tests drive both implementations with the same calls.
#include "vbd_oracle.h"
#include "oracle_ccd.h"
#include <algorithm>
#include <cstdio>

using namespace orc;

namespace orc {
void collision_set_brute_force(Scene& s, bool on);
std::vector<int> morton_leaf_order(const Scene& s, int which);
void probe_boundary(const Scene& s, const Mesh& m, int iV, V3& force, M3& hessian, bool apply_friction);
void probe_collision_per_collision(const ColResult& r, int order, V3& force, M3& hessian);
}

extern "C" {

void* oracle_create(const gaia_physics_params* phys, const gaia_collision_params* col, int nThreads) {
    Scene* s = new Scene();
    s->phys = *phys; s->col = *col;
    // BasePhysicFramework::timeStepEqualize (BasePhysicsFramework.cpp:245-249)
    s->dt = s->phys.timeStep / s->phys.numSubsteps;
    s->dtSqrReciprocal = 1.f / (s->dt * s->dt);
    std::memset(&s->stats, 0, sizeof(s->stats));
    if (nThreads <= 0) nThreads = (int)std::thread::hardware_concurrency();
    s->pool = new Pool(nThreads);
    return s;
}
void oracle_destroy(void* h) { delete (Scene*)h; }
int oracle_num_threads(void* h) { return ((Scene*)h)->pool->size(); }

int oracle_add_tet_mesh(void* h, int32_t nV, int32_t nT, const float* xyz, const int32_t* tets, const gaia_object_params* obj,
                        int32_t nColors, const int32_t* colorOffsets, const int32_t* colorVerts, int32_t nFixed, const int32_t* fixed) {
    Scene* s = (Scene*)h;
    if (s->finalized) return GAIA_ERR_STATE;
    s->meshes.emplace_back();
    Mesh& m = s->meshes.back();
    m.id = (int)s->meshes.size() - 1;
    gaia_object_params o = *obj;
    // VBDPhysics::timeStepEqualize (VBDPhysics.cpp:794-807)
    o.exponentialVelDamping = std::pow(o.exponentialVelDamping, s->dt);
    o.constantVelDamping = o.constantVelDamping * s->dt;
    mesh_initialize(m, nV, nT, xyz, tets, o, nColors, colorOffsets, colorVerts, nFixed, fixed, s->keepColourOrder);
    return m.id;
}
static int add_deformer(void* h, int type, int mesh, const int32_t* verts, int32_t n, const float* origin, const float* axis, const float* speed,
                        double w, double endTime) {
    Scene* s = (Scene*)h;
    if (s->finalized) return GAIA_ERR_STATE;
    Mesh& m = s->meshes[mesh];
    Deformer d;
    d.type = type; d.mesh = mesh; d.verts.assign(verts, verts + n); d.angularVelocity = w; d.endTime = endTime;
    if (origin) d.origin = V3(origin[0], origin[1], origin[2]);
    if (axis) d.axis = V3(axis[0], axis[1], axis[2]);
    if (speed) d.speed = V3(speed[0], speed[1], speed[2]);
    for (int i = 0; i < n; i++) {
        const int v = verts[i];
        if (!m.fixedMask[v]) { m.fixedMask[v] = 1; m.fixedPoints.push_back(v); }
        const V3 P = m.pos[v];
        if (type == 0) {
            const V3 OP = P - d.origin;
            const V3 root = dot(OP, d.axis) * d.axis;
            d.roots.push_back(root); d.rootPs.push_back(P - root);
        } else { d.roots.push_back(P); d.rootPs.push_back(P); }
    }
    s->deformers.push_back(d);
    return GAIA_OK;
}
int oracle_add_rotator(void* h, int mesh, const int32_t* verts, int32_t n, const float* origin, const float* axis, double w, double endTime) {
    return add_deformer(h, 0, mesh, verts, n, origin, axis, nullptr, w, endTime);
}
int oracle_add_translator(void* h, int mesh, const int32_t* verts, int32_t n, const float* speed, double endTime) {
    return add_deformer(h, 1, mesh, verts, n, nullptr, nullptr, speed, 0.0, endTime);
}

// multi-process tests: a partition keeps the global colour order; ghosts behave like fixed points (never solved, overwritten
// by the exchange)
int oracle_set_partition_mode(void* h, int32_t, int32_t) { ((Scene*)h)->keepColourOrder = true; return GAIA_OK; }
int oracle_set_ghosts(void* h, int mesh, const int32_t* verts, int32_t n) {
    Mesh& m = ((Scene*)h)->meshes[mesh];
    for (int i = 0; i < n; i++) if (!m.fixedMask[verts[i]]) { m.fixedMask[verts[i]] = 1; m.fixedPoints.push_back(verts[i]); }
    return GAIA_OK;
}
int oracle_set_line_search(void* h, int32_t enabled, float tau, float tauGD, int32_t maxIters) {
    Scene* s = (Scene*)h;
    s->lineSearch.enabled = enabled != 0; s->lineSearch.tau = tau; s->lineSearch.tauGD = tauGD; s->lineSearch.maxIters = maxIters;
    return GAIA_OK;
}
// parity surface of the line search's merit energy: inertia and elastic part with vertex v of `mesh` at x
int oracle_vertex_merit_energy(void* h, int mesh, int v, const float* x3, float* inertiaOut, float* elasticOut) {
    Scene* s = (Scene*)h;
    if (mesh < 0 || mesh >= (int)s->meshes.size() || v < 0 || v >= s->meshes[mesh].nV) return GAIA_ERR_INVALID_ARG;
    float in = 0.f, el = 0.f;
    vertex_merit_energy_parts(*s, s->meshes[mesh], v, V3(x3[0], x3[1], x3[2]), in, el);
    *inertiaOut = in; *elasticOut = el;
    return GAIA_OK;
}
int64_t oracle_line_search_shrinks(void* h) { return ((Scene*)h)->lineSearchShrinks; }
// probe: VBDTetMeshNeoHookean::accumlateMaterialForce for every vertex of a mesh at the current positions ([nV][3])
int oracle_neohookean_forces(void* h, int mesh, float* out) {
    const Mesh& m = ((Scene*)h)->meshes[mesh];
    for (int v = 0; v < m.nV; v++) { const V3 f = neohookean_force(m, v); out[3 * v] = f.x; out[3 * v + 1] = f.y; out[3 * v + 2] = f.z; }
    return GAIA_OK;
}
// schedule trace of run_step (tests): start / stop recording; read the events back (4 ints each). Single-threaded scenes only.
int oracle_trace_begin(void* h) {
    Scene* s = (Scene*)h;
    if (s->pool->size() != 1) return GAIA_ERR_STATE;
    delete s->trace; s->trace = new std::vector<int32_t>();
    return GAIA_OK;
}
int64_t oracle_trace_read(void* h, int32_t* out, int64_t capacity) {
    Scene* s = (Scene*)h;
    if (!s->trace) return 0;
    const int64_t n = (int64_t)s->trace->size();
    if (out && capacity >= n) std::memcpy(out, s->trace->data(), n * sizeof(int32_t));
    return n;
}
int oracle_phase_begin_substep(void* h) { enable_models(*(Scene*)h); phase_begin_substep(*(Scene*)h); return GAIA_OK; }
int oracle_phase_prepare(void* h) { phase_prepare(*(Scene*)h); return GAIA_OK; }
int oracle_phase_solve_group(void* h, int g, int applyFriction) { phase_solve_group(*(Scene*)h, g, applyFriction != 0); return GAIA_OK; }
int oracle_phase_intermediate_cd(void* h) { phase_intermediate_cd(*(Scene*)h); return GAIA_OK; }
int oracle_phase_end_substep(void* h) { phase_end_substep(*(Scene*)h); return GAIA_OK; }

int oracle_finalize(void* h) { scene_finalize(*(Scene*)h); return GAIA_OK; }

int oracle_run_frames(void* h, int32_t nFrames) {
    Scene* s = (Scene*)h;
    if (!s->finalized) return GAIA_ERR_STATE;
    for (int f = 0; f < nFrames; f++) {
        enable_models(*s);
        run_step(*s);
        ++s->frameId;
        if (!s->err.empty()) return GAIA_ERR_NAN;
    }
    return GAIA_OK;
}
int oracle_run_iterations(void* h, int32_t nIters) {
    Scene* s = (Scene*)h;
    if (!s->finalized) return GAIA_ERR_STATE;
    enable_models(*s);
    run_iterations(*s, nIters);
    return GAIA_OK;
}
int oracle_frame_id(void* h) { return ((Scene*)h)->frameId; }
int oracle_num_meshes(void* h) { return (int)((Scene*)h)->meshes.size(); }
int oracle_mesh_sizes(void* h, int mesh, int32_t* nV, int32_t* nT, int32_t* nSV, int32_t* nSF) {
    Mesh& m = ((Scene*)h)->meshes[mesh];
    *nV = m.nV; *nT = m.nT; *nSV = m.nSurfV(); *nSF = m.nSurfF();
    return GAIA_OK;
}
static void copy_out(const std::vector<V3>& v, float* out) { std::memcpy(out, v.data(), v.size() * sizeof(V3)); }
int oracle_get_positions(void* h, int mesh, float* xyz) { copy_out(((Scene*)h)->meshes[mesh].pos, xyz); return GAIA_OK; }
int oracle_get_prev_positions(void* h, int mesh, float* xyz) { copy_out(((Scene*)h)->meshes[mesh].prev, xyz); return GAIA_OK; }
int oracle_get_velocities(void* h, int mesh, float* xyz) { copy_out(((Scene*)h)->meshes[mesh].vel, xyz); return GAIA_OK; }
int oracle_get_inertia(void* h, int mesh, float* xyz) { copy_out(((Scene*)h)->meshes[mesh].inertia, xyz); return GAIA_OK; }
int oracle_set_state(void* h, int mesh, const float* pos, const float* vel) {
    Mesh& m = ((Scene*)h)->meshes[mesh];
    if (pos) std::memcpy(m.pos.data(), pos, m.pos.size() * sizeof(V3));
    if (vel) std::memcpy(m.vel.data(), vel, m.vel.size() * sizeof(V3));
    return GAIA_OK;
}
int oracle_set_prev_positions(void* h, int mesh, const float* prev) {
    Mesh& m = ((Scene*)h)->meshes[mesh];
    std::memcpy(m.prev.data(), prev, m.prev.size() * sizeof(V3));
    return GAIA_OK;
}
int oracle_get_surface_positions(void* h, int mesh, float* xyz) {
    Mesh& m = ((Scene*)h)->meshes[mesh];
    for (int i = 0; i < m.nSurfV(); i++) { V3 p = m.pos[m.surfaceVIds[i]]; xyz[3 * i] = p.x; xyz[3 * i + 1] = p.y; xyz[3 * i + 2] = p.z; }
    return GAIA_OK;
}
int oracle_get_world_bounds(void* h, float* mn, float* mx) {
    Scene* s = (Scene*)h;
    for (int d = 0; d < 3; d++) { mn[d] = s->phys.worldBoundsMin[d]; mx[d] = s->phys.worldBoundsMax[d]; }
    return GAIA_OK;
}
int oracle_num_parallel_groups(void* h) { return (int)((Scene*)h)->vertexParallelGroups.size(); }
int oracle_get_parallel_group(void* h, int g, int32_t* out, int32_t* nPairs) {
    auto& v = ((Scene*)h)->vertexParallelGroups[g];
    *nPairs = (int32_t)v.size() / 2;
    if (out) std::memcpy(out, v.data(), v.size() * sizeof(int32_t));
    return GAIA_OK;
}
int oracle_get_surface_vertex_ids(void* h, int mesh, int32_t* out) {
    auto& v = ((Scene*)h)->meshes[mesh].surfaceVIds; std::memcpy(out, v.data(), v.size() * 4); return GAIA_OK;
}
int oracle_get_surface_faces(void* h, int mesh, int32_t* out) {
    auto& v = ((Scene*)h)->meshes[mesh].surfFaces; std::memcpy(out, v.data(), v.size() * 4); return GAIA_OK;
}
int oracle_get_vertex_masses(void* h, int mesh, float* out) {
    auto& v = ((Scene*)h)->meshes[mesh].mass; std::memcpy(out, v.data(), v.size() * 4); return GAIA_OK;
}
int oracle_get_dminv(void* h, int mesh, float* out) {
    auto& v = ((Scene*)h)->meshes[mesh].DmInv; std::memcpy(out, v.data(), v.size() * 4); return GAIA_OK;
}
int oracle_get_rest_volumes(void* h, int mesh, float* out) {
    auto& v = ((Scene*)h)->meshes[mesh].restVol; std::memcpy(out, v.data(), v.size() * 4); return GAIA_OK;
}
int oracle_get_vertex_tet_csr(void* h, int mesh, int32_t* infos, int32_t* tets, int32_t* orders) {
    Mesh& m = ((Scene*)h)->meshes[mesh];
    std::memcpy(infos, m.vNbrInfos.data(), m.vNbrInfos.size() * 4);
    std::memcpy(tets, m.vNbrTets.data(), m.vNbrTets.size() * 4);
    std::memcpy(orders, m.vNbrOrder.data(), m.vNbrOrder.size() * 4);
    return GAIA_OK;
}
// parity probes of the tetrahedral traversal (tests/test_oracle_vs_ref.py): the topology it walks, and one walk
int oracle_get_tet_topology(void* h, int mesh, int32_t* tets, int32_t* nbrTets, int32_t* xorSums) {
    Mesh& m = ((Scene*)h)->meshes[mesh];
    if (tets) std::memcpy(tets, m.tets.data(), m.tets.size() * 4);
    if (nbrTets) std::memcpy(nbrTets, m.tetNbrTets.data(), m.tetNbrTets.size() * 4);
    if (xorSums) std::memcpy(xorSums, m.tetXorSums.data(), m.tetXorSums.size() * 4);
    return GAIA_OK;
}
int oracle_tet_traverse(void* h, int mesh, const float* origin3, const float* dir3, float maxDis, int startTet, int startFace, int targetTet, float eps,
                        int32_t* numTraversed) {
    Mesh& m = ((Scene*)h)->meshes[mesh];
    int n = 0;
    const bool ok = tetrahedral_traverse_to(m, V3(origin3[0], origin3[1], origin3[2]), V3(dir3[0], dir3[1], dir3[2]), maxDis, startTet, startFace, targetTet, eps, &n);
    if (numTraversed) *numTraversed = n;
    return ok ? 1 : 0;
}
// surface topology the feasible-region test reads: neighbour face across each edge [3 nF], tet vertex -> surface vertex [nV],
// CSR of the neighbouring surface vertices (tet-mesh ids) of each surface vertex. nbrVerts == nullptr: only *nNbr is written
int oracle_get_surface_topology(void* h, int mesh, int32_t* face3Nbr, int32_t* tetVertToSurf, int32_t* nbrOff, int32_t* nbrVerts, int64_t* nNbr) {
    Mesh& m = ((Scene*)h)->meshes[mesh];
    if (face3Nbr) std::memcpy(face3Nbr, m.surfFace3Nbr.data(), m.surfFace3Nbr.size() * 4);
    if (tetVertToSurf) std::memcpy(tetVertToSurf, m.tetVertToSurf.data(), m.tetVertToSurf.size() * 4);
    int64_t run = 0;
    for (size_t sv = 0; sv < m.surfVertNbrSurfVerts.size(); sv++) {
        if (nbrOff) nbrOff[sv] = (int32_t)run;
        for (int v : m.surfVertNbrSurfVerts[sv]) { if (nbrVerts) nbrVerts[run] = v; run++; }
    }
    if (nbrOff) nbrOff[m.surfVertNbrSurfVerts.size()] = (int32_t)run;
    if (nNbr) *nNbr = run;
    return GAIA_OK;
}
int oracle_check_feasible_region(void* h, int mesh, const float* p3, int faceId, int type, float eps) {
    return probe_feasible_region(((Scene*)h)->meshes[mesh], V3(p3[0], p3[1], p3[2]), faceId, type, eps) ? 1 : 0;
}
int oracle_enable_mesh(void* h, int mesh, int col, int solve) {
    Mesh& m = ((Scene*)h)->meshes[mesh];
    m.activeForCollision = col != 0; m.activeForMaterialSolve = solve != 0;
    return GAIA_OK;
}
int oracle_set_brute_force(void* h, int on) { collision_set_brute_force(*(Scene*)h, on != 0); return GAIA_OK; }

int oracle_detect_collisions(void* h, int doDCD, int doCCD) {
    Scene* s = (Scene*)h;
    gaia_collision_params saved = s->col;
    s->col.allowDCD = saved.allowDCD && doDCD; s->col.allowCCD = saved.allowCCD && doCCD;
    s->substep = 0;
    dcd(*s, false);
    ccd(*s);
    prepare_collision_data(*s);
    s->col = saved;
    return GAIA_OK;
}
int oracle_get_contacts(void* h, gaia_contact* out, int64_t capacity, int64_t* n) {
    Scene* s = (Scene*)h;
    int64_t cnt = 0;
    for (size_t im = 0; im < s->meshes.size(); im++) {
        Mesh& m = s->meshes[im];
        for (int sv = 0; sv < m.nSurfV(); sv++) {
            ColResult& r = m.colResults[sv];
            if (!r.numIntersections) continue;
            if (out && cnt < capacity) {
                gaia_contact c;
                c.queryMesh = r.idTMQuery; c.queryVertex = r.idVQuery; c.hitMesh = r.intersectedMeshId;
                c.hitTet = r.fromCCD ? -1 : r.intersectedElement; c.hitFace = r.closestSurfaceFaceId;
                c.fromCCD = r.fromCCD; c.shortestPathFound = r.shortestPathFound; c.closestPointType = r.closestPointType;
                out[cnt] = c;
            }
            cnt++;
        }
    }
    *n = cnt;
    return GAIA_OK;
}
// a15 parity surfaces (twins of gaia_b200_get_active_list / get_relations): prepare_collision_data's outputs, sorted
int oracle_get_active_list(void* h, int group, int32_t* out, int64_t capacity, int64_t* n) {
    Scene* s = (Scene*)h;
    if (group < 0 || group >= (int)s->activeCollisionsEachGroup.size()) { *n = 0; return GAIA_OK; }
    const std::vector<int>& g = s->activeCollisionsEachGroup[group];
    std::vector<std::pair<int, int>> pairs;
    for (size_t i = 0; i + 1 < g.size(); i += 2) pairs.push_back({g[i], g[i + 1]});
    std::sort(pairs.begin(), pairs.end());
    *n = (int64_t)pairs.size();
    if (out) for (int64_t i = 0; i < *n && i < capacity; i++) { out[2 * i] = pairs[i].first; out[2 * i + 1] = pairs[i].second; }
    return GAIA_OK;
}
// raw views of the prepared collision data, in the order prepare_collision_data built them (the pin against the reference's own
// prepareCollisionDataCPU / addToActiveCollisionList compares order and all): active pairs of a group; per vertex of a mesh its
// relation rows (meshId, surfaceVertexId, collisionId, collisionType, collisionVertexOrder) behind relOff [nV + 1]; the mask;
// the group of every vertex; the detection result of every surface vertex
int64_t oracle_get_active_list_raw(void* h, int group, int32_t* out, int64_t capacity) {
    Scene* s = (Scene*)h;
    const std::vector<int>& g = s->activeCollisionsEachGroup[group];
    if (out && capacity >= (int64_t)g.size()) for (size_t i = 0; i < g.size(); i++) out[i] = g[i];
    return (int64_t)g.size() / 2;
}
int64_t oracle_get_relations_raw(void* h, int mesh, int32_t* relOff, int32_t* rows, int64_t capacityRows, uint8_t* mask, int32_t* vertexGroup) {
    Scene* s = (Scene*)h;
    Mesh& m = s->meshes[mesh];
    int64_t n = 0;
    for (int v = 0; v < m.nV; v++) {
        if (relOff) relOff[v] = (int32_t)n;
        for (auto& r : m.colRelations[v]) {
            if (rows && n < capacityRows) { rows[5 * n] = r.meshId; rows[5 * n + 1] = r.surfaceVertexId; rows[5 * n + 2] = r.collisionId; rows[5 * n + 3] = r.collisionType; rows[5 * n + 4] = r.collisionVertexOrder; }
            n++;
        }
        if (mask) mask[v] = m.activeCollisionMask[v];
        if (vertexGroup) vertexGroup[v] = m.vertexParallelGroup[v];
    }
    if (relOff) relOff[m.nV] = (int32_t)n;
    return n;
}
int oracle_get_col_results(void* h, int mesh, uint8_t* has, uint8_t* fromCCD, uint8_t* pathFound, int32_t* hitMesh, int32_t* hitFace) {
    Mesh& m = ((Scene*)h)->meshes[mesh];
    for (int sv = 0; sv < m.nSurfV(); sv++) {
        const ColResult& r = m.colResults[sv];
        has[sv] = r.numIntersections ? 1 : 0; fromCCD[sv] = r.fromCCD ? 1 : 0; pathFound[sv] = r.shortestPathFound ? 1 : 0;
        hitMesh[sv] = r.intersectedMeshId; hitFace[sv] = r.closestSurfaceFaceId;
    }
    return GAIA_OK;
}
int oracle_get_relations(void* h, int mesh, int32_t* out, int64_t capacity, int64_t* n) {
    Scene* s = (Scene*)h;
    Mesh& m = s->meshes[mesh];
    struct Row { int32_t v, qm, qv, k; };
    std::vector<Row> rows;
    for (int v = 0; v < (int)m.colRelations.size(); v++)
        for (auto& r : m.colRelations[v]) rows.push_back({v, r.meshId, s->meshes[r.meshId].surfaceVIds[r.surfaceVertexId], r.collisionVertexOrder});
    std::sort(rows.begin(), rows.end(), [](const Row& a, const Row& b) {
        if (a.v != b.v) return a.v < b.v; if (a.qm != b.qm) return a.qm < b.qm; if (a.qv != b.qv) return a.qv < b.qv; return a.k < b.k; });
    *n = (int64_t)rows.size();
    if (out) for (int64_t i = 0; i < *n && i < capacity; i++) { out[4 * i] = rows[i].v; out[4 * i + 1] = rows[i].qm; out[4 * i + 2] = rows[i].qv; out[4 * i + 3] = rows[i].k; }
    return GAIA_OK;
}
int oracle_get_bvh_leaf_order(void* h, int which, int32_t* out, int64_t capacity, int64_t* n) {
    std::vector<int> o = morton_leaf_order(*(Scene*)h, which);
    *n = (int64_t)o.size();
    if (out) { if (capacity < *n) return GAIA_ERR_INVALID_ARG; std::memcpy(out, o.data(), o.size() * 4); }
    return GAIA_OK;
}
// per-vertex force / Hessian probe for pinning against oracle/_ref (not part of the product ABI)
int oracle_vertex_material_force_hessian(void* h, int mesh, int v, float* force3, float* hessian9) {
    Scene* s = (Scene*)h;
    V3 f; M3 H;
    neohookean_force_hessian(*s, s->meshes[mesh], v, f, H);
    force3[0] = f.x; force3[1] = f.y; force3[2] = f.z;
    std::memcpy(hessian9, H.m, 36);
    return GAIA_OK;
}
// one vertex-face contact evaluated the CPU way (updateCollisionInfo + accumlateCollisionForceAndHessianPerCollision):
// query vertex p in mesh 0, triangle tri (3 vertices) in mesh 1; order 0..2 = face vertex, 3 = the query vertex
int oracle_contact_force_hessian(const float* p, const float* pPrev, const float* tri, const float* triPrev, int order, float k,
                                 float mu0, float mu1, float epsV0, float epsV1, float dt, float* force3, float* hessian9, float* depth) {
    Scene s;
    std::memset(&s.phys, 0, sizeof(s.phys));
    s.phys.collisionStiffness = k; s.dt = dt;
    s.meshes.resize(2);
    Mesh& a = s.meshes[0]; Mesh& b = s.meshes[1];
    a.pos = {V3(p[0], p[1], p[2])}; a.prev = {V3(pPrev[0], pPrev[1], pPrev[2])};
    a.obj.frictionDynamic = mu0; a.obj.frictionEpsV = epsV0;
    for (int i = 0; i < 3; i++) { b.pos.push_back(V3(tri[3 * i], tri[3 * i + 1], tri[3 * i + 2])); b.prev.push_back(V3(triPrev[3 * i], triPrev[3 * i + 1], triPrev[3 * i + 2])); }
    b.surfFaces = {0, 1, 2};
    b.obj.frictionDynamic = mu1; b.obj.frictionEpsV = epsV1;
    ColResult r;
    r.numIntersections = 1; r.idVQuery = 0; r.idTMQuery = 0; r.intersectedMeshId = 1; r.closestSurfaceFaceId = 0; r.shortestPathFound = true;
    update_collision_info(s, r);
    V3 f; M3 H;
    probe_collision_per_collision(r, order, f, H);
    force3[0] = f.x; force3[1] = f.y; force3[2] = f.z;
    std::memcpy(hessian9, H.m, 36);
    if (depth) *depth = r.penetrationDepth;
    return GAIA_OK;
}
int oracle_boundary_force_hessian(const float* p, const float* pPrev, const float* boundsMin, const float* boundsMax, float kB, float muB,
                                  float epsVB, float dt, int applyFriction, float* force3, float* hessian9) {
    Scene s;
    std::memset(&s.phys, 0, sizeof(s.phys));
    s.phys.usePlaneGround = 1; s.phys.boundaryCollisionStiffness = kB; s.phys.boundaryFrictionDynamic = muB; s.phys.boundaryFrictionEpsV = epsVB;
    for (int d = 0; d < 3; d++) { s.phys.worldBoundsMin[d] = boundsMin[d]; s.phys.worldBoundsMax[d] = boundsMax[d]; }
    s.dt = dt;
    Mesh m;
    m.pos = {V3(p[0], p[1], p[2])}; m.prev = {V3(pPrev[0], pPrev[1], pPrev[2])};
    V3 f; M3 H;
    probe_boundary(s, m, 0, f, H, applyFriction != 0);
    force3[0] = f.x; force3[1] = f.y; force3[2] = f.z;
    std::memcpy(hessian9, H.m, 36);
    return GAIA_OK;
}
// the bowl ground (useBowlGround) for one vertex
int oracle_boundary_bowl_force_hessian(const float* p, const float* pPrev, const float* center, float radius, int cap, float kB, float muB,
                                       float epsVB, float dt, int applyFriction, float* force3, float* hessian9) {
    Scene s;
    std::memset(&s.phys, 0, sizeof(s.phys));
    s.phys.useBowlGround = 1; s.phys.usePlaneGround = 1; s.phys.boundaryCollisionStiffness = kB; s.phys.boundaryFrictionDynamic = muB; s.phys.boundaryFrictionEpsV = epsVB;
    s.bowl.center = V3(center[0], center[1], center[2]); s.bowl.radius = radius; s.bowl.cap = cap != 0;
    s.dt = dt;
    Mesh m;
    m.pos = {V3(p[0], p[1], p[2])}; m.prev = {V3(pPrev[0], pPrev[1], pPrev[2])};
    V3 f; M3 H;
    probe_boundary(s, m, 0, f, H, applyFriction != 0);
    force3[0] = f.x; force3[1] = f.y; force3[2] = f.z;
    std::memcpy(hessian9, H.m, 36);
    return GAIA_OK;
}
int oracle_set_bowl_ground(void* h, const float* center, float radius, int32_t cap) {
    Scene* s = (Scene*)h;
    if (!center || !(radius > 0.f)) return GAIA_ERR_INVALID_ARG;
    s->bowl.center = V3(center[0], center[1], center[2]); s->bowl.radius = radius; s->bowl.cap = cap != 0;
    return GAIA_OK;
}
double oracle_cur_time(void* h) { return (double)((Scene*)h)->curTime; }
// probe: the Chebyshev weights omega_1 .. omega_n of the tet step
void oracle_accelerator_omegas(float rho, int n, float* out) { float w = 1.f; for (int k = 1; k <= n; k++) { w = accelerator_omega(k, rho, w); out[k - 1] = w; } }
// probe: the accelerator's position update for n vertices (v in / out)
void oracle_accelerate(int n, float* v, const float* ypp, float omega) {
    for (int i = 0; i < n; i++) { const V3 r = accelerate_position(V3(v[3 * i], v[3 * i + 1], v[3 * i + 2]), V3(ypp[3 * i], ypp[3 * i + 1], ypp[3 * i + 2]), omega); v[3 * i] = r.x; v[3 * i + 1] = r.y; v[3 * i + 2] = r.z; }
}
// probe: the inertia term for n vertices, added onto given force / Hessian accumulators (in / out)
int oracle_inertia_terms(int n, const float* mass, const float* inertia, const float* x, float dtSqrReciprocal, float* f, float* h) {
    for (int i = 0; i < n; i++) {
        V3 force(f[3 * i], f[3 * i + 1], f[3 * i + 2]); M3 H; std::memcpy(H.m, h + 9 * i, 36);
        inertia_force_hessian(mass[i], V3(inertia[3 * i], inertia[3 * i + 1], inertia[3 * i + 2]), V3(x[3 * i], x[3 * i + 1], x[3 * i + 2]), dtSqrReciprocal, force, H);
        f[3 * i] = force.x; f[3 * i + 1] = force.y; f[3 * i + 2] = force.z; std::memcpy(h + 9 * i, H.m, 36);
    }
    return GAIA_OK;
}
// probe: the tail of the vertex step on n (force, Hessian, position) triples; flags [n]: 0 no move, 1 Newton step, 2 gradient step
int oracle_vertex_step_tail(int n, const float* f, const float* h, float* x, float stepSize, float stepSizeGD, int32_t* flags) {
    for (int i = 0; i < n; i++) {
        M3 H; std::memcpy(H.m, h + 9 * i, 36);
        V3 dir; float st; bool ok;
        const int rc = vertex_step_direction(stepSize, stepSizeGD, V3(f[3 * i], f[3 * i + 1], f[3 * i + 2]), H, dir, st, ok);
        flags[i] = rc <= 0 ? rc : (ok ? 1 : 2);
        if (rc > 0) { const V3 p = V3(x[3 * i], x[3 * i + 1], x[3 * i + 2]) + st * dir; x[3 * i] = p.x; x[3 * i + 1] = p.y; x[3 * i + 2] = p.z; }
    }
    return GAIA_OK;
}
int oracle_solve3x3(const float* m9, const float* b3, float* out3) { return solve3x3_psd_stable(m9, b3, out3) ? 1 : 0; }
int oracle_get_stats(void* h, gaia_stats* out) { *out = ((Scene*)h)->stats; return GAIA_OK; }
int oracle_reset_stats(void* h) { std::memset(&((Scene*)h)->stats, 0, sizeof(gaia_stats)); return GAIA_OK; }

// stand-alone predicate probes (tests pin them against oracle/_ref and against the CUDA predicates)
int oracle_ccd_tri_point(const float* x0 /*9*/, const float* x1 /*9*/, const float* p0, const float* p1, float* t, float* bary3) {
    ccdm::Vec3 x[2][3], p[2];
    for (int k = 0; k < 3; k++) { x[0][k] = ccdm::Vec3(x0[3 * k], x0[3 * k + 1], x0[3 * k + 2]); x[1][k] = ccdm::Vec3(x1[3 * k], x1[3 * k + 1], x1[3 * k + 2]); }
    p[0] = ccdm::Vec3(p0[0], p0[1], p0[2]); p[1] = ccdm::Vec3(p1[0], p1[1], p1[2]);
    float tt = -1; ccdm::Vec3 b;
    bool hit = ccdm::IntersectContinuousTriPoint(tt, x, p, b);
    *t = tt; bary3[0] = b.x; bary3[1] = b.y; bary3[2] = b.z;
    return hit ? 1 : 0;
}
int oracle_cubic_roots(const float* coef4, float x0, float x1, float* roots3) { return ccdm::cubic_roots(roots3, coef4, x0, x1); }

}  // extern "C"

namespace orc {
bool point_in_tet(V3 p, V3 a, V3 b, V3 c, V3 d);
V3 closest_point_triangle(V3 p, V3 a, V3 b, V3 c, V3& bary, int& type);
}
extern "C" {
int oracle_point_in_tet(const float* p, const float* a, const float* b, const float* c, const float* d) {
    return point_in_tet(V3(p[0], p[1], p[2]), V3(a[0], a[1], a[2]), V3(b[0], b[1], b[2]), V3(c[0], c[1], c[2]), V3(d[0], d[1], d[2])) ? 1 : 0;
}
int oracle_closest_point_triangle(const float* p, const float* a, const float* b, const float* c, float* cp3, float* bary3) {
    V3 bary; int type;
    V3 cp = closest_point_triangle(V3(p[0], p[1], p[2]), V3(a[0], a[1], a[2]), V3(b[0], b[1], b[2]), V3(c[0], c[1], c[2]), bary, type);
    cp3[0] = cp.x; cp3[1] = cp.y; cp3[2] = cp.z; bary3[0] = bary.x; bary3[1] = bary.y; bary3[2] = bary.z;
    return type;
}
}
