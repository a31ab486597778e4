// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h, oracle_cloth.h).
// Cloth contact detection: ClothContactDetector::contactQueryVF / contactQueryEE and their Embree callbacks, the
// feasible-region tests and VBDClothSimulationFramework::applyCollisionDetection. Embree only supplies the candidate
// set of a radius query (every primitive whose box meets the query sphere); it is reproduced with a uniform grid (or by
// brute force), and candidates are visited in ascending (mesh, primitive) order (rule R5).
#include "oracle_cloth.h"
#include <unordered_map>

namespace orc {

static const float CMP_EPSILON = 1e-5f;
static const float CMP_EPSILON2 = 1e-10f;
static const float ABSOLUTE_RELAXIATION = 1e-6f;   // M/CollisionDetector/TriMeshCollisionGeometry.cpp:4

// checkEdgeFeasibleRegion (M/CollisionDetector/TriMeshCollisionGeometry.cpp:92-151)
static bool edge_feasible(V3 p, const TriMesh& m, int faceId, int edgeOrderInFace, float eps) {
    const ClothEdge& E = m.edges[m.face3Edges[3 * faceId + edgeOrderInFace]];
    const V3 v1 = m.pos[E.eV1], v2 = m.pos[E.eV2];
    const V3 AP = p - v1, BP = p - v2, AB = v2 - v1;
    const float relaxed = eps + ABSOLUTE_RELAXIATION;
    if (dot(AP, AB) < -relaxed) return false;
    if (dot(BP, AB) > relaxed) return false;
    const float ABSqrNorm = sqnorm(AB);
    if (ABSqrNorm < CMP_EPSILON2) return false;
    for (int side = 0; side < 2; side++) {
        const int vn = side == 0 ? E.eV12Next : E.eV21Next;
        if (vn == -1) continue;
        const V3 vNext = m.pos[vn];
        const float t = dot(AB, vNext - v1) / ABSqrNorm;
        const V3 foot = v1 + t * AB;
        if (dot(vNext - foot, p - foot) > relaxed) return false;
    }
    return true;
}

// checkVertexFeasibleRegion (:153-184)
static bool vertex_feasible(V3 p, const TriMesh& m, int vId, float eps) {
    const V3 A = m.pos[vId];
    const V3 AP = p - A;
    const float relaxed = -eps - ABSOLUTE_RELAXIATION;
    for (int i = m.vNbrVOff[vId]; i < m.vNbrVOff[vId + 1]; i++) {
        const V3 BA = A - m.pos[m.vNbrV[i]];
        if (dot(AP, BA) < relaxed) return false;
    }
    return true;
}

// checkFeasibleRegion (:7-41)
static bool feasible_region(V3 p, const TriMesh& m, int faceId, int type, float eps) {
    const int* fv = &m.faces[3 * faceId];
    switch (type) {
        case cAtInterior: return true;
        case cAtAB: return edge_feasible(p, m, faceId, 0, eps);
        case cAtBC: return edge_feasible(p, m, faceId, 1, eps);
        case cAtAC: return edge_feasible(p, m, faceId, 2, eps);
        case cAtA: return vertex_feasible(p, m, fv[0], eps);
        case cAtB: return vertex_feasible(p, m, fv[1], eps);
        case cAtC: return vertex_feasible(p, m, fv[2], eps);
        default: return false;
    }
}

bool cloth_probe_feasible_region(const TriMesh& m, V3 p, int faceId, int type, float eps) { return feasible_region(p, m, faceId, type, eps); }   // parity probe

// get_closest_points_between_segments (M/CollisionDetector/CollisionGeometry.h:160-267; Eberly / Godot)
void closest_points_segments(V3 p0, V3 p1, V3 q0, V3 q1, V3& ps, V3& qt, float& s, float& t) {
    const V3 p = p1 - p0, q = q1 - q0, r = p0 - q0;
    const float a = dot(p, p), b = dot(p, q), c = dot(q, q), d = dot(p, r), e = dot(q, r);
    s = 0.0f; t = 0.0f;
    const float det = a * c - b * b;
    if (det > CMP_EPSILON) {
        const float bte = b * e, ctd = c * d;
        if (bte <= ctd) {
            if (e <= 0.0f) { s = (-d >= a ? 1 : (-d > 0.0f ? -d / a : 0.0f)); t = 0.0f; }
            else if (e < c) { s = 0.0f; t = e / c; }
            else { s = (b - d >= a ? 1 : (b - d > 0.0f ? (b - d) / a : 0.0f)); t = 1; }
        } else {
            s = bte - ctd;
            if (s >= det) {
                if (b + e <= 0.0f) { s = (-d <= 0.0f ? 0.0f : (-d < a ? -d / a : 1)); t = 0.0f; }
                else if (b + e < c) { s = 1; t = (b + e) / c; }
                else { s = (b - d <= 0.0f ? 0.0f : (b - d < a ? (b - d) / a : 1)); t = 1; }
            } else {
                const float ate = a * e, btd = b * d;
                if (ate <= btd) { s = (-d <= 0.0f ? 0.0f : (-d >= a ? 1 : -d / a)); t = 0.0f; }
                else {
                    t = ate - btd;
                    if (t >= det) { s = (b - d <= 0.0f ? 0.0f : (b - d >= a ? 1 : (b - d) / a)); t = 1; }
                    else { s /= det; t /= det; }
                }
            }
        }
    } else {
        if (e <= 0.0f) { s = (-d <= 0.0f ? 0.0f : (-d >= a ? 1 : -d / a)); t = 0.0f; }
        else if (e >= c) { s = (b - d <= 0.0f ? 0.0f : (b - d >= a ? 1 : (b - d) / a)); t = 1; }
        else { s = 0.0f; t = e / c; }
    }
    ps = (1 - s) * p0 + s * p1;
    qt = (1 - t) * q0 + t * q1;
}

// ---------------------------------------------------------------------------------------------- candidate grid
// Uniform hash grid over primitive boxes inflated by `pad`; a query box returns every primitive whose inflated box
// overlaps it, as sorted global ids (mesh-major). A conservative superset of Embree's candidate set.
struct Grid {
    float h = 1.f;
    std::unordered_map<uint64_t, std::vector<int>> cells;
    static uint64_t key(int x, int y, int z) { return ((uint64_t)(uint32_t)(x & 0x1FFFFF) << 42) | ((uint64_t)(uint32_t)(y & 0x1FFFFF) << 21) | (uint32_t)(z & 0x1FFFFF); }
    int c(float v) const { return (int)std::floor(v / h); }
    void insert(int id, V3 lo, V3 hi) {
        for (int x = c(lo.x); x <= c(hi.x); x++) for (int y = c(lo.y); y <= c(hi.y); y++) for (int z = c(lo.z); z <= c(hi.z); z++) cells[key(x, y, z)].push_back(id);
    }
    void query(V3 lo, V3 hi, std::vector<int>& out) const {
        out.clear();
        for (int x = c(lo.x); x <= c(hi.x); x++) for (int y = c(lo.y); y <= c(hi.y); y++) for (int z = c(lo.z); z <= c(hi.z); z++) {
            auto it = cells.find(key(x, y, z));
            if (it != cells.end()) out.insert(out.end(), it->second.begin(), it->second.end());
        }
        std::sort(out.begin(), out.end());
        out.erase(std::unique(out.begin(), out.end()), out.end());
    }
};
static inline V3 vmin(V3 a, V3 b) { return V3(std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z)); }
static inline V3 vmax(V3 a, V3 b) { return V3(std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z)); }

// triMeshVFRadiusQueryWithTopologyFilteringAndFaceMinDisCaculatingFunc[Collider] (ClothContactDetector.cpp:144-288, 440-589)
// for one candidate face; returns through `res`; the face-side effects (min distance, back-links) are applied by the caller
// afterwards so that the query loop can run in parallel over vertices
static void vf_candidate(const ClothScene& s, int qMesh, int qV, V3 queryPt, int fMesh, int f, std::vector<VFContact>& res, float& minDis,
                         std::vector<std::pair<int64_t, float>>& faceDis) {
    const TriMesh& tm = s.meshes[fMesh];
    if (fMesh >= s.numSim && qMesh >= s.numSim) return;                 // collider callback: only simulated meshes query colliders
    const int* fv = &tm.faces[3 * f];
    if (qMesh == fMesh && (fv[0] == qV || fv[1] == qV || fv[2] == qV)) return;
    // ...WithSewing (:290-438), attached to meshes that have seams (:969-974): the filtered vertex-face pairs of that mesh are skipped
    if (qMesh == fMesh && !tm.sewingVertices.empty() && !tm.sewingVF.empty() && std::binary_search(tm.sewingVF[qV].begin(), tm.sewingVF[qV].end(), f)) return;
    V3 bary; int type;
    const V3 cp = closest_point_triangle(queryPt, tm.pos[fv[0]], tm.pos[fv[1]], tm.pos[fv[2]], bary, type);
    const float d = norm(queryPt - cp);
    if (!(d < s.phys.maxQueryDis)) return;
    minDis = std::min(d, minDis);
    faceDis.push_back({((int64_t)fMesh << 32) | (uint32_t)f, d});
    int primId = -1;
    switch (type) {
        case cAtA: primId = fv[0]; break;
        case cAtB: primId = fv[1]; break;
        case cAtC: primId = fv[2]; break;
        case cAtAB: primId = tm.face3Edges[3 * f + 0]; break;
        case cAtBC: primId = tm.face3Edges[3 * f + 1]; break;
        case cAtAC: primId = tm.face3Edges[3 * f + 2]; break;
        case cAtInterior: primId = f; break;
        default: break;
    }
    const int primType = type <= cAtC ? 0 : (type <= cAtAC ? 1 : 2);
    // de-duplication by closest primitive. NB the reference compares (primitiveType, primitiveId) only - not the mesh
    for (auto& c : res) if (c.primType == primType && c.primId == primId) return;
    if (!feasible_region(queryPt, tm, f, type, 1e-3f)) return;
    VFContact c;
    c.vMesh = qMesh; c.v = qV; c.fMesh = fMesh; c.f = f; c.closestType = type; c.primType = primType; c.primId = primId;
    c.d = d; c.cp = cp; c.bary = bary;
    res.push_back(c);
}

// triMeshEdgeContactQueryFunc (ClothContactDetector.cpp:731-813) for one candidate edge
static void ee_candidate(const ClothScene& s, int qMesh, int qE, int tMesh, int tE, std::vector<EEContact>& res, float& minDis) {
    const TriMesh& qm = s.meshes[qMesh];
    const TriMesh& tm = s.meshes[tMesh];
    const ClothEdge& Q = qm.edges[qE];
    const ClothEdge& T = tm.edges[tE];
    if (qMesh == tMesh) {
        if (tE == qE) return;
        if (Q.eV1 == T.eV1 || Q.eV1 == T.eV2 || Q.eV2 == T.eV1 || Q.eV2 == T.eV2) return;
        // triMeshEdgeContactQueryFuncWithSewing (:815-901, attached at :1011-1016): the filtered edge pairs
        if (!tm.sewingVertices.empty() && !tm.sewingEE.empty() && std::binary_search(tm.sewingEE[qE].begin(), tm.sewingEE[qE].end(), tE)) return;
    }
    const V3 p1 = qm.pos[Q.eV1], p2 = qm.pos[Q.eV2], q1 = tm.pos[T.eV1], q2 = tm.pos[T.eV2];
    V3 c1, c2; float mua, mub;
    closest_points_segments(p1, p2, q1, q2, c1, c2, mua, mub);
    const float d = norm(c1 - c2);
    minDis = std::min(d, minDis);
    if (d < s.phys.maxQueryDis) {
        bool feasible = false;
        const V3 eVec1 = p2 - p1, eVec2 = q2 - q1;
        if (dot(c2 - p1, eVec1) >= 0 && dot(p2 - c2, eVec1) >= 0) feasible = true;
        else if (dot(c1 - q1, eVec2) >= 0 && dot(q2 - c1, eVec2) >= 0) feasible = true;
        if (feasible) res.push_back({qMesh, qE, tMesh, tE, mua, mub, c1, c2, d});
    }
}

// the last loop of applyCollisionDetection (M/VBDCloth/VBDClothPhysics.cpp:775-819) for one simulated mesh: fold what the queries left
// per vertex / face / edge over each vertex's incident faces and edges -> conservative bounds, active mask (only ever raised),
// positions at this detection
void cloth_conservative_bounds(const ClothScene& s, TriMesh& m) {
    const float R = s.phys.maxQueryDis;
    for (int v = 0; v < m.nV; v++) {
        float dMin = R;
        dMin = std::min(dMin, m.vfMinDis[v]);
        bool active = !m.vf[v].empty();
        for (int i = m.vNbrFOff[v]; i < m.vNbrFOff[v + 1]; i++) {
            dMin = std::min(dMin, m.faceMinDis[m.vNbrF[i]]);
            active = active || !m.faceLinks[m.vNbrF[i]].empty();
        }
        for (int i = m.vNbrEOff[v]; i < m.vNbrEOff[v + 1]; i++) {
            dMin = std::min(dMin, m.eeMinDis[m.vNbrE[i]]);
            active = active || !m.ee[m.vNbrE[i]].empty();
        }
        dMin = std::max(0.f, dMin - s.phys.thickness);
        m.bounds[v] = dMin * s.phys.conservativeStepRelaxation;
        if (active) m.activeCollisionMask[v] = 1;
    }
    m.posAtCD = m.pos;
}

// VBDClothSimulationFramework::applyCollisionDetection (M/VBDCloth/VBDClothPhysics.cpp:735-820): VF queries for the vertices
// of ALL meshes, EE queries for the edges of simulated meshes, conservative bounds, active mask.
void cloth_detect(ClothScene& s) {
    const float R = s.phys.maxQueryDis;
    const int nM = (int)s.meshes.size();
    std::vector<int64_t> faceBase(nM + 1, 0), edgeBase(nM + 1, 0);
    for (int i = 0; i < nM; i++) { faceBase[i + 1] = faceBase[i] + s.meshes[i].nF; edgeBase[i + 1] = edgeBase[i] + s.meshes[i].nE; }
    auto mesh_of = [&](const std::vector<int64_t>& base, int gid) { int mi = 0; while (gid >= base[mi + 1]) mi++; return mi; };
    // resetFaceContactInfo (ClothContactDetector.h:144-152)
    for (auto& m : s.meshes) { m.faceMinDis.assign(m.nF, R); for (auto& l : m.faceLinks) l.clear(); }
    Grid gf, ge;
    if (!s.bruteForce) {
        double sum = 0; int64_t cnt = 0;
        for (auto& m : s.meshes) for (auto& e : m.edges) { sum += norm(m.pos[e.eV1] - m.pos[e.eV2]); cnt++; }
        const float h = std::max(R, cnt ? (float)(1.5 * sum / cnt) : R);
        gf.h = ge.h = h;
        for (int mi = 0; mi < nM; mi++) {
            const TriMesh& m = s.meshes[mi];
            for (int f = 0; f < m.nF; f++) {
                const V3 a = m.pos[m.faces[3 * f]], b = m.pos[m.faces[3 * f + 1]], c = m.pos[m.faces[3 * f + 2]];
                const V3 pad(1.01f * R, 1.01f * R, 1.01f * R);
                gf.insert((int)(faceBase[mi] + f), vmin(a, vmin(b, c)) - pad, vmax(a, vmax(b, c)) + pad);
            }
            for (int e = 0; e < m.nE; e++) {
                const V3 a = m.pos[m.edges[e].eV1], b = m.pos[m.edges[e].eV2];
                const V3 pad(1.01f * R, 1.01f * R, 1.01f * R);
                ge.insert((int)(edgeBase[mi] + e), vmin(a, b) - pad, vmax(a, b) + pad);
            }
        }
    }
    // ---- VF (contactQueryVF, ClothContactDetector.cpp:1052-1076)
    for (int mi = 0; mi < nM; mi++) {
        TriMesh& m = s.meshes[mi];
        std::vector<std::vector<std::pair<int64_t, float>>> faceDis(m.nV);
        s.pool->parallel_for(m.nV, [&](int64_t b, int64_t e) {
            std::vector<int> cand;
            for (int64_t v = b; v < e; v++) {
                m.vf[v].clear();
                float minDis = R;
                const V3 p = m.pos[v];
                if (s.bruteForce) {
                    for (int tm = 0; tm < nM; tm++) for (int f = 0; f < s.meshes[tm].nF; f++) vf_candidate(s, mi, (int)v, p, tm, f, m.vf[v], minDis, faceDis[v]);
                } else {
                    gf.query(p, p, cand);
                    for (int gid : cand) { const int tm = mesh_of(faceBase, gid); vf_candidate(s, mi, (int)v, p, tm, (int)(gid - faceBase[tm]), m.vf[v], minDis, faceDis[v]); }
                }
                m.vfMinDis[v] = minDis;
            }
        }, 64);
        // face-side effects, serially: update_min(faceMinDisToVertices) and the back-links (:181, 263-276)
        for (int v = 0; v < m.nV; v++) {
            for (auto& fd : faceDis[v]) { TriMesh& tm = s.meshes[(int)(fd.first >> 32)]; float& fm = tm.faceMinDis[(int)(fd.first & 0xFFFFFFFF)]; fm = std::min(fm, fd.second); }
            for (size_t ci = 0; ci < m.vf[v].size(); ci++) s.meshes[m.vf[v][ci].fMesh].faceLinks[m.vf[v][ci].f].push_back({mi, v, (int)ci});
        }
    }
    // ---- EE (contactQueryEE, :1106-1135): simulated meshes only
    for (int mi = 0; mi < s.numSim; mi++) {
        TriMesh& m = s.meshes[mi];
        s.pool->parallel_for(m.nE, [&](int64_t b, int64_t e) {
            std::vector<int> cand;
            for (int64_t ei = b; ei < e; ei++) {
                m.ee[ei].clear();
                float minDis = R;
                if (s.bruteForce) {
                    for (int tm = 0; tm < nM; tm++) for (int te = 0; te < s.meshes[tm].nE; te++) ee_candidate(s, mi, (int)ei, tm, te, m.ee[ei], minDis);
                } else {
                    const V3 a = m.pos[m.edges[ei].eV1], c = m.pos[m.edges[ei].eV2];
                    ge.query(vmin(a, c), vmax(a, c), cand);
                    for (int gid : cand) { const int tm = mesh_of(edgeBase, gid); ee_candidate(s, mi, (int)ei, tm, (int)(gid - edgeBase[tm]), m.ee[ei], minDis); }
                }
                m.eeMinDis[ei] = minDis;
            }
        }, 64);
    }
    // ---- conservative bounds + active mask (:775-819)
    int64_t nVF = 0, nEE = 0;
    for (auto& m : s.meshes) { for (auto& l : m.vf) nVF += l.size(); for (auto& l : m.ee) nEE += l.size(); }
    for (int mi = 0; mi < s.numSim; mi++) cloth_conservative_bounds(s, s.meshes[mi]);
    s.collisionDetectionRequired = false;
    s.stats.numberCollisionDetection++;
    s.stats.numContactsVF = nVF; s.stats.numContactsEE = nEE;
}

}  // namespace orc
