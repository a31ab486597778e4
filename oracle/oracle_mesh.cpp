// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// Mesh initialisation + topology + colour-group merge, restating
//   TetMeshFEM::initialize            M/TetMesh/TetMeshFEM.cpp:332-443
//   TetMeshTopology::initialize       M/TetMesh/TetMeshFEM.cpp:40-330
//   MeshFrame2 orderings              3P/MeshFrame2/.../TetMesh/basetmesh.h:600-900, 1881-1930,
//                                     SurfaceMesh/SurfaceMeshStatic.h:15-55, TriMesh/BaseMesh.h:1484-1530
//   VBDPhysics::initialize (groups)   M/VBD/VBDPhysics.cpp:99-122, 236-274
#include "vbd_oracle.h"
#include <map>
#include <cassert>

namespace orc {

static const int kTetFaces[4][3] = {{1, 2, 3}, {2, 0, 3}, {0, 1, 3}, {1, 0, 2}};  // TetMeshFEM.cpp:22

// Eigen::AngleAxisf::toRotationMatrix (Rodrigues), float
static void angle_axis_to_matrix(float angle, V3 axis, M3& R) {
    float s = std::sin(angle), c = std::cos(angle);
    V3 sin_axis = s * axis;
    V3 cos1_axis = (1.f - c) * axis;
    float tmp;
    tmp = cos1_axis.x * axis.y; R(0, 1) = tmp - sin_axis.z; R(1, 0) = tmp + sin_axis.z;
    tmp = cos1_axis.x * axis.z; R(0, 2) = tmp + sin_axis.y; R(2, 0) = tmp - sin_axis.y;
    tmp = cos1_axis.y * axis.z; R(1, 2) = tmp - sin_axis.x; R(2, 1) = tmp + sin_axis.x;
    R(0, 0) = cos1_axis.x * axis.x + c;
    R(1, 1) = cos1_axis.y * axis.y + c;
    R(2, 2) = cos1_axis.z * axis.z + c;
}

// Eigen 3x3 determinant (bruteforce_det3_helper)
static inline float det3(const M3& m) {
    auto h = [&](int a, int b, int c) { return m(0, a) * (m(1, b) * m(2, c) - m(1, c) * m(2, b)); };
    return h(0, 1, 2) - h(1, 0, 2) + h(2, 0, 1);
}
// Eigen 3x3 inverse (compute_inverse_size3_helper: cofactors * 1/det)
static inline void inv3(const M3& m, M3& r) {
    auto cof = [&](int i, int j) {
        int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
        return m(i1, j1) * m(i2, j2) - m(i1, j2) * m(i2, j1);
    };
    float c00 = cof(0, 0), c10 = cof(1, 0), c20 = cof(2, 0);
    float det = c00 * m(0, 0) + c10 * m(1, 0) + c20 * m(2, 0);
    float invdet = 1.f / det;
    r(0, 0) = c00 * invdet; r(0, 1) = c10 * invdet; r(0, 2) = c20 * invdet;
    r(1, 0) = cof(0, 1) * invdet; r(1, 1) = cof(1, 1) * invdet; r(1, 2) = cof(2, 1) * invdet;
    r(2, 0) = cof(0, 2) * invdet; r(2, 1) = cof(1, 2) * invdet; r(2, 2) = cof(2, 2) * invdet;
}

static void build_topology(Mesh& m) {
    const int nV = m.nV, nT = m.nT;
    // --- CSR vertex -> incident tets: MeshFrame2 keeps per-vertex tvertex lists in tet creation order
    //     (basetmesh.h:779-800), so incident tets are ascending in tet id.
    m.vNbrInfos.assign(2 * nV, 0);
    std::vector<int> cnt(nV, 0);
    for (int t = 0; t < nT; t++) for (int k = 0; k < 4; k++) cnt[m.tets[4 * t + k]]++;
    int run = 0;
    for (int v = 0; v < nV; v++) { m.vNbrInfos[2 * v] = run; m.vNbrInfos[2 * v + 1] = cnt[v]; run += cnt[v]; }
    m.vNbrTets.assign(run, 0); m.vNbrOrder.assign(run, 0);
    std::vector<int> fill(nV, 0);
    for (int t = 0; t < nT; t++) for (int k = 0; k < 4; k++) {
        int v = m.tets[4 * t + k];
        int at = m.vNbrInfos[2 * v] + fill[v]++;
        m.vNbrTets[at] = t;
        // vertex order = first k with tetVIds(k,t)==v (TetMeshFEM.cpp:299-306)
        int ord = -1;
        for (int q = 0; q < 4; q++) if (m.tets[4 * t + q] == v) { ord = q; break; }
        m.vNbrOrder[at] = ord;
    }
    // --- half-faces; face i of a tet is opposite vertex i with vertex order kTetFaces[i]
    struct HF { int a, b, c, t, i; };
    std::vector<HF> hfs(4 * (size_t)nT);
    for (int t = 0; t < nT; t++) for (int i = 0; i < 4; i++) {
        int k[3] = {m.tets[4 * t + kTetFaces[i][0]], m.tets[4 * t + kTetFaces[i][1]], m.tets[4 * t + kTetFaces[i][2]]};
        std::sort(k, k + 3);
        hfs[4 * (size_t)t + i] = {k[0], k[1], k[2], t, i};
    }
    std::vector<size_t> ord(hfs.size());
    for (size_t q = 0; q < ord.size(); q++) ord[q] = q;
    std::sort(ord.begin(), ord.end(), [&](size_t x, size_t y) {
        const HF& A = hfs[x]; const HF& B = hfs[y];
        if (A.a != B.a) return A.a < B.a;
        if (A.b != B.b) return A.b < B.b;
        if (A.c != B.c) return A.c < B.c;
        return x < y;
    });
    m.tetNbrTets.assign(4 * (size_t)nT, -1);
    std::vector<uint8_t> isBoundary(hfs.size(), 1);
    for (size_t q = 0; q + 1 < ord.size(); q++) {
        const HF& A = hfs[ord[q]]; const HF& B = hfs[ord[q + 1]];
        if (A.a == B.a && A.b == B.b && A.c == B.c) {
            m.tetNbrTets[4 * (size_t)A.t + A.i] = B.t;
            m.tetNbrTets[4 * (size_t)B.t + B.i] = A.t;
            isBoundary[ord[q]] = 0; isBoundary[ord[q + 1]] = 0;
        }
    }
    m.tetXorSums.resize(nT);
    for (int t = 0; t < nT; t++) m.tetXorSums[t] = m.tets[4 * t] ^ m.tets[4 * t + 1] ^ m.tets[4 * t + 2] ^ m.tets[4 * t + 3];
    // --- surface: boundary half-faces in creation order (tet asc, local face asc); surface vertices in
    //     ascending tet-mesh vertex id (std::set with id compare, basetmesh.h:1887-1923)
    m.surfFaces.clear(); m.surfFaceTet.clear(); m.surfFaceIdAtTet.clear();
    std::vector<uint8_t> onSurf(nV, 0);
    for (int t = 0; t < nT; t++) for (int i = 0; i < 4; i++) {
        if (!isBoundary[4 * (size_t)t + i]) continue;
        int p0 = m.tets[4 * t + kTetFaces[i][0]], p1 = m.tets[4 * t + kTetFaces[i][1]], p2 = m.tets[4 * t + kTetFaces[i][2]];
        // trimesh createFace: half-edge 0 has source pVs[2]; FHEIterator lists sources => (p2, p0, p1)
        m.surfFaces.push_back(p2); m.surfFaces.push_back(p0); m.surfFaces.push_back(p1);
        m.surfFaceTet.push_back(t);
        m.surfFaceIdAtTet.push_back(i);  // the tet vertex not on the face (TetMeshFEM.cpp:196-213)
        onSurf[p0] = onSurf[p1] = onSurf[p2] = 1;
    }
    m.surfaceVIds.clear(); m.tetVertToSurf.assign(nV, -1);
    for (int v = 0; v < nV; v++) if (onSurf[v]) { m.tetVertToSurf[v] = (int)m.surfaceVIds.size(); m.surfaceVIds.push_back(v); }
    const int nF = m.nSurfF();
    // neighbour face across directed edge (fv[k] -> fv[k+1]) is the face holding the reversed half-edge
    std::map<std::pair<int, int>, int> he2face;
    for (int f = 0; f < nF; f++) for (int k = 0; k < 3; k++)
        he2face[{m.surfFaces[3 * f + k], m.surfFaces[3 * f + (k + 1) % 3]}] = f;
    m.surfFace3Nbr.assign(3 * (size_t)nF, -1);
    for (int f = 0; f < nF; f++) for (int k = 0; k < 3; k++) {
        auto it = he2face.find({m.surfFaces[3 * f + (k + 1) % 3], m.surfFaces[3 * f + k]});
        if (it != he2face.end()) m.surfFace3Nbr[3 * f + k] = it->second;
    }
    m.surfVertNbrSurfVerts.assign(m.nSurfV(), {}); m.surfVertNbrSurfFaces.assign(m.nSurfV(), {});
    for (int f = 0; f < nF; f++) for (int k = 0; k < 3; k++) {
        int sv = m.tetVertToSurf[m.surfFaces[3 * f + k]];
        m.surfVertNbrSurfFaces[sv].push_back(f);
        for (int q = 1; q < 3; q++) {
            int o = m.surfFaces[3 * f + (k + q) % 3];
            auto& L = m.surfVertNbrSurfVerts[sv];
            if (std::find(L.begin(), L.end(), o) == L.end()) L.push_back(o);
        }
    }
    for (auto& L : m.surfVertNbrSurfVerts) std::sort(L.begin(), L.end());
}

void mesh_initialize(Mesh& m, int nV, int nT, const float* xyz, const int32_t* tets, const gaia_object_params& obj,
                     int nColors, const int32_t* colorOffsets, const int32_t* colorVerts, int nFixed, const int32_t* fixed,
                     bool keepColourOrder) {
    m.nV = nV; m.nT = nT; m.obj = obj;
    m.pos.resize(nV);
    for (int v = 0; v < nV; v++) m.pos[v] = V3(xyz[3 * v], xyz[3 * v + 1], xyz[3 * v + 2]);
    m.tets.assign(tets, tets + 4 * (size_t)nT);
    build_topology(m);
    // colours: size-descending (TetMeshFEM.cpp:122-133). std::sort is unstable on ties; the documented rule
    // here is a STABLE sort (ties keep file order), mirrored by the product.
    m.colors.clear();
    for (int c = 0; c < nColors; c++)
        m.colors.emplace_back(colorVerts + colorOffsets[c], colorVerts + colorOffsets[c + 1]);
    if (!keepColourOrder)
        std::stable_sort(m.colors.begin(), m.colors.end(),
                         [](const std::vector<int>& a, const std::vector<int>& b) { return a.size() > b.size(); });
    // rotation -> scale -> translation (TetMeshFEM.cpp:369-384)
    V3 rot(obj.rotation[0], obj.rotation[1], obj.rotation[2]);
    float ang = norm(rot);
    if (ang > 0) {
        M3 R; angle_axis_to_matrix(ang, rot / ang, R);
        for (int v = 0; v < nV; v++) {
            V3 p = m.pos[v];
            m.pos[v] = V3(R(0, 0) * p.x + R(0, 1) * p.y + R(0, 2) * p.z,
                          R(1, 0) * p.x + R(1, 1) * p.y + R(1, 2) * p.z,
                          R(2, 0) * p.x + R(2, 1) * p.y + R(2, 2) * p.z);
        }
    }
    for (int v = 0; v < nV; v++) {
        V3& p = m.pos[v];
        p.x *= obj.scale[0]; p.y *= obj.scale[1]; p.z *= obj.scale[2];
        p.x += obj.translation[0]; p.y += obj.translation[1]; p.z += obj.translation[2];
    }
    m.vel.assign(nV, V3(obj.initialVelocity[0], obj.initialVelocity[1], obj.initialVelocity[2]));
    m.velPrev = m.vel;   // VBDBaseTetMesh::initialize: mVelocitiesPrev = mVelocity (VBD_BaseMaterial.h:333-334)
    m.prev.assign(nV, V3());
    m.inertia.assign(nV, V3());
    m.accel.assign(nV, V3());
    m.mass.assign(nV, 0.f);
    m.restVol.resize(nT); m.DmInv.resize(9 * (size_t)nT);
    for (int t = 0; t < nT; t++) {
        M3 Dm, Di;
        V3 v0 = m.pos[m.tets[4 * t]];
        for (int c = 0; c < 3; c++) {
            V3 d = m.pos[m.tets[4 * t + c + 1]] - v0;
            Dm(0, c) = d.x; Dm(1, c) = d.y; Dm(2, c) = d.z;
        }
        inv3(Dm, Di);
        for (int q = 0; q < 9; q++) m.DmInv[9 * (size_t)t + q] = Di.m[q];
        float vol = det3(Dm) / 6;
        m.restVol[t] = vol;
        // vertexMass += 0.25 * vol * density, with 0.25 a double literal (TetMeshFEM.cpp:426)
        for (int k = 0; k < 4; k++) {
            int v = m.tets[4 * t + k];
            m.mass[v] = (float)((double)m.mass[v] + 0.25 * (double)vol * (double)obj.density);
        }
    }
    m.fixedMask.assign(nV, 0);
    m.fixedPoints.assign(fixed, fixed + nFixed);
    for (int i = 0; i < nFixed; i++) m.fixedMask[fixed[i]] = 1;
    m.activeCollisionMask.assign(nV, 0); m.penetratedMask.assign(nV, 0);
    m.colResults.assign(m.nSurfV(), ColResult());
    m.colRelations.assign(nV, {});
    m.vertexParallelGroup.assign(nV, -1);
    // disableModelsLatterToAppear (BasePhysicsFramework.cpp:211-223)
    if (obj.frameToAppear > 0) { m.activeForCollision = false; m.activeForMaterialSolve = false; }
}

void scene_finalize(Scene& s) {
    // updateWorldBox (BasePhysicsFramework.cpp:73-98)
    if (s.phys.checkAndUpdateWorldBounds) {
        for (auto& m : s.meshes) for (int d = 0; d < 3; d++) {
            float mn = 1e30f, mx = -1e30f;
            for (int v = 0; v < m.nV; v++) { mn = std::min(mn, m.pos[v][d]); mx = std::max(mx, m.pos[v][d]); }
            if (m.nV && s.phys.worldBoundsMin[d] > mn) s.phys.worldBoundsMin[d] = mn;
            if (m.nV && s.phys.worldBoundsMax[d] < mx) s.phys.worldBoundsMax[d] = mx;
        }
    }
    // colour-group merge (VBDPhysics.cpp:236-274)
    size_t nGroups = 0;
    for (auto& m : s.meshes) nGroups = std::max(nGroups, m.colors.size());
    s.vertexParallelGroups.assign(nGroups, {});
    for (size_t iMesh = 0; iMesh < s.meshes.size(); iMesh++) {
        Mesh& m = s.meshes[iMesh];
        std::vector<int> available(m.colors.size());
        for (size_t c = 0; c < available.size(); c++) available[c] = (int)c;
        for (size_t c = 0; c < m.colors.size(); c++) {
            // findSmallestParallelGroup(available, groups, remove=true) (VBDPhysics.cpp:99-122)
            size_t best = s.vertexParallelGroups[available[0]].size();
            int bestId = available[0]; size_t bestI = 0;
            // (a partition keeps the global group of each colour = the colour's index: oracle_set_partition_mode)
            for (size_t i = 1; i < available.size() && !s.keepColourOrder; i++)
                if (s.vertexParallelGroups[available[i]].size() < best) {
                    best = s.vertexParallelGroups[available[i]].size(); bestId = available[i]; bestI = i;
                }
            available.erase(available.begin() + bestI);
            for (int v : m.colors[c]) {
                s.vertexParallelGroups[bestId].push_back((int)iMesh);
                s.vertexParallelGroups[bestId].push_back(v);
                m.vertexParallelGroup[v] = bestId;
            }
        }
    }
    s.activeCollisionsEachGroup.assign(nGroups, {});
    collision_initialize(s);
    s.finalized = true;
}

Scene::~Scene() {
    collision_free(*this);
    delete pool;
    delete trace;
}

}  // namespace orc
