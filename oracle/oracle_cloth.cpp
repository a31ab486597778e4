// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h, oracle_cloth.h).
// CPU restatement of the reference VBD cloth step. Every function cites the reference file:line it follows
// (paths relative to Simulator/Modules/ = "M/", Simulator/3rdParty/ = "3P/").
//
// Canonical rules where the reference's outcome depends on Embree's traversal order or on a data race
// (the CUDA path applies the same rules):
//   R5  contact candidates are visited in ascending (mesh, primitive id): the face that represents a shared closest
//       primitive (ClothContactDetector.cpp:186-202) is the feasible one with the smallest (faceMesh, faceId); contact
//       lists are therefore sorted.
//   R6  `collisionDetectionRequired = truncationHappened` (VBDClothPhysics.cpp:1147, 909) is a last-writer-wins race in
//       the reference; here it is the OR over all vertices since the last detection (the stated intent, :449-450).
//   R7  per-vertex neighbour lists are walked in ascending id (the reference walks MeshFrame's half-edge lists): the
//       sums are the same sets in a different floating-point order.
#include "oracle_cloth.h"

namespace orc {

static const float CMP_EPSILON = 1e-5f;    // M/Types/Types.h:20
static const float CMP_EPSILON2 = 1e-10f;  // M/Types/Types.h:21
static const int GRAVITY_AXIS = 1;         // M/Types/Types.h:9

// Eigen::AngleAxisf::toRotationMatrix (Rodrigues), float
static void cloth_angle_axis(float angle, V3 axis, M3& R) {
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

// CuMatrix::buildOrthonormalBasis (3P/CuMatrix/CuMatrix/Geometry/Geometry.h:181-206)
static void build_orthonormal_basis(V3 n, V3& b1, V3& b2) {
    if (n.z < 0.) {
        const float a = 1.0f / (1.0f - n.z);
        const float b = n.x * n.y * a;
        b1 = V3(1.0f - n.x * n.x * a, -b, n.x);
        b2 = V3(b, n.y * n.y * a - 1.0f, -n.y);
    } else {
        const float a = 1.0f / (1.0f + n.z);
        const float b = -n.x * n.y * a;
        b1 = V3(1.0f - n.x * n.x * a, b, -n.x);
        b2 = V3(b, 1.0f - n.y * n.y * a, -n.y);
    }
}

static inline V3 face_normal_raw(const TriMesh& m, int f) {   // TriMeshFEM::computeNormal(f, false) (M/TriMesh/TriMesh.h:677-690)
    const V3 a = m.pos[m.faces[3 * f]];
    return cross(m.pos[m.faces[3 * f + 1]] - a, m.pos[m.faces[3 * f + 2]] - a);
}

// TriMeshTopology::initialize (M/TriMesh/TriMesh.cpp:481-753). Edge ids follow MeshFrame's creation order
// (3P/MeshFrame2/.../TriMesh/BaseMesh.h:1484-1531): faces in order; per face the half-edges (v2->v0), (v0->v1), (v1->v2);
// a half-edge whose opposite does not exist yet creates the next edge, eV1 = its source, eV2 = its target.
static void build_cloth_topology(TriMesh& m) {
    const int nV = m.nV, nF = m.nF;
    std::vector<std::vector<std::pair<int, int>>> out(nV);   // per source vertex: (target, edge id)
    m.edges.clear();
    m.face3Edges.assign(3 * nF, -1);
    m.face3NbrFaces.assign(3 * nF, -1);
    auto find = [&](int s, int t) { for (auto& pr : out[s]) if (pr.first == t) return pr.second; return -1; };
    for (int f = 0; f < nF; f++) {
        const int* fv = &m.faces[3 * f];
        for (int i = 0; i < 3; i++) {
            const int tgt = fv[i], src = fv[(i + 2) % 3];
            int e = find(tgt, src);
            if (e >= 0) {   // second side of an existing edge (orientation eV2 -> eV1)
                m.edges[e].fId2 = f;
                m.edges[e].eV21Next = fv[(i + 1) % 3];
            } else {
                e = (int)m.edges.size();
                m.edges.push_back({src, tgt, fv[(i + 1) % 3], -1, f, -1});
            }
            out[src].push_back({tgt, e});
            // the half-edge src->tgt is edge slot (i+2)%3 of the face: (v[k] -> v[k+1]) with k = (i+2)%3
            m.face3Edges[3 * f + (i + 2) % 3] = e;
        }
    }
    m.nE = (int)m.edges.size();
    for (int f = 0; f < nF; f++)
        for (int k = 0; k < 3; k++) {
            const ClothEdge& e = m.edges[m.face3Edges[3 * f + k]];
            m.face3NbrFaces[3 * f + k] = (e.fId1 == f) ? e.fId2 : e.fId1;
        }
    // per-vertex lists (R7: ascending ids)
    std::vector<std::vector<int>> nv(nV), nf(nV), nfo(nV), ne(nV), neo(nV), nb(nV), nbo(nV);
    for (int f = 0; f < nF; f++)
        for (int k = 0; k < 3; k++) { nf[m.faces[3 * f + k]].push_back(f); nfo[m.faces[3 * f + k]].push_back(k); }
    for (int e = 0; e < m.nE; e++) {
        const ClothEdge& E = m.edges[e];
        ne[E.eV1].push_back(e); neo[E.eV1].push_back(0);
        ne[E.eV2].push_back(e); neo[E.eV2].push_back(1);
        nv[E.eV1].push_back(E.eV2); nv[E.eV2].push_back(E.eV1);
    }
    // relevant bendings (:664-722): non-boundary incident edges (order 0/1), then per incident face the opposite edge
    // when it is not a boundary edge (order 2: vertex is eV12Next, 3: eV21Next)
    for (int v = 0; v < nV; v++) {
        for (size_t i = 0; i < ne[v].size(); i++)
            if (m.edges[ne[v][i]].fId2 != -1) { nb[v].push_back(ne[v][i]); nbo[v].push_back(neo[v][i]); }
        for (size_t i = 0; i < nf[v].size(); i++) {
            const int f = nf[v][i], k = nfo[v][i];
            const int e = m.face3Edges[3 * f + (k + 1) % 3];     // edge (v[k+1] -> v[k+2]) is opposite to v[k]
            if (m.edges[e].fId2 == -1) continue;
            nb[v].push_back(e); nbo[v].push_back(m.edges[e].fId1 == f ? 2 : 3);
        }
    }
    auto pack = [&](std::vector<std::vector<int>>& src, std::vector<int>& off, std::vector<int>& dst) {
        off.assign(nV + 1, 0); dst.clear();
        for (int v = 0; v < nV; v++) { off[v] = (int)dst.size(); dst.insert(dst.end(), src[v].begin(), src[v].end()); }
        off[nV] = (int)dst.size();
    };
    for (int v = 0; v < nV; v++) std::sort(nv[v].begin(), nv[v].end());
    pack(nv, m.vNbrVOff, m.vNbrV);
    pack(nf, m.vNbrFOff, m.vNbrF); { std::vector<int> o; pack(nfo, o, m.vNbrFOrder); }
    pack(ne, m.vNbrEOff, m.vNbrE); { std::vector<int> o; pack(neo, o, m.vNbrEOrder); }
    pack(nb, m.vBendOff, m.vBend); { std::vector<int> o; pack(nbo, o, m.vBendOrder); }
}

static inline float cotangent(V3 v1, V3 v2) { return dot(v1, v2) / norm(cross(v1, v2)); }   // VBDTriMeshStVK.cpp:46-48

// VBDTriMeshStVk::intializeEdgeQuadraticForms (M/VBDCloth/VBDTriMeshStVK.cpp:50-92)
static void build_bending_forms(TriMesh& m) {
    m.Q.assign(m.nE, std::array<float, 16>{});
    for (int e = 0; e < m.nE; e++) {
        const ClothEdge& E = m.edges[e];
        if (E.fId2 == -1) continue;
        const V3 x0 = m.pos[E.eV1], x1 = m.pos[E.eV2], x2 = m.pos[E.eV12Next], x3 = m.pos[E.eV21Next];
        const V3 e01 = x1 - x0, e03 = x3 - x0, e02 = x2 - x0, e12 = x2 - x1, e13 = x3 - x1;
        const float c01 = cotangent(e02, e01), c02 = cotangent(e01, e03), c03 = cotangent(-e01, e12), c04 = cotangent(-e01, e13);
        const float K0[4] = {c03 + c04, c01 + c02, -c01 - c03, -c02 - c04};
        const float A0 = norm(face_normal_raw(m, E.fId1)), A1 = norm(face_normal_raw(m, E.fId2));
        for (int r = 0; r < 4; r++)
            for (int c = 0; c < 4; c++) m.Q[e][4 * r + c] = 3.f * K0[r] * K0[c] / (A0 + A1);
    }
}

// TriMeshFEM::initialize + applyRotationScalingTranslation + computeRestposeTrianglesFrom3D (M/TriMesh/TriMesh.cpp:15-132,
// 213-244), VBDBaseTriMesh::initialize (M/VBDCloth/VBDBaseTriMesh.cpp:28-58), VBDTriMeshStVk::initialize
// applyRotationScalingTranslation (M/TriMesh/TriMesh.cpp:113-132): rotation (axis-angle), scale, translation
void cloth_transform_points(const gaia_cloth_object_params& obj, std::vector<V3>& pos) {
    const V3 rot(obj.rotation[0], obj.rotation[1], obj.rotation[2]);
    const float ang = norm(rot);
    if (ang > CMP_EPSILON) {
        M3 R; cloth_angle_axis(ang, rot / ang, R);
        for (auto& p : pos) p = V3(R(0, 0) * p.x + R(0, 1) * p.y + R(0, 2) * p.z, R(1, 0) * p.x + R(1, 1) * p.y + R(1, 2) * p.z,
                                   R(2, 0) * p.x + R(2, 1) * p.y + R(2, 2) * p.z);
    }
    for (auto& p : pos) p = V3(p.x * obj.scale[0] + obj.translation[0], p.y * obj.scale[1] + obj.translation[1], p.z * obj.scale[2] + obj.translation[2]);
}

// ColliderTrimeshSequence::initialize (M/SpatialQuery/ColiiderTriMeshSequence.h:131-144)
void cloth_set_collider_sequence(ClothScene& s, int mesh, int nKeys, const int32_t* keyFrames, const float* xyz, bool interpolateSubstep, int interpolateIter) {
    const TriMesh& m = s.meshes[mesh];
    ColliderSequence c;
    c.mesh = mesh;
    for (int i = 0; i < nKeys; i++) c.keyFrames.push_back(keyFrames ? keyFrames[i] : i);
    c.interpolateIter = interpolateIter;
    c.interpolateSubstep = interpolateSubstep || interpolateIter > 1;
    for (int k = 0; k < nKeys; k++) {
        std::vector<V3> p(m.nV);
        for (int v = 0; v < m.nV; v++) p[v] = V3(xyz[((size_t)k * m.nV + v) * 3], xyz[((size_t)k * m.nV + v) * 3 + 1], xyz[((size_t)k * m.nV + v) * 3 + 2]);
        cloth_transform_points(m.obj, p);
        c.keyPos.push_back(p);
    }
    s.colliderSeqs.push_back(c);
}

void cloth_mesh_initialize(TriMesh& m, bool simulated, int nV, int nF, const float* xyz, const int32_t* faces,
                           const gaia_cloth_object_params& obj, int nColors, const int32_t* colorOffsets, const int32_t* colorVerts,
                           int nFixed, const int32_t* fixed) {
    m.nV = nV; m.nF = nF; m.simulated = simulated; m.obj = obj;
    m.pos.resize(nV);
    for (int i = 0; i < nV; i++) m.pos[i] = V3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    m.faces.assign(faces, faces + 3 * nF);
    cloth_transform_points(obj, m.pos);
    m.prev.assign(nV, V3()); m.vel.assign(nV, V3()); m.velPrev.assign(nV, V3()); m.inertia.assign(nV, V3());
    m.accel.assign(nV, V3()); m.next = m.pos; m.posAtCD.assign(nV, V3()); m.prevIter = m.pos;
    // rest pose from the 3-D triangles
    m.DmInv.assign(4 * (size_t)nF, 0.f); m.restArea.assign(nF, 0.f); m.mass.assign(nV, 0.f);
    for (int f = 0; f < nF; f++) {
        V3 n = face_normal_raw(m, f);
        n = n / norm(n);
        V3 a0, a1; build_orthonormal_basis(n, a0, a1);
        const V3 v1 = m.pos[m.faces[3 * f]];
        const V3 d1 = m.pos[m.faces[3 * f + 1]] - v1, d2 = m.pos[m.faces[3 * f + 2]] - v1;
        // Dm = [axesT*d1, axesT*d2] (2x2, column-major)
        const float D00 = dot(a0, d1), D10 = dot(a1, d1), D01 = dot(a0, d2), D11 = dot(a1, d2);
        const float det = D00 * D11 - D01 * D10;            // Eigen 2x2 determinant
        m.restArea[f] = det * 0.5f;
        for (int k = 0; k < 3; k++) m.mass[m.faces[3 * f + k]] += obj.density * m.restArea[f] / 3.f;
        const float invdet = 1.f / det;                     // Eigen compute_inverse_size2_helper
        float* I = &m.DmInv[4 * (size_t)f];
        I[0] = D11 * invdet; I[1] = -D10 * invdet; I[2] = -D01 * invdet; I[3] = D00 * invdet;   // (0,0),(1,0),(0,1),(1,1)
    }
    build_cloth_topology(m);
    m.fixedMask.assign(nV, 0);
    m.fixedPoints.assign(fixed, fixed + nFixed);
    for (int i = 0; i < nFixed; i++) m.fixedMask[fixed[i]] = 1;
    if (!simulated) { m.fixedMask.assign(nV, 1); m.fixedPoints.resize(nV); for (int i = 0; i < nV; i++) m.fixedPoints[i] = i; }
    m.globalColor.assign(nV, -1);
    m.colors.clear();
    if (simulated) {
        for (int c = 0; c < nColors; c++) m.colors.emplace_back(colorVerts + colorOffsets[c], colorVerts + colorOffsets[c + 1]);
        // "we sort them from large to small" (M/TriMesh/TriMesh.cpp:747-750; std::sort is unstable on ties: stable here, as for tets)
        std::stable_sort(m.colors.begin(), m.colors.end(), [](const std::vector<int>& a, const std::vector<int>& b) { return a.size() > b.size(); });
    }
    build_bending_forms(m);
    m.bounds.assign(nV, 0.f);
    m.activeCollisionMask.assign(nV, 0);
    m.vf.assign(nV, {}); m.vfMinDis.assign(nV, 0.f);
    m.ee.assign(simulated ? m.nE : 0, {}); m.eeMinDis.assign(simulated ? m.nE : 0, 0.f);
    m.faceMinDis.assign(nF, 0.f); m.faceLinks.assign(nF, {});
}

// VBDClothSimulationFramework::initializeParallelGroups + findSmallestParallelGroup (M/VBDCloth/VBDClothPhysics.cpp:72-138).
// NB: unlike the tet framework the colour is NOT removed from availableColors (removeSmallestAvailableColor = false), so two
// colours of one mesh may land in the same group: restated as is.
void cloth_finalize(ClothScene& s) {
    size_t nGroups = 0;
    s.numSim = 0;
    for (auto& m : s.meshes) if (m.simulated) { s.numSim++; nGroups = std::max(nGroups, m.colors.size()); }
    s.groups.assign(nGroups, {});
    for (int im = 0; im < s.numSim; im++) {
        TriMesh& m = s.meshes[im];
        const int numColors = (int)m.colors.size();
        std::vector<int> available(numColors);
        for (int c = 0; c < numColors; c++) available[c] = c;
        for (int c = 0; c < numColors; c++) {
            size_t best = s.groups[available[0]].size(); int bestId = available[0];
            for (int i = 1; i < numColors; i++)
                if (s.groups[available[i]].size() < best) { best = s.groups[available[i]].size(); bestId = available[i]; }
            for (int v : m.colors[c]) { s.groups[bestId].push_back(im); s.groups[bestId].push_back(v); m.globalColor[v] = bestId; }
        }
    }
    s.finalized = true;
}

}  // namespace orc
