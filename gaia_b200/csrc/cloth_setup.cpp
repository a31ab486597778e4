// Host setup of the cloth path: what the reference does once per scene before the first step.
//   TriMeshFEM::initialize, applyRotationScalingTranslation, computeRestposeTrianglesFrom3D   M/TriMesh/TriMesh.cpp:15-132, 213-244
//   TriMeshTopology::initialize (edges, per-vertex lists, relevant bendings)                 M/TriMesh/TriMesh.cpp:481-753
//   MeshFrame edge creation order                                                            3P/MeshFrame2/.../TriMesh/BaseMesh.h:1484-1531
//   VBDTriMeshStVk::intializeEdgeQuadraticForms                                              M/VBDCloth/VBDTriMeshStVK.cpp:46-92
//   VBDClothSimulationFramework::initializeParallelGroups / findSmallestParallelGroup        M/VBDCloth/VBDClothPhysics.cpp:72-138
//   CuMatrix::buildOrthonormalBasis                                                          3P/CuMatrix/CuMatrix/Geometry/Geometry.h:181-206
// Compiled without FMA contraction: rest-shape quantities round like the CPU path's.
#include "cloth_scene.h"
#include <algorithm>
#include <cmath>
#include <cstring>

namespace gbc {

namespace {
struct F3 { float x, y, z; };
inline F3 f3(float x, float y, float z) { return F3{x, y, z}; }
inline F3 operator-(F3 a, F3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline F3 operator-(F3 a) { return f3(-a.x, -a.y, -a.z); }
inline F3 operator*(float s, F3 a) { return f3(a.x * s, a.y * s, a.z * s); }
inline F3 operator/(F3 a, float s) { return f3(a.x / s, a.y / s, a.z / s); }
inline float dot(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline F3 cross(F3 a, F3 b) { return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float norm(F3 a) { return std::sqrt(dot(a, a)); }
inline F3 P(const std::vector<float>& p, int v) { return f3(p[3 * (size_t)v], p[3 * (size_t)v + 1], p[3 * (size_t)v + 2]); }

void orthonormal_basis(F3 n, F3& b1, F3& b2) {
    if (n.z < 0.) {
        const float a = 1.0f / (1.0f - n.z), b = n.x * n.y * a;
        b1 = f3(1.0f - n.x * n.x * a, -b, n.x);
        b2 = f3(b, n.y * n.y * a - 1.0f, -n.y);
    } else {
        const float a = 1.0f / (1.0f + n.z), b = -n.x * n.y * a;
        b1 = f3(1.0f - n.x * n.x * a, b, -n.x);
        b2 = f3(b, 1.0f - n.y * n.y * a, -n.y);
    }
}

// Eigen::AngleAxisf::toRotationMatrix
void rotation_matrix(float angle, F3 ax, float R[3][3]) {
    const float s = std::sin(angle), c = std::cos(angle);
    const F3 sa = s * ax, ca = (1.f - c) * ax;
    float t;
    t = ca.x * ax.y; R[0][1] = t - sa.z; R[1][0] = t + sa.z;
    t = ca.x * ax.z; R[0][2] = t + sa.y; R[2][0] = t - sa.y;
    t = ca.y * ax.z; R[1][2] = t - sa.x; R[2][1] = t + sa.x;
    R[0][0] = ca.x * ax.x + c; R[1][1] = ca.y * ax.y + c; R[2][2] = ca.z * ax.z + c;
}

inline F3 raw_normal(const ClothMesh& m, int f) {
    const F3 a = P(m.pos0, m.faces[3 * (size_t)f]);
    return cross(P(m.pos0, m.faces[3 * (size_t)f + 1]) - a, P(m.pos0, m.faces[3 * (size_t)f + 2]) - a);
}
inline float cotan(F3 a, F3 b) { return dot(a, b) / norm(cross(a, b)); }

void flatten(const std::vector<std::vector<int32_t>>& src, std::vector<int32_t>& off, std::vector<int32_t>& dst) {
    off.assign(src.size() + 1, 0);
    dst.clear();
    for (size_t v = 0; v < src.size(); v++) { off[v] = (int32_t)dst.size(); dst.insert(dst.end(), src[v].begin(), src[v].end()); }
    off[src.size()] = (int32_t)dst.size();
}

// Edge ids in MeshFrame's creation order: faces in order, per face the half-edges (v2->v0), (v0->v1), (v1->v2); a half-edge
// whose opposite does not exist yet opens the next edge with eV1 = source, eV2 = target.
void build_topology(ClothMesh& m) {
    const int nV = m.nV, nF = m.nF;
    std::vector<std::vector<std::pair<int32_t, int32_t>>> halfOut(nV);   // source -> (target, edge)
    m.edges.clear();
    m.face3Edges.assign(3 * (size_t)nF, -1);
    for (int f = 0; f < nF; f++) {
        const int32_t* fv = &m.faces[3 * (size_t)f];
        for (int i = 0; i < 3; i++) {
            const int32_t to = fv[i], from = fv[(i + 2) % 3], third = fv[(i + 1) % 3];
            int32_t e = -1;
            for (auto& h : halfOut[to]) if (h.first == from) { e = h.second; break; }
            if (e >= 0) {                       // the twin exists: this face is the second side
                m.edges[6 * (size_t)e + 5] = f;
                m.edges[6 * (size_t)e + 3] = third;
            } else {
                e = (int32_t)(m.edges.size() / 6);
                const int32_t rec[6] = {from, to, third, -1, f, -1};
                m.edges.insert(m.edges.end(), rec, rec + 6);
            }
            halfOut[from].push_back({to, e});
            m.face3Edges[3 * (size_t)f + (i + 2) % 3] = e;   // half-edge v[k]->v[k+1], k = (i+2)%3, is edge slot k of the face
        }
    }
    m.nE = (int)(m.edges.size() / 6);
    std::vector<std::vector<int32_t>> nv(nV), nf(nV), nfo(nV), ne(nV), neo(nV), nb(nV), nbo(nV);
    for (int f = 0; f < nF; f++)
        for (int k = 0; k < 3; k++) { const int32_t v = m.faces[3 * (size_t)f + k]; nf[v].push_back(f); nfo[v].push_back(k); }
    for (int e = 0; e < m.nE; e++) {
        const int32_t a = m.edges[6 * (size_t)e], b = m.edges[6 * (size_t)e + 1];
        ne[a].push_back(e); neo[a].push_back(0);
        ne[b].push_back(e); neo[b].push_back(1);
        nv[a].push_back(b); nv[b].push_back(a);
    }
    // relevant bendings: interior incident edges (vertex is eV1 / eV2), then per incident face its opposite edge when
    // interior (vertex is eV12Next / eV21Next)
    for (int v = 0; v < nV; v++) {
        for (size_t i = 0; i < ne[v].size(); i++)
            if (m.edges[6 * (size_t)ne[v][i] + 5] != -1) { nb[v].push_back(ne[v][i]); nbo[v].push_back(neo[v][i]); }
        for (size_t i = 0; i < nf[v].size(); i++) {
            const int32_t f = nf[v][i], k = nfo[v][i];
            const int32_t e = m.face3Edges[3 * (size_t)f + (k + 1) % 3];
            if (m.edges[6 * (size_t)e + 5] == -1) continue;
            nb[v].push_back(e); nbo[v].push_back(m.edges[6 * (size_t)e + 4] == f ? 2 : 3);
        }
        std::sort(nv[v].begin(), nv[v].end());
    }
    std::vector<int32_t> tmp;
    flatten(nv, m.vNbrVOff, m.vNbrV);
    flatten(nf, m.vNbrFOff, m.vNbrF); flatten(nfo, tmp, m.vNbrFOrder);
    flatten(ne, m.vNbrEOff, m.vNbrE); flatten(neo, tmp, m.vNbrEOrder);
    flatten(nb, m.vBendOff, m.vBend); flatten(nbo, tmp, m.vBendOrder);
}

void build_quadratic_forms(ClothMesh& m) {
    m.Q.assign(16 * (size_t)m.nE, 0.f);
    for (int e = 0; e < m.nE; e++) {
        const int32_t* E = &m.edges[6 * (size_t)e];
        if (E[5] == -1) continue;
        const F3 x0 = P(m.pos0, E[0]), x1 = P(m.pos0, E[1]), x2 = P(m.pos0, E[2]), x3 = P(m.pos0, E[3]);
        const F3 e01 = x1 - x0, e03 = x3 - x0, e02 = x2 - x0, e12 = x2 - x1, e13 = x3 - x1;
        const float c01 = cotan(e02, e01), c02 = cotan(e01, e03), c03 = cotan(-e01, e12), c04 = cotan(-e01, e13);
        const float K0[4] = {c03 + c04, c01 + c02, -c01 - c03, -c02 - c04};
        const float A0 = norm(raw_normal(m, E[4])), A1 = norm(raw_normal(m, E[5]));
        for (int r = 0; r < 4; r++)
            for (int c = 0; c < 4; c++) m.Q[16 * (size_t)e + 4 * r + c] = 3.f * K0[r] * K0[c] / (A0 + A1);
    }
}
}  // namespace

// applyRotationScalingTranslation (M/TriMesh/TriMesh.cpp:113-132): axis-angle rotation, per-axis scale, translation
void cloth_transform_points(const gaia_cloth_object_params& obj, int32_t nV, std::vector<float>& pos) {
    const F3 rot = f3(obj.rotation[0], obj.rotation[1], obj.rotation[2]);
    const float ang = norm(rot);
    if (ang > 1e-5f) {
        float R[3][3];
        rotation_matrix(ang, rot / ang, R);
        for (int v = 0; v < nV; v++) {
            const F3 p = P(pos, v);
            for (int r = 0; r < 3; r++) pos[3 * (size_t)v + r] = R[r][0] * p.x + R[r][1] * p.y + R[r][2] * p.z;
        }
    }
    for (int v = 0; v < nV; v++)
        for (int d = 0; d < 3; d++) pos[3 * (size_t)v + d] = pos[3 * (size_t)v + d] * obj.scale[d] + obj.translation[d];
}

void cloth_mesh_init(ClothMesh& m, bool simulated, int32_t nV, int32_t nF, const float* xyz, const int32_t* faces,
                     const gaia_cloth_object_params& obj, int32_t nColors, const int32_t* colorOffsets, const int32_t* colorVerts,
                     int32_t nFixed, const int32_t* fixed) {
    m.nV = nV; m.nF = nF; m.simulated = simulated; m.obj = obj;
    m.pos0.assign(xyz, xyz + 3 * (size_t)nV);
    m.faces.assign(faces, faces + 3 * (size_t)nF);
    cloth_transform_points(obj, nV, m.pos0);
    m.DmInv.assign(4 * (size_t)nF, 0.f); m.restArea.assign(nF, 0.f); m.mass.assign(nV, 0.f);
    for (int f = 0; f < nF; f++) {
        F3 n = raw_normal(m, f);
        n = n / norm(n);
        F3 a0, a1;
        orthonormal_basis(n, a0, a1);
        const F3 o = P(m.pos0, m.faces[3 * (size_t)f]);
        const F3 d1 = P(m.pos0, m.faces[3 * (size_t)f + 1]) - o, d2 = P(m.pos0, m.faces[3 * (size_t)f + 2]) - o;
        const float D00 = dot(a0, d1), D10 = dot(a1, d1), D01 = dot(a0, d2), D11 = dot(a1, d2);
        const float det = D00 * D11 - D01 * D10;
        m.restArea[f] = det * 0.5f;
        for (int k = 0; k < 3; k++) m.mass[m.faces[3 * (size_t)f + k]] += obj.density * m.restArea[f] / 3.f;
        const float inv = 1.f / det;
        float* I = &m.DmInv[4 * (size_t)f];
        I[0] = D11 * inv; I[1] = -D10 * inv; I[2] = -D01 * inv; I[3] = D00 * inv;
    }
    build_topology(m);
    m.fixedMask.assign(nV, simulated ? 0 : 1);
    if (simulated) for (int i = 0; i < nFixed; i++) m.fixedMask[fixed[i]] = 1;
    m.colors.clear();
    if (simulated) {
        for (int c = 0; c < nColors; c++) m.colors.emplace_back(colorVerts + colorOffsets[c], colorVerts + colorOffsets[c + 1]);
        // colours large -> small (M/TriMesh/TriMesh.cpp:747-750), ties keep their order
        std::stable_sort(m.colors.begin(), m.colors.end(), [](const std::vector<int32_t>& a, const std::vector<int32_t>& b) { return a.size() > b.size(); });
    }
    build_quadratic_forms(m);
}

bool cloth_host_finalize(ClothHost& s, std::string& err) {
    size_t G = 0;
    s.numSim = 0;
    s.N = s.nF = s.nE = s.nSimV = s.nSimE = s.nSimF = 0;
    for (auto& m : s.meshes) {
        m.vBase = s.N; m.fBase = s.nF; m.eBase = s.nE;
        s.N += m.nV; s.nF += m.nF; s.nE += m.nE;
        if (m.simulated) { s.numSim++; G = std::max(G, m.colors.size()); s.nSimV += m.nV; s.nSimE += m.nE; s.nSimF += m.nF; }
    }
    if (s.N >= (int64_t)1 << 27) { err = "too many vertices for the packed contact ids"; return false; }
    // every simulated vertex must be coloured exactly once
    for (int im = 0; im < s.numSim; im++) {
        ClothMesh& m = s.meshes[im];
        std::vector<uint8_t> seen(m.nV, 0);
        size_t cnt = 0;
        for (auto& c : m.colors) for (int32_t v : c) { if (v < 0 || v >= m.nV || seen[v]) { err = "bad colouring"; return false; } seen[v] = 1; cnt++; }
        if (cnt != (size_t)m.nV) { err = "colouring does not cover every vertex"; return false; }
    }
    // parallel groups: each colour of a mesh goes to the currently smallest group (the colour is NOT removed from the
    // candidates, as in the reference's cloth framework)
    s.groups.assign(G, {});
    for (int im = 0; im < s.numSim; im++) {
        ClothMesh& m = s.meshes[im];
        const int nc = (int)m.colors.size();
        for (int c = 0; c < nc; c++) {
            int best = 0;
            for (int i = 1; i < nc; i++) if (s.groups[i].size() < s.groups[best].size()) best = i;
            for (int32_t v : m.colors[c]) { s.groups[best].push_back(im); s.groups[best].push_back(v); }
        }
    }
    // sorted order
    s.groupStart.assign(G + 1, 0);
    s.sortedToOrig.assign(s.N, 0);
    for (auto& m : s.meshes) m.sortedId.assign(m.nV, -1);
    int32_t cur = 0;
    for (size_t g = 0; g < G; g++) {
        s.groupStart[g] = cur;
        for (size_t i = 0; i < s.groups[g].size(); i += 2) {
            ClothMesh& m = s.meshes[s.groups[g][i]];
            const int32_t v = s.groups[g][i + 1];
            m.sortedId[v] = cur;
            s.sortedToOrig[cur] = (int32_t)(m.vBase + v);
            cur++;
        }
    }
    s.groupStart[G] = cur;
    for (auto& m : s.meshes)
        if (!m.simulated)
            for (int v = 0; v < m.nV; v++) { m.sortedId[v] = cur; s.sortedToOrig[cur] = (int32_t)(m.vBase + v); cur++; }
    // ELL widths + algorithmic bytes per group (cloth formula)
    s.ellF.assign(G, 0); s.ellH.assign(G, 0); s.ellFBase.assign(G, 0); s.ellHBase.assign(G, 0);
    s.groupSolved.assign(G, 0); s.groupAlgBytes.assign(G, 0);
    s.nRecF = s.nRecH = 0;
    std::vector<int32_t> stamp(s.N, -1);
    for (size_t g = 0; g < G; g++) {
        int64_t nFaceInc = 0, nHingeInc = 0, distinct = 0, solved = 0;
        for (size_t i = 0; i < s.groups[g].size(); i += 2) {
            const ClothMesh& m = s.meshes[s.groups[g][i]];
            const int32_t v = s.groups[g][i + 1];
            const int nf = m.vNbrFOff[v + 1] - m.vNbrFOff[v], nh = m.vBendOff[v + 1] - m.vBendOff[v];
            s.ellF[g] = std::max(s.ellF[g], nf); s.ellH[g] = std::max(s.ellH[g], nh);
            if (m.fixedMask[v]) continue;
            solved++; nFaceInc += nf; nHingeInc += nh;
            auto touch = [&](int32_t w) { const int32_t sw = m.sortedId[w]; if (w != v && stamp[sw] != (int32_t)g) { stamp[sw] = (int32_t)g; distinct++; } };
            for (int q = m.vNbrFOff[v]; q < m.vNbrFOff[v + 1]; q++) for (int k = 0; k < 3; k++) touch(m.faces[3 * (size_t)m.vNbrF[q] + k]);
            for (int q = m.vBendOff[v]; q < m.vBendOff[v + 1]; q++) for (int k = 0; k < 4; k++) touch(m.edges[6 * (size_t)m.vBend[q] + k]);
        }
        const int64_t n = (int64_t)s.groups[g].size() / 2;
        s.ellFBase[g] = s.nRecF; s.ellHBase[g] = s.nRecH;
        s.nRecF += n * s.ellF[g]; s.nRecH += n * s.ellH[g];
        s.groupSolved[g] = solved;
        s.groupAlgBytes[g] = 60 * solved + 32 * nFaceInc + 80 * nHingeInc + 12 * distinct;
    }
    return true;
}

}  // namespace gbc
