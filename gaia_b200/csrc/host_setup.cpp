// Host-side mesh preparation of libgaia_b200.so: everything VBDPhysics::initialize does before the first step,
// re-derived for a colour-sorted SoA device layout.
//   transform / rest volume / lumped mass      TetMeshFEM::initialize, M/TetMesh/TetMeshFEM.cpp:369-429
//   surface vertex / face orderings            MeshFrame2 basetmesh.h:1881-1930, SurfaceMeshStatic.h:15-55,
//                                              TriMesh/BaseMesh.h:1484-1530; TetMeshTopology::initialize TetMeshFEM.cpp:136-221
//   colour sort + group merge                  TetMeshFEM.cpp:122-133, M/VBD/VBDPhysics.cpp:99-122, 236-274
//   world box update                           BasePhysicFramework::updateWorldBox, BasePhysicsFramework.cpp:73-98
#include "scene_impl.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>

namespace gb {

static const int kFaceOf[4][3] = {{1, 2, 3}, {2, 0, 3}, {0, 1, 3}, {1, 0, 2}};   // face i is opposite vertex i

// ---- float helpers written to round exactly like the reference's Eigen expressions (no FMA: -ffp-contract=off)
static void rodrigues(const float rot[3], float R[9] /*row-major*/) {
    const float ang = std::sqrt(rot[0] * rot[0] + rot[1] * rot[1] + rot[2] * rot[2]);
    const float ax[3] = {rot[0] / ang, rot[1] / ang, rot[2] / ang};
    const float s = std::sin(ang), c = std::cos(ang);
    const float sx = s * ax[0], sy = s * ax[1], sz = s * ax[2];
    const float cx = (1.f - c) * ax[0], cy = (1.f - c) * ax[1], cz = (1.f - c) * ax[2];
    float t;
    t = cx * ax[1]; R[1] = t - sz; R[3] = t + sz;
    t = cx * ax[2]; R[2] = t + sy; R[6] = t - sy;
    t = cy * ax[2]; R[5] = t - sx; R[7] = t + sx;
    R[0] = cx * ax[0] + c; R[4] = cy * ax[1] + c; R[8] = cz * ax[2] + c;
}

static inline float det3_cols(const float a[3], const float b[3], const float c[3]) {
    // Eigen bruteforce_det3 on the matrix with columns a,b,c
    auto h = [&](const float* p, const float* q, const float* r) { return p[0] * (q[1] * r[2] - r[1] * q[2]); };
    return h(a, b, c) - h(b, a, c) + h(c, a, b);
}

static void surface_topology(HostMesh& m) {
    const int nV = m.nV, nT = m.nT;
    // bucket the 4*nT half-faces by their smallest vertex, then match duplicates inside each bucket
    struct Half { int32_t lo, mid, hi, tet, face; };
    std::vector<int32_t> bucketOff(nV + 1, 0);
    std::vector<Half> halves(4 * (size_t)nT);
    for (int t = 0; t < nT; t++)
        for (int i = 0; i < 4; i++) {
            int32_t k0 = m.tets[4 * t + kFaceOf[i][0]], k1 = m.tets[4 * t + kFaceOf[i][1]], k2 = m.tets[4 * t + kFaceOf[i][2]];
            if (k0 > k1) std::swap(k0, k1);
            if (k1 > k2) std::swap(k1, k2);
            if (k0 > k1) std::swap(k0, k1);
            halves[4 * (size_t)t + i] = {k0, k1, k2, t, i};
            bucketOff[k0 + 1]++;
        }
    for (int v = 0; v < nV; v++) bucketOff[v + 1] += bucketOff[v];
    std::vector<int32_t> slot(halves.size());
    {
        std::vector<int32_t> cur(bucketOff.begin(), bucketOff.end() - 1);
        for (size_t q = 0; q < halves.size(); q++) slot[cur[halves[q].lo]++] = (int32_t)q;
    }
    m.tetNbrTets.assign(4 * (size_t)nT, -1);
    for (int v = 0; v < nV; v++)
        for (int x = bucketOff[v]; x < bucketOff[v + 1]; x++) {
            const Half& A = halves[slot[x]];
            if (m.tetNbrTets[4 * (size_t)A.tet + A.face] != -1) continue;
            for (int y = x + 1; y < bucketOff[v + 1]; y++) {
                const Half& B = halves[slot[y]];
                if (A.mid == B.mid && A.hi == B.hi) {
                    m.tetNbrTets[4 * (size_t)A.tet + A.face] = B.tet;
                    m.tetNbrTets[4 * (size_t)B.tet + B.face] = A.tet;
                    break;
                }
            }
        }
    // boundary half-faces in creation order; face vertices rotate to (p2,p0,p1) (trimesh createFace + FHEIterator)
    std::vector<uint8_t> onSurface(nV, 0);
    for (int t = 0; t < nT; t++)
        for (int i = 0; i < 4; i++) {
            if (m.tetNbrTets[4 * (size_t)t + i] != -1) continue;
            const int32_t p0 = m.tets[4 * t + kFaceOf[i][0]], p1 = m.tets[4 * t + kFaceOf[i][1]], p2 = m.tets[4 * t + kFaceOf[i][2]];
            m.surfFaces.insert(m.surfFaces.end(), {p2, p0, p1});
            m.surfFaceTet.push_back(t);
            m.surfFaceIdAtTet.push_back(i);
            onSurface[p0] = onSurface[p1] = onSurface[p2] = 1;
        }
    m.tetVertToSurf.assign(nV, -1);
    for (int v = 0; v < nV; v++)
        if (onSurface[v]) { m.tetVertToSurf[v] = (int32_t)m.surfaceVIds.size(); m.surfaceVIds.push_back(v); }
    const int nF = m.nSurfF();
    // neighbour face across edge k = face that owns the opposite directed edge; sort packed directed edges
    std::vector<std::pair<uint64_t, int32_t>> edges(3 * (size_t)nF);
    for (int f = 0; f < nF; f++)
        for (int k = 0; k < 3; k++) {
            const uint64_t a = (uint32_t)m.surfFaces[3 * f + k], b = (uint32_t)m.surfFaces[3 * f + (k + 1) % 3];
            edges[3 * (size_t)f + k] = {(a << 32) | b, f};
        }
    std::sort(edges.begin(), edges.end());
    m.surfFace3Nbr.assign(3 * (size_t)nF, -1);
    for (int f = 0; f < nF; f++)
        for (int k = 0; k < 3; k++) {
            const uint64_t a = (uint32_t)m.surfFaces[3 * f + k], b = (uint32_t)m.surfFaces[3 * f + (k + 1) % 3];
            const uint64_t key = (b << 32) | a;
            auto it = std::lower_bound(edges.begin(), edges.end(), std::make_pair(key, (int32_t)-1));
            // several faces may own the same directed edge on non-manifold input: the last one created wins
            // (BaseMesh.h:1497-1511 overwrites he_sym)
            int32_t found = -1;
            for (; it != edges.end() && it->first == key; ++it) found = it->second;
            m.surfFace3Nbr[3 * f + k] = found;
        }
    // neighbour surface vertices per surface vertex (ascending ids), CSR
    std::vector<std::vector<int32_t>> nb(m.nSurfV());
    for (int f = 0; f < nF; f++)
        for (int k = 0; k < 3; k++) {
            auto& L = nb[m.tetVertToSurf[m.surfFaces[3 * f + k]]];
            L.push_back(m.surfFaces[3 * f + (k + 1) % 3]);
            L.push_back(m.surfFaces[3 * f + (k + 2) % 3]);
        }
    m.surfNbrOff.assign(m.nSurfV() + 1, 0);
    for (int sv = 0; sv < m.nSurfV(); sv++) {
        auto& L = nb[sv];
        std::sort(L.begin(), L.end());
        L.erase(std::unique(L.begin(), L.end()), L.end());
        m.surfNbrOff[sv + 1] = m.surfNbrOff[sv] + (int32_t)L.size();
        m.surfNbrVerts.insert(m.surfNbrVerts.end(), L.begin(), L.end());
    }
}

void host_mesh_init(HostMesh& m, int32_t nV, int32_t nT, const float* xyz, const int32_t* tets, const gaia_object_params& obj,
                    int32_t nColors, const int32_t* colorOffsets, const int32_t* colorVerts, int32_t nFixed, const int32_t* fixed,
                    bool keepColourOrder) {
    m.nV = nV; m.nT = nT; m.obj = obj;
    m.tets.assign(tets, tets + 4 * (size_t)nT);
    m.pos0.assign(xyz, xyz + 3 * (size_t)nV);
    const float rn2 = obj.rotation[0] * obj.rotation[0] + obj.rotation[1] * obj.rotation[1] + obj.rotation[2] * obj.rotation[2];
    if (std::sqrt(rn2) > 0) {
        float R[9];
        rodrigues(obj.rotation, R);
        for (int v = 0; v < nV; v++) {
            float* p = &m.pos0[3 * (size_t)v];
            const float x = p[0], y = p[1], z = p[2];
            p[0] = R[0] * x + R[1] * y + R[2] * z;
            p[1] = R[3] * x + R[4] * y + R[5] * z;
            p[2] = R[6] * x + R[7] * y + R[8] * z;
        }
    }
    for (int v = 0; v < nV; v++)
        for (int d = 0; d < 3; d++) {
            float& c = m.pos0[3 * (size_t)v + d];
            c *= obj.scale[d];
            c += obj.translation[d];
        }
    m.restVol.resize(nT);
    m.mass.assign(nV, 0.f);
    for (int t = 0; t < nT; t++) {
        const float* x0 = &m.pos0[3 * (size_t)m.tets[4 * t]];
        float e[3][3];
        for (int c = 0; c < 3; c++) {
            const float* xc = &m.pos0[3 * (size_t)m.tets[4 * t + c + 1]];
            for (int d = 0; d < 3; d++) e[c][d] = xc[d] - x0[d];
        }
        const float vol = det3_cols(e[0], e[1], e[2]) / 6;
        m.restVol[t] = vol;
        for (int k = 0; k < 4; k++) {
            float& mv = m.mass[m.tets[4 * t + k]];
            mv = (float)((double)mv + 0.25 * (double)vol * (double)obj.density);
        }
    }
    m.fixedMask.assign(nV, 0);
    m.ghostMask.assign(nV, 0);
    m.fixedPoints.assign(fixed, fixed + nFixed);
    for (int i = 0; i < nFixed; i++) m.fixedMask[fixed[i]] = 1;
    for (int c = 0; c < nColors; c++) m.colors.emplace_back(colorVerts + colorOffsets[c], colorVerts + colorOffsets[c + 1]);
    // size-descending; ties keep file order (documented rule: stable sort)
    // (a partition of a multi-GPU scene keeps the given order: every partition must sweep the same global colour together)
    if (!keepColourOrder)
        std::stable_sort(m.colors.begin(), m.colors.end(),
                         [](const std::vector<int32_t>& a, const std::vector<int32_t>& b) { return a.size() > b.size(); });
    surface_topology(m);
    if (obj.frameToAppear > 0) { m.activeForCollision = false; m.activeForMaterialSolve = false; }
}

// Merge per-mesh colours into global groups, assign sorted ids, build incidence streams. Returns false on bad input.
bool host_finalize(Scene& s) {
    auto& M = s.meshes;
    if (s.phys.checkAndUpdateWorldBounds) {
        for (auto& m : M)
            for (int d = 0; d < 3; d++) {
                if (!m.nV) continue;
                float mn = m.pos0[d], mx = m.pos0[d];
                for (int v = 1; v < m.nV; v++) { mn = std::min(mn, m.pos0[3 * (size_t)v + d]); mx = std::max(mx, m.pos0[3 * (size_t)v + d]); }
                if (s.phys.worldBoundsMin[d] > mn) s.phys.worldBoundsMin[d] = mn;
                if (s.phys.worldBoundsMax[d] < mx) s.phys.worldBoundsMax[d] = mx;
            }
    }
    size_t nGroups = 0;
    for (auto& m : M) nGroups = std::max(nGroups, m.colors.size());
    s.groups.assign(nGroups, {});
    for (size_t im = 0; im < M.size(); im++) {
        HostMesh& m = M[im];
        m.groupOfVertex.assign(m.nV, -1);
        std::vector<int> free_(m.colors.size());
        std::iota(free_.begin(), free_.end(), 0);
        for (auto& colour : m.colors) {
            size_t pick = 0;
            // (a partition of a multi-GPU scene is handed the GLOBAL group of each colour as the colour's index: every
            //  partition must sweep the same global group together, so nothing is re-balanced locally)
            if (!s.partitionMode)
                for (size_t i = 1; i < free_.size(); i++)
                    if (s.groups[free_[i]].size() < s.groups[free_[pick]].size()) pick = i;   // first smallest
            const int g = free_[pick];
            free_.erase(free_.begin() + pick);
            for (int32_t v : colour) {
                if (v < 0 || v >= m.nV) { s.err = "colouring references a vertex out of range"; return false; }
                if (m.groupOfVertex[v] != -1) { s.err = "vertex coloured twice"; return false; }
                s.groups[g].push_back((int32_t)im); s.groups[g].push_back(v);
                m.groupOfVertex[v] = g;
            }
        }
        for (int v = 0; v < m.nV; v++)
            if (m.groupOfVertex[v] == -1) { s.err = "vertex without a colour"; return false; }
    }
    // bases and sorted ids
    int64_t vb = 0, tb = 0, svb = 0, sfb = 0;
    for (auto& m : M) {
        m.vertBase = vb; m.tetBase = tb; m.surfVBase = svb; m.surfFBase = sfb;
        vb += m.nV; tb += m.nT; svb += m.nSurfV(); sfb += m.nSurfF();
        m.sortedId.assign(m.nV, -1);
    }
    s.nVerts = vb; s.nTets = tb; s.nSurfVerts = svb; s.nSurfFaces = sfb;
    if (vb >= (int64_t)1 << 31 || 4 * tb >= (int64_t)1 << 31) { s.err = "scene too large for 32-bit indices"; return false; }
    s.groupStart.assign(nGroups + 1, 0);
    s.sortedToOrig.resize(vb);
    {
        int32_t sid = 0;
        for (size_t g = 0; g < nGroups; g++) {
            s.groupStart[g] = sid;
            const auto& G = s.groups[g];
            for (size_t i = 0; i < G.size(); i += 2) {
                HostMesh& m = M[G[i]];
                m.sortedId[G[i + 1]] = sid;
                s.sortedToOrig[sid] = (int32_t)(m.vertBase + G[i + 1]);
                sid++;
            }
        }
        s.groupStart[nGroups] = sid;
    }
    // incidence streams: per sorted vertex, incident tets in ascending tet id (MeshFrame2 order). Each record is the
    // tet re-ordered by an EVEN permutation that puts the owning vertex first: (x, a, b, c).
    static const int kEven[4][4] = {{0, 1, 2, 3}, {1, 0, 3, 2}, {2, 0, 1, 3}, {3, 0, 2, 1}};
    s.h_incOff.assign(vb + 1, 0);
    for (auto& m : M)
        for (int t = 0; t < m.nT; t++)
            for (int k = 0; k < 4; k++) s.h_incOff[m.sortedId[m.tets[4 * t + k]] + 1]++;
    for (int64_t i = 0; i < vb; i++) s.h_incOff[i + 1] += s.h_incOff[i];
    const int64_t nInc = s.h_incOff[vb];
    s.nInc = nInc;
    s.h_incIdx.resize((size_t)nInc);
    s.h_incW.resize(4 * (size_t)nInc);
    // distinct neighbours of every vertex -> local slots referenced by the incidence records. The list is ordered by the
    // neighbours' ORIGINAL vertex ids (all of them live in the vertex's own mesh): slot order decides which lane sums which
    // neighbour in the sweep, and original ids keep their relative order when a mesh is partitioned, sorted ids do not.
    s.h_nbOff.assign(vb + 1, 0);
    std::vector<int32_t> nbOrig;      // parallel to h_nbList: original (mesh-local) ids
    {
        // incidence CSR (mesh-local tet id, local vertex k) filled in ascending tet order
        std::vector<int32_t> incTet((size_t)nInc);
        std::vector<int32_t> fillAt(s.h_incOff.begin(), s.h_incOff.end() - 1);
        for (auto& m : M)
            for (int t = 0; t < m.nT; t++)
                for (int k = 0; k < 4; k++) incTet[fillAt[m.sortedId[m.tets[4 * (size_t)t + k]]]++] = 4 * t + k;
        std::vector<int32_t> tmp;
        s.h_nbList.reserve((size_t)nInc); nbOrig.reserve((size_t)nInc);
        // vertices are visited in sorted order so nbList is laid out exactly like the device reads it
        for (int64_t sid = 0; sid < vb; sid++) {
            const int64_t orig = s.sortedToOrig[sid];
            // find mesh by base (meshes are few; linear scan from the back is fine, binary search for many)
            size_t lo = 0, hi = M.size();
            while (hi - lo > 1) { size_t mid = (lo + hi) / 2; if (M[mid].vertBase <= orig) lo = mid; else hi = mid; }
            const HostMesh& m = M[lo];
            tmp.clear();
            for (int32_t j = s.h_incOff[sid]; j < s.h_incOff[sid + 1]; j++) {
                const int t = incTet[j] >> 2, k = incTet[j] & 3;
                for (int q = 0; q < 4; q++) if (q != k) tmp.push_back(m.tets[4 * (size_t)t + q]);
            }
            std::sort(tmp.begin(), tmp.end());
            tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
            if (s.h_incOff[sid + 1] - s.h_incOff[sid] > 65535) { s.err = "a vertex has more than 65535 incident tets"; return false; }
            if (tmp.size() > 255) { s.err = "a vertex has more than 255 distinct neighbours"; return false; }
            s.h_nbOff[sid + 1] = s.h_nbOff[sid] + (int32_t)tmp.size();
            nbOrig.insert(nbOrig.end(), tmp.begin(), tmp.end());
            for (int32_t o : tmp) s.h_nbList.push_back(m.sortedId[o]);
        }
    }
    s.nNb = s.h_nbOff[vb];
    auto local_slot = [&](int32_t owner, int32_t nbOriginal) {
        const int32_t* b0 = nbOrig.data() + s.h_nbOff[owner];
        const int32_t* e0 = nbOrig.data() + s.h_nbOff[owner + 1];
        return (uint32_t)(std::lower_bound(b0, e0, nbOriginal) - b0);
    };
    std::vector<int32_t> cursor(s.h_incOff.begin(), s.h_incOff.end() - 1);
    for (auto& m : M) {
        m.nSolved = 0; m.nIncSolved = 0;
        for (int t = 0; t < m.nT; t++) {
            const int32_t* tv = &m.tets[4 * (size_t)t];
            for (int k = 0; k < 4; k++) {
                const int32_t own = tv[kEven[k][0]], ia = tv[kEven[k][1]], ib = tv[kEven[k][2]], ic = tv[kEven[k][3]];
                const float* X = &m.pos0[3 * (size_t)own];
                const float* A = &m.pos0[3 * (size_t)ia];
                const float* B = &m.pos0[3 * (size_t)ib];
                const float* C = &m.pos0[3 * (size_t)ic];
                // Dm' = [A-X, B-X, C-X]; B' = Dm'^-1; m0 = -B'^T 1; w = B' m0   (double, rounded once)
                double D[3][3];
                for (int d = 0; d < 3; d++) { D[d][0] = (double)A[d] - X[d]; D[d][1] = (double)B[d] - X[d]; D[d][2] = (double)C[d] - X[d]; }
                const double det = D[0][0] * (D[1][1] * D[2][2] - D[1][2] * D[2][1]) - D[0][1] * (D[1][0] * D[2][2] - D[1][2] * D[2][0]) +
                                   D[0][2] * (D[1][0] * D[2][1] - D[1][1] * D[2][0]);
                double Bi[3][3];
                Bi[0][0] = (D[1][1] * D[2][2] - D[1][2] * D[2][1]) / det; Bi[0][1] = (D[0][2] * D[2][1] - D[0][1] * D[2][2]) / det; Bi[0][2] = (D[0][1] * D[1][2] - D[0][2] * D[1][1]) / det;
                Bi[1][0] = (D[1][2] * D[2][0] - D[1][0] * D[2][2]) / det; Bi[1][1] = (D[0][0] * D[2][2] - D[0][2] * D[2][0]) / det; Bi[1][2] = (D[0][2] * D[1][0] - D[0][0] * D[1][2]) / det;
                Bi[2][0] = (D[1][0] * D[2][1] - D[1][1] * D[2][0]) / det; Bi[2][1] = (D[0][1] * D[2][0] - D[0][0] * D[2][1]) / det; Bi[2][2] = (D[0][0] * D[1][1] - D[0][1] * D[1][0]) / det;
                double m0[3], w[3];
                for (int c = 0; c < 3; c++) m0[c] = -(Bi[0][c] + Bi[1][c] + Bi[2][c]);
                for (int r = 0; r < 3; r++) w[r] = Bi[r][0] * m0[0] + Bi[r][1] * m0[1] + Bi[r][2] * m0[2];
                // the record carries A*w' and det(B') = 1/(6A) with A the reference's float rest volume:
                //   A*mu*F m0 = mu*Ds'(A w');  A*|m0|^2 = -(1,1,1).(A w');  A*g = -n/6;  A*g g^T = det(B')/6 * n n^T
                const double vol = (double)m.restVol[t];
                const int32_t sOwn = m.sortedId[own];
                const int32_t at = cursor[sOwn]++;
                float* rw = &s.h_incW[4 * (size_t)at];
                rw[0] = (float)(vol * w[0]); rw[1] = (float)(vol * w[1]); rw[2] = (float)(vol * w[2]); rw[3] = (float)(1.0 / (6.0 * vol));
                s.h_incIdx[at] = local_slot(sOwn, ia) | (local_slot(sOwn, ib) << 8) | (local_slot(sOwn, ic) << 16);
            }
        }
        for (int v = 0; v < m.nV; v++)
            if (!m.fixedMask[v] && !m.ghostMask[v]) {
                m.nSolved++;
                m.nIncSolved += s.h_incOff[m.sortedId[v] + 1] - s.h_incOff[m.sortedId[v]];
            }
        // |N(c)|: distinct neighbour vertices read by this mesh's part of each group
        m.nbrDistinctPerGroup.assign(nGroups, 0);
        std::vector<int32_t> stamp(m.nV, -1);
        for (size_t g = 0; g < nGroups; g++) {
            int64_t cnt = 0;
            for (int t = 0; t < m.nT; t++) {
                const int32_t* tv = &m.tets[4 * (size_t)t];
                bool touches = false;
                for (int k = 0; k < 4; k++) if (m.groupOfVertex[tv[k]] == (int)g && !m.fixedMask[tv[k]] && !m.ghostMask[tv[k]]) touches = true;
                if (!touches) continue;
                for (int q = 0; q < 4; q++)
                    if (m.groupOfVertex[tv[q]] != (int)g && stamp[tv[q]] != (int)g) { stamp[tv[q]] = (int)g; cnt++; }
            }
            m.nbrDistinctPerGroup[g] = cnt;
        }
    }
    return true;
}

// ---- record order of one vertex: triangle strips over its link.
// The incident tets of a vertex, seen from the vertex, are the triangles (a,b,c) of its link. The sweep keeps the three
// neighbour positions of a lane's current tet in registers; when the lane's next tet shares an edge with the current one in
// the strict triangle-strip pattern (it replaces the OLDEST of the three vertices) only ONE position has to be read from
// shared memory instead of three. plan_strips orders the triangles into strips (greedy: start at the unused triangle with
// the fewest unused neighbours, try its six vertex orders, keep the longest run), emit_lane_records replays a lane's share
// of that order against the register file the kernel will hold and writes either a continuation (one slot) or a restart
// (three slots). Ties are broken by the triangle's position in ascending tet order and by ORIGINAL vertex ids only, so a
// partition of the mesh (order-preserving renumbering) plans the same strips for the vertices it owns.
struct LinkTri { int32_t v[3]; };          // link-local vertex indices (rank of the original id among the neighbours), oriented

static void plan_strips(const std::vector<LinkTri>& tri, int nLinkV, std::vector<int>& order, std::vector<int>& scratch) {
    const int n = (int)tri.size();
    order.clear();
    // edge -> up to two triangles (further ones on a non-manifold edge are simply not linked)
    std::vector<int>& edgeTri = scratch;
    edgeTri.assign((size_t)nLinkV * nLinkV * 2, -1);
    auto ekey = [&](int a, int b) { return (size_t)(std::min(a, b) * nLinkV + std::max(a, b)) * 2; };
    for (int t = 0; t < n; t++)
        for (int e = 0; e < 3; e++) {
            const size_t k = ekey(tri[t].v[e], tri[t].v[(e + 1) % 3]);
            if (edgeTri[k] < 0) edgeTri[k] = t; else if (edgeTri[k + 1] < 0) edgeTri[k + 1] = t;
        }
    auto across = [&](int t, int a, int b) { const size_t k = ekey(a, b); return edgeTri[k] == t ? edgeTri[k + 1] : (edgeTri[k + 1] == t ? edgeTri[k] : -1); };
    auto third = [&](int t, int a, int b) { for (int e = 0; e < 3; e++) if (tri[t].v[e] != a && tri[t].v[e] != b) return tri[t].v[e]; return -1; };
    std::vector<char> used(n, 0);
    std::vector<int> run, best;
    static const int kPerm[6][3] = {{0, 1, 2}, {1, 2, 0}, {2, 0, 1}, {0, 2, 1}, {2, 1, 0}, {1, 0, 2}};
    int left = n;
    while (left) {
        int start = -1, fewest = 4;
        for (int t = 0; t < n; t++) {
            if (used[t]) continue;
            int c = 0;
            for (int e = 0; e < 3; e++) { const int o = across(t, tri[t].v[e], tri[t].v[(e + 1) % 3]); if (o >= 0 && !used[o]) c++; }
            if (c < fewest) { fewest = c; start = t; }
        }
        best.clear();
        for (int pidx = 0; pidx < 6; pidx++) {
            run.clear(); run.push_back(start);
            used[start] = 1;
            int v1 = tri[start].v[kPerm[pidx][1]], v2 = tri[start].v[kPerm[pidx][2]], cur = start;
            for (;;) {
                const int nx = across(cur, v1, v2);
                if (nx < 0 || used[nx]) break;
                const int v3 = third(nx, v1, v2);
                if (v3 < 0) break;
                used[nx] = 1; run.push_back(nx);
                v1 = v2; v2 = v3; cur = nx;
            }
            for (int t : run) used[t] = 0;
            if (run.size() > best.size()) best = run;
        }
        for (int t : best) { used[t] = 1; order.push_back(t); }
        left -= (int)best.size();
    }
}

// Lane's share `seq` of the planned order -> records. reg[i] = link vertex held by register i of the kernel; step t replaces
// register t % 3. Returns the number of shared-memory position reads the lane will do.
static int emit_lane_records(const std::vector<LinkTri>& tri, const int* seq, int cnt, const int* slotOfLinkV, const float* detB,
                             uint2* out, int stride) {
    int reg[3] = {-1, -1, -1}, reads = 0;
    auto bits = [](float f) { uint32_t i; std::memcpy(&i, &f, 4); return i; };
    for (int t = 0; t < cnt; t++) {
        const LinkTri& T = tri[seq[t]];
        const int r0 = t % 3, r1 = (t + 1) % 3, r2 = (t + 2) % 3;
        int place[3] = {-1, -1, -1};      // link vertex for registers r0 (new), r1, r2
        bool cont = false;
        if (t > 0) {
            // continuation: the triangle holds the vertices in r1 and r2; its third vertex goes to r0
            int nw = -1, hit = 0;
            for (int e = 0; e < 3; e++) { if (T.v[e] == reg[r1] || T.v[e] == reg[r2]) hit++; else nw = T.v[e]; }
            if (hit == 2 && reg[r1] != reg[r2] && nw >= 0) { cont = true; place[0] = nw; place[1] = reg[r1]; place[2] = reg[r2]; }
        }
        if (!cont) {
            // restart: put the edge shared with the NEXT triangle of the lane into r2 and r0 (r1 is replaced next), if any
            int o = 0;
            if (t + 1 < cnt) {
                const LinkTri& N = tri[seq[t + 1]];
                for (int e = 0; e < 3; e++) {
                    bool inN = false;
                    for (int f = 0; f < 3; f++) inN |= (N.v[f] == T.v[e]);
                    if (!inN) { o = e; break; }           // the vertex the next triangle does not have goes to r1
                }
            }
            // the vertex after next drops r2: prefer the strict-strip order (v0 -> r1, v1 -> r2, v2 -> r0)
            place[1] = T.v[o]; place[2] = T.v[(o + 1) % 3]; place[0] = T.v[(o + 2) % 3];
            if (t + 2 < cnt && t + 1 < cnt) {
                const LinkTri& N2 = tri[seq[t + 2]];
                bool keeps = false;                        // the triangle after next must still hold what sits in r0
                for (int f = 0; f < 3; f++) keeps |= (N2.v[f] == place[0]);
                if (!keeps) std::swap(place[0], place[2]);
            }
        }
        reg[r0] = place[0]; reg[r1] = place[1]; reg[r2] = place[2];
        // orientation of (reg0, reg1, reg2) against the triangle's own (a, b, c)
        int pos[3];
        for (int i = 0; i < 3; i++) for (int e = 0; e < 3; e++) if (reg[i] == T.v[e]) pos[i] = e;
        const bool even = (pos[1] == (pos[0] + 1) % 3);
        uint32_t I = (uint32_t)slotOfLinkV[place[0]];
        if (!cont) I |= ((uint32_t)slotOfLinkV[place[1]] << 8) | ((uint32_t)slotOfLinkV[place[2]] << 16) | (1u << 30);
        if (!even) I |= 1u << 31;
        out[(size_t)t * stride] = make_uint2(bits(detB[seq[t]]), I);
        reads += cont ? 1 : 3;
    }
    return reads;
}

// Tile streams of the colour sweep (sweep.cu). The Neo-Hookean vertex force splits into a part that is LINEAR in the
// neighbour positions, -mu * sum_tets Ds'(A w') = -mu * sum_nb W_nb (x_nb - x) with W_nb the sum of A w' over the tets
// that share the neighbour, and the volumetric part that needs the tet (det F and the opposite face's normal). So a
// vertex carries one 8-byte entry per distinct neighbour {sorted id, W_nb} and one 8-byte record per incident tet
// {det B', three neighbour slots}; A*sum|m0|^2 = -sum_nb W_nb is a per-vertex constant in the header. Tiles are runs of
// at most tileV vertices of one colour whose entries fit the shared-memory stage; sums are formed in double in ascending
// tet order and rounded once, so a partition of the mesh builds the same numbers for the vertices it owns.
bool build_tile_streams(Scene& s, int tileV) {
    const int64_t N = s.nVerts;
    const size_t nGroups = s.groupStart.size() - 1;
    tileV = std::max(1, std::min(tileV, kTileV));
    s.h_vtx2.assign(N, make_int4(0, 0, 0, (int)0x80000000));
    s.h_rec.clear(); s.h_nbw.clear(); s.h_tiles.clear();
    s.h_rec.reserve((size_t)s.nInc + 2 * (size_t)(N / tileV + nGroups + 1));
    s.h_nbw.reserve((size_t)s.nNb + 2 * (size_t)(N / tileV + nGroups + 1));
    s.groupTileStart.assign(nGroups + 1, 0);
    std::vector<int32_t> info(N);
    for (auto& m : s.meshes)
        for (int v = 0; v < m.nV; v++)
            info[m.sortedId[v]] = (m.id & 0x00FFFFFF) | ((m.fixedMask[v] || m.ghostMask[v]) ? (int32_t)0x80000000 : 0);
    std::vector<double> wsum;
    std::vector<LinkTri> tri;
    std::vector<float> detB;
    std::vector<int> order, scratch;
    int identity[256];
    for (int i = 0; i < 256; i++) identity[i] = i;
    s.sweepPosReads = 0;
    auto bits = [](float f) { int32_t i; std::memcpy(&i, &f, 4); return i; };
    for (size_t g = 0; g < nGroups; g++) {
        s.groupTileStart[g] = (int32_t)s.h_tiles.size();
        int64_t v = s.groupStart[g];
        const int64_t ve = s.groupStart[g + 1];
        while (v < ve) {
            TileDesc t;
            std::memset(&t, 0, sizeof(t));
            t.vBegin = (int32_t)v; t.recBegin = (int32_t)s.h_rec.size(); t.nbBegin = (int32_t)s.h_nbw.size();
            int nRec = 0, nNb = 0;
            while (v < ve && t.nV < tileV) {
                const bool solved = info[v] >= 0;
                const int r = solved ? s.h_incOff[v + 1] - s.h_incOff[v] : 0;
                const int n = solved ? s.h_nbOff[v + 1] - s.h_nbOff[v] : 0;
                if (r > kTileRecCap || n > kTileNbCap) { s.err = "a vertex has more incident tets than a sweep tile holds"; return false; }
                if (nRec + r > kTileRecCap || nNb + n > kTileNbCap) break;
                wsum.assign(n, 0.0);
                double s1 = 0.0;
                tri.resize(r); detB.resize(r);
                for (int q = 0; q < r; q++) {
                    const int j = s.h_incOff[v] + q;
                    const float* W = &s.h_incW[4 * (size_t)j];
                    const uint32_t I = s.h_incIdx[j];
                    wsum[I & 255] += W[0]; wsum[(I >> 8) & 255] += W[1]; wsum[(I >> 16) & 255] += W[2];
                    s1 -= (double)W[0] + (double)W[1] + (double)W[2];
                    tri[q].v[0] = (int)(I & 255); tri[q].v[1] = (int)((I >> 8) & 255); tri[q].v[2] = (int)((I >> 16) & 255);
                    detB[q] = W[3];
                }
                // records in lane-strided strip order: lane l walks records l, l + 4, ... = its share of the planned order
                plan_strips(tri, n, order, scratch);
                const size_t base = s.h_rec.size();
                s.h_rec.resize(base + r);
                for (int l = 0, at = 0; l < kSweepLanes; l++) {
                    const int cnt = (r - l + kSweepLanes - 1) / kSweepLanes;
                    if (cnt <= 0) break;
                    s.sweepPosReads += emit_lane_records(tri, order.data() + at, cnt, identity, detB.data(), s.h_rec.data() + base + l, kSweepLanes);
                    at += cnt;
                }
                for (int q = 0; q < n; q++) s.h_nbw.push_back(make_int2(s.h_nbList[s.h_nbOff[v] + q], bits((float)wsum[q])));
                s.h_vtx2[v] = make_int4(nRec | (r << 16), nNb | (n << 16), bits((float)s1), info[v]);
                nRec += r; nNb += n; t.nV++; v++;
            }
            if (nRec & 1) { s.h_rec.push_back(make_uint2(0, 0)); nRec++; }
            if (nNb & 1) { s.h_nbw.push_back(make_int2(0, 0)); nNb++; }
            t.recCount = nRec; t.nbCount = nNb;
            s.h_tiles.push_back(t);
        }
    }
    s.groupTileStart[nGroups] = (int32_t)s.h_tiles.size();
    if (s.h_rec.size() >= ((size_t)1 << 31) || s.h_nbw.size() >= ((size_t)1 << 31)) { s.err = "scene too large for 32-bit stream offsets"; return false; }
    return true;
}

}  // namespace gb
