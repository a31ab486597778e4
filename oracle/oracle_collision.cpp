// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// Collision detection of the VBD step, restating
//   VBDPhysics::dcd / intermediateDCD / ccd / intermediateCCD      M/VBD/VBDPhysics.cpp:1958-2156
//   prepareCollisionDataCPU + ActiveCollisionList::addToActiveCollisionList
//                                                                   M/VBD/VBDPhysics.cpp:1709-1810, M/VBD/ActiveCollisionList.cpp:43-110
//   updateCollisionInfo                                             M/VBD/VBDPhysics.cpp:2205-2311
//   tetIntersectionFunc / closestPointQueryFunc / feasible regions  M/CollisionDetector/DiscreteCollisionDetector.cpp:12-319, 664-815
//   pointInTet / closestPointTriangle                               M/CollisionDetector/CollisionGeometry.h:56-157
//   tetrahedralTraverseTo (+Dynamic) and helpers                    M/TetMesh/TetMeshFEM.cpp:503-742, 909-..., TetMeshFEM.h:423-560
//   continuousTriPointIntersectionFunc / vertexContinuousCollisionDetection
//                                                                   M/CollisionDetector/ContinuousCollisionDetector.cpp:70-274, 352-381
//   cy::IntersectContinuousTriPoint (restated in oracle_ccd.h)      M/CollisionDetector/CCDSolver.h:103-202, 3P/CyCodeBase/cyPolynomial.h
//
// Embree (3.13, absent here) only supplies BVH culling and the ORDER in which candidates reach the callbacks.
// Where the reference's outcome depends on that order the oracle applies these CANONICAL RULES (the product
// applies the same ones; DESIGN.md §"canonical contact rules"):
//   R1  DCD phase 1 ("first tet found", maxNumberOfCollisions = 1): the containing tet with the smallest
//       (meshId, tetId).
//   R2  closest surface point: smallest distance d among the faces that pass the feasibility + traversal tests;
//       ties -> smallest face id.  maxNumberOfBVHQuery (an Embree-order-dependent early-out) is applied in its order-independent
//       form R2': the query fails when at least that many faces are strictly nearer than the nearest acceptable one.
//   R3  CCD: smallest penetrationDepth = t*|p1-p0| over ALL faces of active meshes; ties -> smallest
//       (meshId, faceId).  (Any true hit lies inside the query sphere, so the sphere only culls.)
//   R4  tetrahedral traversal uses unbounded stacks (= static version when it does not overflow, = the
//       dynamic fallback otherwise).
// The BVH below (median-split AABB tree) is an accelerator with conservative culling: results equal brute force
// (tests/test_oracle_collision.py checks that).
#include "vbd_oracle.h"
#include "oracle_ccd.h"
#include <limits>

namespace orc {

static const int kTetFaces[4][3] = {{1, 2, 3}, {2, 0, 3}, {0, 1, 3}, {1, 0, 2}};
static const int kPossibleExit3Faces[3][3] = {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}};   // TetMeshFEM.cpp:25
static const float CMP_EPSILON = 1e-5f;
static const int GRAVITY_AXIS = 1;
enum { AtA = 0, AtB, AtC, AtAB, AtBC, AtAC, AtInterior, NotFound };

// ---------------------------------------------------------------- geometry predicates
bool point_in_tet(V3 p, V3 a, V3 b, V3 c, V3 d) {
    const V3 AB = b - a, AC = c - a, AD = d - a;
    const float vol = dot(AB, cross(AC, AD));
    if (std::fabs(vol) < CMP_EPSILON) return false;
    const V3 vs[4] = {a, b, c, d};
    for (int i = 0; i < 4; i++) {
        const V3 v1 = vs[kTetFaces[i][0]], v2 = vs[kTetFaces[i][1]], v3 = vs[kTetFaces[i][2]];
        const V3 v = p - v1;
        if (dot(v, cross(v2 - v1, v3 - v2)) * vol >= 0) return false;
    }
    return true;
}

V3 closest_point_triangle(V3 p, V3 a, V3 b, V3 c, V3& bary, int& type) {
    const V3 ab = b - a, ac = c - a, ap = p - a;
    const float d1 = dot(ab, ap), d2 = dot(ac, ap);
    if (d1 <= 0.f && d2 <= 0.f) { type = AtA; bary = V3(1, 0, 0); return a; }
    const V3 bp = p - b;
    const float d3 = dot(ab, bp), d4 = dot(ac, bp);
    if (d3 >= 0.f && d4 <= d3) { type = AtB; bary = V3(0, 1, 0); return b; }
    const V3 cp = p - c;
    const float d5 = dot(ab, cp), d6 = dot(ac, cp);
    if (d6 >= 0.f && d5 <= d6) { type = AtC; bary = V3(0, 0, 1); return c; }
    const float vc = d1 * d4 - d3 * d2;
    if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) {
        const float v = d1 / (d1 - d3);
        type = AtAB; bary = V3(1.0f - v, v, 0.f);
        return a + v * ab;
    }
    const float vb = d5 * d2 - d1 * d6;
    if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) {
        const float v = d2 / (d2 - d6);
        type = AtAC; bary = V3(1.0f - v, 0.f, v);
        return a + v * ac;
    }
    const float va = d3 * d6 - d5 * d4;
    if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) {
        type = AtBC;
        const float v = (d4 - d3) / ((d4 - d3) + (d5 - d6));
        bary = V3(0.f, 1.f - v, v);
        return b + v * (c - b);
    }
    const float denom = 1.f / (va + vb + vc);
    const float v = vb * denom, w = vc * denom;
    type = AtInterior; bary = V3(1.f - v - w, v, w);
    return (a + v * ab) + w * ac;
}

static inline V3 face_normal(const Mesh& m, int f) {
    V3 a = m.pos[m.surfFaces[3 * f]], b = m.pos[m.surfFaces[3 * f + 1]], c = m.pos[m.surfFaces[3 * f + 2]];
    V3 n = cross(b - a, c - a);
    return n / norm(n);
}

static bool check_edge_feasible(V3 p, const Mesh& m, int faceId, int edgeId, int e1, int e2, float eps) {
    const int nbr = m.surfFace3Nbr[3 * faceId + edgeId];
    const V3 v1 = m.pos[e1], v2 = m.pos[e2];
    if (nbr == -1) return true;
    const V3 fN1 = -face_normal(m, faceId), fN2 = -face_normal(m, nbr);
    const V3 AP = p - v1, BP = p - v2, AB = v2 - v1, BA = -AB;
    const float relaxed = -(sqnorm(AB)) * eps - 0.f;
    if (dot(AP, AB) < relaxed) return false;
    if (dot(BP, BA) < relaxed) return false;
    if (dot(AP, cross(fN1, AB)) < relaxed) return false;
    if (dot(AP, cross(fN2, BA)) < relaxed) return false;
    return true;
}

static bool check_vertex_feasible(V3 p, const Mesh& m, int vId, float eps) {
    const V3 A = m.pos[vId], AP = p - A;
    const int sv = m.tetVertToSurf[vId];
    for (int nb : m.surfVertNbrSurfVerts[sv]) {
        const V3 BA = A - m.pos[nb];
        const float relaxed = -(dot(BA, BA)) * eps - 0.f;
        if (dot(AP, BA) < relaxed) return false;
    }
    return true;
}

static bool check_feasible_region(V3 p, const Mesh& m, int faceId, int type, float eps) {
    const int* fv = &m.surfFaces[3 * faceId];
    switch (type) {
        case AtInterior: return true;
        case AtAB: return check_edge_feasible(p, m, faceId, 0, fv[0], fv[1], eps);
        case AtBC: return check_edge_feasible(p, m, faceId, 1, fv[1], fv[2], eps);
        case AtAC: return check_edge_feasible(p, m, faceId, 2, fv[2], fv[0], eps);
        case AtA: return check_vertex_feasible(p, m, fv[0], eps);
        case AtB: return check_vertex_feasible(p, m, fv[1], eps);
        case AtC: return check_vertex_feasible(p, m, fv[2], eps);
        default: return false;
    }
}

bool probe_feasible_region(const Mesh& m, V3 p, int faceId, int type, float eps) { return check_feasible_region(p, m, faceId, type, eps); }   // parity probe

// ---------------------------------------------------------------- tetrahedral traversal
static void build_orthonormal_basis(V3 n, V3& b1, V3& b2) {   // CuMatrix Geometry.h:183-206
    if (n.z < 0.f) {
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
struct P2 { float x, y; };
static inline float cross2(P2 a, P2 b) { return a.x * b.y - a.y * b.x; }

static void project_2d(const Mesh& m, P2 out[4], int tetId, int inFace, V3 ax0, V3 ax1, V3 origin) {
    const int* tv = &m.tets[4 * (size_t)tetId];
    const int ids[4] = {tv[kTetFaces[inFace][0]], tv[kTetFaces[inFace][1]], tv[kTetFaces[inFace][2]], tv[inFace]};
    for (int i = 0; i < 4; i++) {
        V3 d = m.pos[ids[i]] - origin;
        out[i].x = dot(ax0, d); out[i].y = dot(ax1, d);
    }
}
static int exit_face_selection(const P2 p[4], int possible[3], float eps) {
    int n = 0;
    possible[0] = possible[1] = possible[2] = 0;
    P2 v1 = {p[1].x - p[0].x, p[1].y - p[0].y}, v2 = {p[2].x - p[0].x, p[2].y - p[0].y};
    const float area = cross2(v1, v2);
    const float sgn = std::copysign(1.0f, area);
    const float d0 = cross2(p[3], p[0]), d1 = cross2(p[3], p[1]), d2 = cross2(p[3], p[2]);
    if (d1 * sgn >= -eps && d2 * sgn <= eps) { possible[0] = 1; ++n; }
    if (d2 * sgn >= -eps && d0 * sgn <= eps) { possible[1] = 1; ++n; }
    if (d0 * sgn >= -eps && d1 * sgn <= eps) { possible[2] = 1; ++n; }
    return n;
}
static void origin_barycentric_2d(P2 a, P2 b, P2 c, float barys[3]) {
    P2 v0 = {b.x - a.x, b.y - a.y}, v1 = {c.x - a.x, c.y - a.y}, v2 = {-a.x, -a.y};
    const float den = v0.x * v1.y - v1.x * v0.y;
    barys[1] = (v2.x * v1.y - v1.x * v2.y) / den;
    barys[2] = (v0.x * v2.y - v2.x * v0.y) / den;
    barys[0] = 1.0f - barys[1] - barys[2];
}

bool tetrahedral_traverse_to(const Mesh& m, V3 rayOrigin, V3 rayDir, float maxDis, int startTet, int startFace,
                             int targetTet, float eps, int* numTraversed) {
    if (numTraversed) *numTraversed = 0;
    if (startTet == targetTet) return true;
    V3 ax0, ax1;
    build_orthonormal_basis(rayDir, ax0, ax1);
    P2 pts[4];
    project_2d(m, pts, startTet, startFace, ax0, ax1, rayOrigin);
    int possible[3];
    exit_face_selection(pts, possible, eps);
    std::vector<int> stackFace, stackTet;
    // CircularArray<int32_t, 32> (MeshFrame Memory/Array.h:363)
    const int CAP = 32;
    int ring[CAP]; int rBegin = 0, rSize = 0;
    auto ring_push = [&](int v) {
        if (rSize < CAP) { ring[(rBegin + rSize) % CAP] = v; rSize++; }
        else { ring[rBegin] = v; rBegin = (rBegin + 1) % CAP; }
    };
    auto ring_has = [&](int v) { for (int i = 0; i < rSize; i++) if (ring[(rBegin + i) % CAP] == v) return true; return false; };
    ring_push(startTet);
    for (int i = 0; i < 3; i++) if (possible[i]) { stackFace.push_back(kTetFaces[startFace][i]); stackTet.push_back(startTet); }
    int nTrav = 0;
    while (!stackFace.empty()) {
        const int hitFace = stackFace.back(); stackFace.pop_back();
        const int prevTet = stackTet.back(); stackTet.pop_back();
        const int* ptv = &m.tets[4 * (size_t)prevTet];
        const int fv[3] = {ptv[kTetFaces[hitFace][0]], ptv[kTetFaces[hitFace][1]], ptv[kTetFaces[hitFace][2]]};
        const int cur = m.tetNbrTets[4 * (size_t)prevTet + hitFace];
        if (cur == targetTet) { if (numTraversed) *numTraversed = nTrav; return true; }
        if (cur == -1) { if (numTraversed) *numTraversed = nTrav; return false; }
        if (ring_has(cur)) continue;
        ring_push(cur);
        const int newV = m.tetXorSums[cur] ^ fv[0] ^ fv[1] ^ fv[2];
        int inFace = 3;
        for (int k = 0; k < 3; k++) if (newV == m.tets[4 * (size_t)cur + k]) { inFace = k; break; }
        project_2d(m, pts, cur, inFace, ax0, ax1, rayOrigin);
        exit_face_selection(pts, possible, eps);
        for (int iF = 0; iF < 3; iF++) {
            if (!possible[iF]) continue;
            const int exitFace = kTetFaces[inFace][iF];
            const int nextTet = m.tetNbrTets[4 * (size_t)cur + exitFace];
            if (nextTet == targetTet) { if (numTraversed) *numTraversed = nTrav; return true; }
            stackFace.push_back(exitFace); stackTet.push_back(cur);
            if ((!(nTrav % 5)) && maxDis > 0.f) {
                float barys[3];
                origin_barycentric_2d(pts[kPossibleExit3Faces[iF][0]], pts[kPossibleExit3Faces[iF][1]], pts[kPossibleExit3Faces[iF][2]], barys);
                const int* ctv = &m.tets[4 * (size_t)cur];
                const int ids[4] = {ctv[kTetFaces[inFace][0]], ctv[kTetFaces[inFace][1]], ctv[kTetFaces[inFace][2]], ctv[inFace]};
                V3 ip = (barys[0] * m.pos[ids[kPossibleExit3Faces[iF][0]]] + barys[1] * m.pos[ids[kPossibleExit3Faces[iF][1]]])
                        + barys[2] * m.pos[ids[kPossibleExit3Faces[iF][2]]];
                if (sqnorm(ip - rayOrigin) > maxDis * maxDis) { if (numTraversed) *numTraversed = nTrav; return false; }
            }
        }
        ++nTrav;
    }
    if (numTraversed) *numTraversed = nTrav;
    return false;
}

// ---------------------------------------------------------------- AABB tree (accelerator only)
struct Box { float lo[3], hi[3]; };
static inline Box box_empty() { Box b; for (int d = 0; d < 3; d++) { b.lo[d] = 1e30f; b.hi[d] = -1e30f; } return b; }
static inline void box_add(Box& b, V3 p) { for (int d = 0; d < 3; d++) { b.lo[d] = std::min(b.lo[d], p[d]); b.hi[d] = std::max(b.hi[d], p[d]); } }
static inline void box_merge(Box& b, const Box& o) { for (int d = 0; d < 3; d++) { b.lo[d] = std::min(b.lo[d], o.lo[d]); b.hi[d] = std::max(b.hi[d], o.hi[d]); } }
static inline float box_dist2(const Box& b, V3 p) {
    float s = 0;
    for (int d = 0; d < 3; d++) { float v = p[d] < b.lo[d] ? b.lo[d] - p[d] : (p[d] > b.hi[d] ? p[d] - b.hi[d] : 0.f); s += v * v; }
    return s;
}
struct Tree {
    struct Node { Box box; int left, right, begin, end; };   // leaf: left == -1, prims [begin,end)
    std::vector<Node> nodes;
    std::vector<int> prims;          // permutation of primitive ids
    std::vector<Box> primBoxes;      // by primitive id
    void build(const std::vector<Box>& boxes) {
        primBoxes = boxes;
        prims.resize(boxes.size());
        for (size_t i = 0; i < prims.size(); i++) prims[i] = (int)i;
        nodes.clear(); nodes.reserve(2 * boxes.size() / 2 + 2);
        if (!boxes.empty()) build_rec(0, (int)boxes.size());
    }
    int build_rec(int b, int e) {
        int id = (int)nodes.size();
        nodes.push_back(Node());
        Box bb = box_empty(), cb = box_empty();
        for (int i = b; i < e; i++) {
            const Box& pb = primBoxes[prims[i]];
            if (pb.lo[0] > pb.hi[0]) continue;
            box_merge(bb, pb);
            box_add(cb, V3(0.5f * (pb.lo[0] + pb.hi[0]), 0.5f * (pb.lo[1] + pb.hi[1]), 0.5f * (pb.lo[2] + pb.hi[2])));
        }
        nodes[id].box = bb; nodes[id].begin = b; nodes[id].end = e; nodes[id].left = nodes[id].right = -1;
        if (e - b <= 4) return id;
        int ax = 0; float ext = -1;
        for (int d = 0; d < 3; d++) if (cb.hi[d] - cb.lo[d] > ext) { ext = cb.hi[d] - cb.lo[d]; ax = d; }
        int mid = (b + e) / 2;
        std::nth_element(prims.begin() + b, prims.begin() + mid, prims.begin() + e, [&](int x, int y) {
            return primBoxes[x].lo[ax] + primBoxes[x].hi[ax] < primBoxes[y].lo[ax] + primBoxes[y].hi[ax];
        });
        int l = build_rec(b, mid); int r = build_rec(mid, e);
        nodes[id].left = l; nodes[id].right = r;
        return id;
    }
    void refit(const std::vector<Box>& boxes) {
        primBoxes = boxes;
        for (int id = (int)nodes.size() - 1; id >= 0; id--) {
            Node& n = nodes[id];
            Box bb = box_empty();
            if (n.left < 0) { for (int i = n.begin; i < n.end; i++) { const Box& pb = primBoxes[prims[i]]; if (pb.lo[0] <= pb.hi[0]) box_merge(bb, pb); } }
            else { bb = nodes[n.left].box; box_merge(bb, nodes[n.right].box); }
            n.box = bb;
        }
    }
    // visit(prim) for every primitive whose box is within sqrt(r2()) of p; r2 may shrink during the query
    template <class R2, class Visit>
    void query(V3 p, R2 r2, Visit visit) const {
        if (nodes.empty()) return;
        int stack[128]; int sp = 0; stack[sp++] = 0;
        while (sp) {
            const Node& n = nodes[stack[--sp]];
            if (n.box.lo[0] > n.box.hi[0]) continue;
            if (box_dist2(n.box, p) > r2()) continue;
            if (n.left < 0) {
                for (int i = n.begin; i < n.end; i++) {
                    const Box& pb = primBoxes[prims[i]];
                    if (pb.lo[0] > pb.hi[0]) continue;
                    if (box_dist2(pb, p) > r2()) continue;
                    visit(prims[i]);
                }
            } else { stack[sp++] = n.left; stack[sp++] = n.right; }
        }
    }
};

struct TetScene { Tree tree; std::vector<int> primMesh, primTet; std::vector<uint8_t> activeAtBuild; bool built = false; };
struct SurfScenes { std::vector<Tree> trees; std::vector<uint8_t> built; };
struct CcdScene { Tree tree; std::vector<int> primMesh, primFace; std::vector<uint8_t> activeAtBuild; bool built = false; };
struct ColAccel { TetScene tet; SurfScenes surf; CcdScene ccd; bool bruteForce = false; };

static const float kInflate = 1e-4f;   // conservative culling margin (relative + absolute), accelerator only
static inline void inflate(Box& b) {
    for (int d = 0; d < 3; d++) {
        float m = kInflate * (1.f + std::max(std::fabs(b.lo[d]), std::fabs(b.hi[d])));
        b.lo[d] -= m; b.hi[d] += m;
    }
}

static void tet_boxes(const Scene& s, TetScene& ts, std::vector<Box>& boxes) {
    boxes.resize(ts.primTet.size());
    for (size_t i = 0; i < boxes.size(); i++) {
        const Mesh& m = s.meshes[ts.primMesh[i]];
        if (!m.activeForCollision) { boxes[i] = box_empty(); continue; }
        const int* tv = &m.tets[4 * (size_t)ts.primTet[i]];
        Box b = box_empty();
        for (int k = 0; k < 4; k++) box_add(b, m.pos[tv[k]]);
        inflate(b);
        boxes[i] = b;
    }
}
static void surf_boxes(const Mesh& m, std::vector<Box>& boxes) {
    boxes.resize(m.nSurfF());
    for (int f = 0; f < m.nSurfF(); f++) {
        Box b = box_empty();
        for (int k = 0; k < 3; k++) box_add(b, m.pos[m.surfFaces[3 * f + k]]);
        inflate(b);
        boxes[f] = b;
    }
}
static void ccd_boxes(const Scene& s, CcdScene& cs, std::vector<Box>& boxes) {
    boxes.resize(cs.primFace.size());
    for (size_t i = 0; i < boxes.size(); i++) {
        const Mesh& m = s.meshes[cs.primMesh[i]];
        if (!m.activeForCollision) { boxes[i] = box_empty(); continue; }
        Box b = box_empty();
        for (int k = 0; k < 3; k++) { int v = m.surfFaces[3 * cs.primFace[i] + k]; box_add(b, m.pos[v]); box_add(b, m.prev[v]); }
        inflate(b);
        boxes[i] = b;
    }
}

void collision_initialize(Scene& s) {
    collision_free(s);
    ColAccel* a = new ColAccel();
    for (size_t im = 0; im < s.meshes.size(); im++) {
        for (int t = 0; t < s.meshes[im].nT; t++) { a->tet.primMesh.push_back((int)im); a->tet.primTet.push_back(t); }
        for (int f = 0; f < s.meshes[im].nSurfF(); f++) { a->ccd.primMesh.push_back((int)im); a->ccd.primFace.push_back(f); }
    }
    a->surf.trees.resize(s.meshes.size()); a->surf.built.assign(s.meshes.size(), 0);
    s.bvhTet = a;
}
void collision_free(Scene& s) {
    delete (ColAccel*)s.bvhTet;
    s.bvhTet = nullptr;
}

static bool activity_changed(const Scene& s, std::vector<uint8_t>& snap) {
    bool ch = snap.size() != s.meshes.size();
    snap.resize(s.meshes.size());
    for (size_t i = 0; i < s.meshes.size(); i++) { uint8_t a = s.meshes[i].activeForCollision; if (snap[i] != a) ch = true; snap[i] = a; }
    return ch;
}

static void update_dcd_bvh(Scene& s, bool rebuildTet, bool rebuildSurf) {
    ColAccel* a = (ColAccel*)s.bvhTet;
    std::vector<Box> boxes;
    tet_boxes(s, a->tet, boxes);
    bool ch = activity_changed(s, a->tet.activeAtBuild);
    if (!a->tet.built || rebuildTet || ch) { a->tet.tree.build(boxes); a->tet.built = true; }
    else a->tet.tree.refit(boxes);
    for (size_t im = 0; im < s.meshes.size(); im++) {
        if (!s.meshes[im].activeForCollision) continue;
        surf_boxes(s.meshes[im], boxes);
        if (!a->surf.built[im] || rebuildSurf) { a->surf.trees[im].build(boxes); a->surf.built[im] = 1; }
        else a->surf.trees[im].refit(boxes);
    }
}
static void update_ccd_bvh(Scene& s, bool rebuild) {
    ColAccel* a = (ColAccel*)s.bvhTet;
    std::vector<Box> boxes;
    ccd_boxes(s, a->ccd, boxes);
    bool ch = activity_changed(s, a->ccd.activeAtBuild);
    if (!a->ccd.built || rebuild || ch) { a->ccd.tree.build(boxes); a->ccd.built = true; }
    else a->ccd.tree.refit(boxes);
}

// DiscreteCollisionDetector::vertexCollisionDetection (rule R1)
static void vertex_collision_detection(Scene& s, int vId, int meshId, ColResult& r) {
    ColAccel* a = (ColAccel*)s.bvhTet;
    const Mesh& mq = s.meshes[meshId];
    const V3 p = mq.pos[vId];
    r.clear();
    r.idTMQuery = meshId; r.idVQuery = vId; r.fromCCD = false;
    int bestMesh = -1, bestTet = -1;
    auto visitTet = [&](int prim) {
        const int g = a->tet.primMesh[prim], t = a->tet.primTet[prim];
        if (!s.col.handleSelfCollision && g == meshId) return;
        const Mesh& mi = s.meshes[g];
        const int* tv = &mi.tets[4 * (size_t)t];
        if (g == meshId) for (int k = 0; k < 4; k++) if (tv[k] == vId) return;
        if (bestMesh != -1 && !(g < bestMesh || (g == bestMesh && t < bestTet))) return;
        if (point_in_tet(p, mi.pos[tv[0]], mi.pos[tv[1]], mi.pos[tv[2]], mi.pos[tv[3]])) { bestMesh = g; bestTet = t; }
    };
    if (a->bruteForce) { for (size_t i = 0; i < a->tet.primTet.size(); i++) if (s.meshes[a->tet.primMesh[i]].activeForCollision) visitTet((int)i); }
    else a->tet.tree.query(p, [] { return 0.f; }, visitTet);
    if (bestMesh >= 0) { r.numIntersections = 1; r.intersectedMeshId = bestMesh; r.intersectedElement = bestTet; }
}

// closestPointQueryFunc evaluated for one candidate face: returns true when the face is an acceptable closest
// face (feasible region + tetrahedral traversal), with its distance/point/barycentrics/type
static bool closest_candidate(const Scene& s, const Mesh& ms, int meshSearch, int face, int meshQuery, int vQuery, V3 q,
                              int embraceTet, float curRadius, int curFace, float& dOut, V3& cpOut, V3& baryOut, int& typeOut) {
    const int* fv = &ms.surfFaces[3 * face];
    const V3 a = ms.pos[fv[0]], b = ms.pos[fv[1]], c = ms.pos[fv[2]];
    int type; V3 bary;
    const V3 cp = closest_point_triangle(q, a, b, c, bary, type);
    const float d = norm(q - cp);
    if (meshSearch == meshQuery) {
        int cpId = -1;
        if (type == AtA) cpId = fv[0]; else if (type == AtB) cpId = fv[1]; else if (type == AtC) cpId = fv[2];
        if (cpId != -1 && cpId == vQuery) return false;
    }
    // rule R2: strictly closer, or equally close with a smaller face id
    if (!(d < curRadius || (d == curRadius && curFace >= 0 && face < curFace))) return false;
    if (s.col.checkFeasibleRegion && !check_feasible_region(q, ms, face, type, s.col.feasibleRegionEpsilon)) return false;
    if (s.col.checkTetTraverse && (meshSearch == meshQuery || s.col.tetrahedralTraverseForNonSelfIntersection)) {
        V3 tracing;
        if (type == AtInterior) tracing = cp;
        else tracing = cp * (1.f - s.col.centerShiftLevel) + (s.col.centerShiftLevel / 3.0f) * ((a + b) + c);
        V3 target;
        if (s.col.shiftQueryPointToCenter) {
            const int* tv = &ms.tets[4 * (size_t)embraceTet];
            V3 cen = (((V3() + ms.pos[tv[0]]) + ms.pos[tv[1]]) + ms.pos[tv[2]]) + ms.pos[tv[3]];
            cen = cen * 0.25f;
            target = (1.f - s.col.centerShiftLevel) * q + s.col.centerShiftLevel * cen;
        } else target = q;
        V3 dir = target - tracing;
        const float len = norm(dir);
        dir = dir / len;
        const float maxDis = s.col.stopTraversingAfterPassingQueryPoint ? s.col.maxSearchDistanceMultiplier * len : -1.f;
        if (!tetrahedral_traverse_to(ms, tracing, dir, maxDis, ms.surfFaceTet[face], ms.surfFaceIdAtTet[face], embraceTet,
                                     s.col.rayTriIntersectionEPSILON, nullptr)) return false;
    }
    dOut = d; cpOut = cp; baryOut = bary; typeOut = type;
    return true;
}

// DiscreteCollisionDetector::closestPointQuery (rule R2)
static void closest_point_query(Scene& s, ColResult& r) {
    ColAccel* acc = (ColAccel*)s.bvhTet;
    const Mesh& mq = s.meshes[r.idTMQuery];
    const V3 q = mq.pos[r.idVQuery];
    const int g = r.intersectedMeshId;
    const Mesh& ms = s.meshes[g];
    float radius = std::numeric_limits<float>::infinity();
    int bestFace = -1; V3 bestCp, bestBary; int bestType = NotFound;
    auto visit = [&](int face) {
        float d; V3 cp, bary; int type;
        if (closest_candidate(s, ms, g, face, r.idTMQuery, r.idVQuery, q, r.intersectedElement, radius, bestFace, d, cp, bary, type)) {
            radius = d; bestFace = face; bestCp = cp; bestBary = bary; bestType = type;
        }
    };
    if (acc->bruteForce) { for (int f = 0; f < ms.nSurfF(); f++) visit(f); }
    else acc->surf.trees[g].query(q, [&] { return radius == std::numeric_limits<float>::infinity() ? 1e30f : radius * radius * 1.0001f + 1e-12f; }, visit);
    // rule R2' (stand-in for maxNumberOfBVHQuery, DiscreteCollisionDetector.cpp:104-110: the reference gives up - found = false -
    // once its callback has run more than maxNumberOfBVHQuery times; which faces those are depends on Embree's order). The
    // order-independent statement: the query fails when at least maxNumberOfBVHQuery faces of the mesh are strictly nearer
    // (exact closest-point distance) than the nearest acceptable one - a nearest-first search would have spent its budget on them.
    if (bestFace >= 0 && s.col.maxNumberOfBVHQuery > 0 && ms.nSurfF() >= s.col.maxNumberOfBVHQuery) {
        int nearer = 0;
        for (int f = 0; f < ms.nSurfF() && nearer < s.col.maxNumberOfBVHQuery; f++) {
            const int* fv = &ms.surfFaces[3 * f];
            int type; V3 bary;
            const V3 cp = closest_point_triangle(q, ms.pos[fv[0]], ms.pos[fv[1]], ms.pos[fv[2]], bary, type);
            if (norm(q - cp) < radius) nearer++;
        }
        if (nearer >= s.col.maxNumberOfBVHQuery) bestFace = -1;
    }
    if (bestFace >= 0) {
        r.shortestPathFound = true; r.closestSurfacePtBarycentrics = bestBary; r.closestSurfacePt = bestCp;
        r.closestSurfaceFaceId = bestFace; r.closestPointType = bestType; r.closestPointNormal = V3();
    } else {
        r.shortestPathFound = false; r.closestSurfacePtBarycentrics = V3(-1, -1, -1); r.closestSurfacePt = V3(-1, -1, -1);
        r.closestSurfaceFaceId = -1; r.closestPointType = NotFound; r.closestPointNormal = V3();
    }
}

// ContinuousCollisionDetector::vertexContinuousCollisionDetection (rule R3)
static void vertex_ccd(Scene& s, int vId, int meshId, ColResult& r) {
    ColAccel* acc = (ColAccel*)s.bvhTet;
    const Mesh& mq = s.meshes[meshId];
    r.clear();
    r.idVQuery = vId; r.idTMQuery = meshId; r.fromCCD = true;
    const V3 p1 = mq.pos[vId], p0 = mq.prev[vId];
    const V3 mid = 0.5f * (p1 + p0);
    const float radius = norm(p1 - p0) * 0.5f;
    float bestDepth = 0.f; int bestMesh = -1, bestFace = -1; V3 bestBary; float bestT = 0;
    const float r2 = radius * radius * 1.0001f + 1e-12f;
    auto visit = [&](int prim) {
        const int g = acc->ccd.primMesh[prim], f = acc->ccd.primFace[prim];
        const Mesh& mi = s.meshes[g];
        if (!mi.activeForCollision) return;
        const int* fv = &mi.surfFaces[3 * f];
        if (g == meshId) for (int k = 0; k < 3; k++) if (fv[k] == vId) return;
        ccdm::Vec3 x[2][3], p[2];
        for (int k = 0; k < 3; k++) {
            x[0][k] = ccdm::Vec3(mi.prev[fv[k]].x, mi.prev[fv[k]].y, mi.prev[fv[k]].z);
            x[1][k] = ccdm::Vec3(mi.pos[fv[k]].x, mi.pos[fv[k]].y, mi.pos[fv[k]].z);
        }
        p[0] = ccdm::Vec3(p0.x, p0.y, p0.z); p[1] = ccdm::Vec3(p1.x, p1.y, p1.z);
        float tt = -1; ccdm::Vec3 bary;
        if (ccdm::IntersectContinuousTriPoint(tt, x, p, bary)) {
            const float depth = tt * norm(p1 - p0);
            bool better = bestMesh < 0 || depth < bestDepth || (depth == bestDepth && (g < bestMesh || (g == bestMesh && f < bestFace)));
            if (better) { bestDepth = depth; bestMesh = g; bestFace = f; bestBary = V3(bary.x, bary.y, bary.z); bestT = tt; }
        }
    };
    if (acc->bruteForce) { for (size_t i = 0; i < acc->ccd.primFace.size(); i++) visit((int)i); }
    else acc->ccd.tree.query(mid, [&] { return r2; }, visit);
    if (bestMesh >= 0) {
        r.numIntersections = 1;
        r.closestSurfaceFaceId = bestFace; r.intersectedMeshId = bestMesh; r.shortestPathFound = true;
        r.closestSurfacePtBarycentrics = bestBary;
        // (1. - tt) * prev + tt * pos (ContinuousCollisionDetector.cpp:146-147); Eigen casts the double scalar to float
        const float w0 = (float)(1. - (double)bestT);
        r.closestSurfacePt = w0 * p0 + bestT * p1;
        r.closestPointType = AtInterior; r.penetrationDepth = bestDepth; r.closestPointNormal = V3();
    }
}

void dcd(Scene& s, bool intermediate) {
    if (s.col.allowDCD) {
        bool rebuildTet = false, rebuildSurf = false;
        if (!intermediate) {
            rebuildTet = (s.substep == 0) && !(s.frameId % s.phys.dcdTetMeshSceneBVHRebuildSteps) && s.frameId;
            rebuildSurf = (s.substep == 0) && !(s.frameId % s.phys.dcdSurfaceSceneBVHRebuildSteps) && s.frameId;
        }
        s.ev(12, rebuildTet, rebuildSurf);
        update_dcd_bvh(s, rebuildTet, rebuildSurf);
        for (size_t im = 0; im < s.meshes.size(); im++) {
            Mesh& m = s.meshes[im];
            if (!m.activeForCollision) continue;
            if (!intermediate) { s.ev(18, (int)im); std::fill(m.penetratedMask.begin(), m.penetratedMask.end(), 0); }
            s.pool->parallel_for(m.nSurfV(), [&](int64_t b, int64_t e) {
                for (int64_t sv = b; sv < e; sv++) {
                    const int v = m.surfaceVIds[sv];
                    bool doIt = intermediate ? (m.penetratedMask[v] || !s.col.allowCCD)
                                             : (m.pos[v][GRAVITY_AXIS] < s.phys.collisionOffHeight);
                    if (!doIt) continue;
                    ColResult& r = m.colResults[sv];
                    vertex_collision_detection(s, v, (int)im, r);
                    s.ev(13, (int)im, v, r.numIntersections);
                    if (r.numIntersections) {
                        if (!intermediate) m.penetratedMask[v] = 1;
                        s.ev(14, (int)im, v);
                        closest_point_query(s, r);
                    }
                }
            }, 32);
        }
    } else if (s.col.allowCCD && !intermediate) {
        for (size_t im = 0; im < s.meshes.size(); im++) for (size_t sv = 0; sv < s.meshes[im].colResults.size(); sv++) { s.ev(17, (int)im, (int)sv); s.meshes[im].colResults[sv].clear(); }
    }
}

// ccd() (M/VBD/VBDPhysics.cpp:2077-2115) and, with intermediate, intermediateCCD() (:2117-2156: the same loop, never a rebuild)
void ccd(Scene& s, bool intermediate) {
    if (!s.col.allowCCD) return;
    const bool rebuild = !intermediate && (s.substep == 0) && !(s.frameId % s.phys.ccdBVHRebuildSteps) && s.frameId;
    s.ev(15, rebuild);
    update_ccd_bvh(s, rebuild);
    for (size_t im = 0; im < s.meshes.size(); im++) {
        Mesh& m = s.meshes[im];
        if (!m.activeForCollision) continue;
        s.pool->parallel_for(m.nSurfV(), [&](int64_t b, int64_t e) {
            for (int64_t sv = b; sv < e; sv++) {
                const int v = m.surfaceVIds[sv];
                if (m.pos[v][GRAVITY_AXIS] < s.phys.collisionOffHeight && !m.penetratedMask[v]) {
                    s.ev(16, (int)im, v);
                    vertex_ccd(s, v, (int)im, m.colResults[sv]);
                }
            }
        }, 32);
    }
}

void prepare_collision_data(Scene& s) {
    for (auto& g : s.activeCollisionsEachGroup) g.clear();
    for (auto& m : s.meshes) for (auto& rl : m.colRelations) rl.clear();
    s.stats.numContactsDCD = 0; s.stats.numContactsCCD = 0;
    for (size_t im = 0; im < s.meshes.size(); im++) {
        Mesh& m = s.meshes[im];
        std::fill(m.activeCollisionMask.begin(), m.activeCollisionMask.end(), 0);
    }
    for (size_t im = 0; im < s.meshes.size(); im++) {
        Mesh& m = s.meshes[im];
        for (int sv = 0; sv < m.nSurfV(); sv++) {
            ColResult& r = m.colResults[sv];
            const int v = m.surfaceVIds[sv];
            if (!r.numIntersections) continue;
            // addToActiveCollisionList
            if (r.shortestPathFound) {
                int added[4]; int nAdded = 0;
                const int gv = s.meshes[r.idTMQuery].vertexParallelGroup[r.idVQuery];
                s.activeCollisionsEachGroup[gv].push_back(r.idTMQuery); s.activeCollisionsEachGroup[gv].push_back(r.idVQuery);
                added[nAdded++] = gv;
                Mesh& mi = s.meshes[r.intersectedMeshId];
                for (int k = 0; k < 3; k++) {
                    const int fvId = mi.surfFaces[3 * r.closestSurfaceFaceId + k];
                    const int gf = mi.vertexParallelGroup[fvId];
                    bool has = false; for (int q = 0; q < nAdded; q++) if (added[q] == gf) has = true;
                    if (!has) {
                        s.activeCollisionsEachGroup[gf].push_back(r.idTMQuery); s.activeCollisionsEachGroup[gf].push_back(r.idVQuery);
                        added[nAdded++] = gf;
                    }
                }
            }
            if (r.fromCCD) s.stats.numContactsCCD++; else s.stats.numContactsDCD++;
            if (r.shortestPathFound) {
                m.activeCollisionMask[v] = 1;
                Mesh& mi = s.meshes[r.intersectedMeshId];
                for (int k = 0; k < 3; k++) {
                    const int fvId = mi.surfFaces[3 * r.closestSurfaceFaceId + k];
                    mi.activeCollisionMask[fvId] = 1;
                    mi.colRelations[fvId].push_back({(int)im, sv, 0, 0, k});
                }
            }
        }
    }
}

void update_collision_info(Scene& s, ColResult& r) {
    if (!r.numIntersections || !r.shortestPathFound) return;
    Mesh& cur = s.meshes[r.idTMQuery];
    Mesh& hit = s.meshes[r.intersectedMeshId];
    const int f = r.closestSurfaceFaceId;
    const int t0 = hit.surfFaces[3 * f], t1 = hit.surfFaces[3 * f + 1], t2 = hit.surfFaces[3 * f + 2];
    const V3 a = hit.pos[t0], b = hit.pos[t1], c = hit.pos[t2], p = cur.pos[r.idVQuery];
    V3 bary; int type;
    const V3 cp = closest_point_triangle(p, a, b, c, bary, type);
    r.closestSurfacePt = cp; r.closestSurfacePtBarycentrics = bary;
    V3 faceN = cross(b - a, c - a); faceN = faceN / norm(faceN);   // TetMeshFEM::computeFaceNormal
    if (type == AtInterior) r.closestPointNormal = faceN;
    else if (type != NotFound) {
        V3 d = cp - p;
        d = d / norm(d);
        if (dot(d, faceN) < 0) d = -d;
        r.closestPointNormal = d;
    }
    const V3 x = p;
    const V3 dx3 = x - cur.prev[r.idVQuery];
    const V3 dx0 = hit.pos[t0] - hit.prev[t0], dx1 = hit.pos[t1] - hit.prev[t1], dx2 = hit.pos[t2] - hit.prev[t2];
    const V3 dx = dx3 - ((bary.x * dx0 + bary.y * dx1) + bary.z * dx2);
    V3 e0 = hit.pos[t1] - hit.pos[t0]; e0 = e0 / norm(e0);
    const V3 n = r.closestPointNormal;
    V3 e1 = cross(e0, n); e1 = e1 / norm(e1);
    const float T[6] = {e0.x, e0.y, e0.z, e1.x, e1.y, e1.z};
    const float u[2] = {dot(e0, dx), dot(e1, dx)};
    // (a + b) * 0.5 with a double literal (VBDPhysics.cpp:2287-2288)
    const float mu = (float)(((double)(hit.obj.frictionDynamic + cur.obj.frictionDynamic)) * 0.5);
    const float epsV = (float)(((double)(hit.obj.frictionEpsV + cur.obj.frictionEpsV)) * 0.5);
    const float epsU = epsV * s.dt;
    const V3 diff = r.closestSurfacePt - x;
    r.penetrationDepth = dot(diff, n);
    const float k = s.phys.collisionStiffness;
    const float lambda = r.penetrationDepth * k;
    r.collisionForce = V3(); r.collisionHessian = M3();
    if (r.penetrationDepth > 0.f) {
        vertex_friction(mu, lambda, T, u, epsU, r.collisionForce, r.collisionHessian);
        r.collisionForce += lambda * n;
        for (int rr = 0; rr < 3; rr++) for (int cc = 0; cc < 3; cc++) r.collisionHessian(rr, cc) += (k * n[rr]) * n[cc];
    }
}

// ---------------------------------------------------------------- Morton leaf order (integer parity surface)
// The product's LBVH sorts primitives by (segment, 30-bit Morton code of the primitive's box centre); codes follow
// lbvh::morton_code (M/BVH/MortonCode.h:8-31), the centre is normalised by the bounds of the active meshes' vertices (scene
// trees) or of the primitive's own mesh (per-mesh surface trees); ties keep ascending primitive id (stable sort).
static inline uint32_t expand_bits(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
static uint32_t morton_of(const Box& b, const Box& nb) {
    uint32_t e[3];
    for (int d = 0; d < 3; d++) {
        const float c = 0.5f * (b.lo[d] + b.hi[d]);
        const float ext = nb.hi[d] - nb.lo[d];
        float u = ext > 0.f ? (c - nb.lo[d]) / ext : 0.f;
        u = std::fmin(std::fmax(u * 1024.0f, 0.0f), 1023.0f);
        e[d] = expand_bits((uint32_t)u);
    }
    return e[0] * 4 + e[1] * 2 + e[2];
}
std::vector<int> morton_leaf_order(const Scene& s, int which) {
    std::vector<Box> meshBox(s.meshes.size());
    Box scene; for (int d = 0; d < 3; d++) { scene.lo[d] = 3.0e38f; scene.hi[d] = -3.0e38f; }
    for (size_t im = 0; im < s.meshes.size(); im++) {
        Box b; for (int d = 0; d < 3; d++) { b.lo[d] = 3.0e38f; b.hi[d] = -3.0e38f; }
        for (const V3& p : s.meshes[im].pos) box_add(b, p);
        meshBox[im] = b;
        if (s.meshes[im].activeForCollision) box_merge(scene, b);
    }
    std::vector<uint64_t> keys;
    for (size_t im = 0; im < s.meshes.size(); im++) {
        const Mesh& m = s.meshes[im];
        const int n = which == 0 ? m.nT : m.nSurfF();
        for (int i = 0; i < n; i++) {
            const uint64_t seg = which == 1 ? ((uint64_t)im << 32) : 0ull;
            if (!m.activeForCollision) { keys.push_back(seg | (1ull << 30)); continue; }
            Box b; for (int d = 0; d < 3; d++) { b.lo[d] = 3e38f; b.hi[d] = -3e38f; }
            if (which == 0) for (int k = 0; k < 4; k++) box_add(b, m.pos[m.tets[4 * (size_t)i + k]]);
            else for (int k = 0; k < 3; k++) {
                box_add(b, m.pos[m.surfFaces[3 * (size_t)i + k]]);
            }
            if (which == 2) for (int k = 0; k < 3; k++) box_add(b, m.prev[m.surfFaces[3 * (size_t)i + k]]);
            keys.push_back(seg | morton_of(b, which == 1 ? meshBox[im] : scene));
        }
    }
    std::vector<int> order(keys.size());
    for (size_t i = 0; i < order.size(); i++) order[i] = (int)i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return keys[a] < keys[b]; });
    return order;
}

// test hook: switch the accelerator off (brute force over all primitives)
void collision_set_brute_force(Scene& s, bool on) { if (s.bvhTet) ((ColAccel*)s.bvhTet)->bruteForce = on; }

}  // namespace orc
