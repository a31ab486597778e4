// The VBD cloth step on the GPU behind include/gaia_b200_cloth.h. Replaces
//   VBDClothSimulationFramework::runStep_CDEveryIter                       M/VBDCloth/VBDClothPhysics.cpp:389-518
//   VBDStepWithExistingCollisions / evaluateVertexForceAndHessian          M/VBDCloth/VBDClothPhysics.cpp:1026-1221
//   VBDTriMeshStVk::accumlateStVKForceAndHessian / accumlateBending...     M/VBDCloth/VBDTriMeshStVK.cpp:721-982
//   accumulateVF/EEContactForceAndHessian, computeFriction, computeContactRepulsiveForce, accumlateBoundary...
//                                                                          M/VBDCloth/VBDClothPhysics.cpp:1248-1579
//   applyCollisionDetection, applyConservativeBoundTruncation, applyAcceleration, updateVelocity
//                                                                          M/VBDCloth/VBDClothPhysics.cpp:735-820, 1223-1246, 893-915, 2318-2359
//   VBDBaseTriMesh::applyInitialStep / forwardStep_intertia                M/VBDCloth/VBDBaseTriMesh.cpp:60-182
//   ClothContactDetector::contactQueryVF / contactQueryEE + callbacks      M/CollisionDetector/ClothContactDetector.cpp:144-288, 731-813, 1052-1135
//   checkFeasibleRegion (edge / vertex), TriMesh                           M/CollisionDetector/TriMeshCollisionGeometry.cpp:7-184
//   get_closest_points_between_segments                                    M/CollisionDetector/CollisionGeometry.h:160-267
// The Embree radius queries become two LBVHs (faces, edges: Morton sort, Karras build, refit at every detection).
//
// Kernels (all HBM-bound gather/scatter; no tensor cores):
//   k_solve_colour    one thread per vertex of the colour; incident-face and hinge records come from the per-colour ELL
//                     streams (coalesced 16-byte loads), neighbour positions are gathered (L2 resident), 3x3 solve,
//                     conservative-bound truncation, in-place position store
//   k_contact_terms   per colour, before the sweep: contact + friction force/Hessian of every vertex that has contacts
//                     (Jacobi within the colour: the reference writes to vertexPositionNext and copies after the sweep)
//   k_vf_query / k_ee_query   stack traversals of the two LBVHs, one query primitive per thread
//
// Canonical rules shared with the CPU path: R5 contact lists are sorted by (mesh, primitive id) and the
// face representing a shared closest primitive is the feasible one with the smallest id; R6 re-detection is requested when
// ANY vertex was truncated since the last detection; R7 per-vertex lists are walked in ascending id.
//
// Compiled with -fmad=false: every predicate and every sum is the same sequence of IEEE single operations as the CPU
// path, so contact sets are bit-identical and positions agree to rounding of the few library calls that differ.
#include "cloth_scene.h"
#include "scene_impl.h"   // DevBuf
#include "geom.cuh"       // V3, closest_point_triangle
#include <cub/cub.cuh>
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>

using gb::DevBuf;
using gb::V3;
using gb::v3;

namespace gbc {

#define CK(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return _e; } while (0)
static const int kB = 128;
static inline int nblk(int64_t n, int per = kB) { return (int)((n + per - 1) / per); }

static const int KV = 24;      // VF contacts kept per vertex (the reference pre-allocates 8 and grows); more -> GAIA_ERR_OVERFLOW
static const int KE = 64;      // EE contacts kept per edge (coarse colliders put ~30 edges within reach of one cloth edge)
static const int kStack = 64;
enum { ERR_VF = 0, ERR_EE, ERR_STACK, ERR_LINK, FLAG_TRUNC, ERR_SOLVER, ERR_NAN, N_FLAGS = 8 };

struct MeshDev {
    int32_t vBase, fBase, eBase, simulated;
    float miu, lambda, dampStVK, kBend, dampBend, frictionDynamic, frictionEpsV;
    float expDamp, constDamp, maxVel, noGravThr;
    int32_t hasNoGravZone, initType;
};

struct StepP {
    float dt, dtSqrRec;
    float g[3], bmin[3], bmax[3];
    float kBoundary, muBoundary, epsVBoundary;
    float stepSize, kContact, radius, thickness, relax, R, degThres2;
    int32_t energyType, handleCollision, assocGravity, numSim;
};

// ------------------------------------------------------------------------------------------------ small vector helpers
#define CDEV __device__ __forceinline__
CDEV void operator+=(V3& a, V3 b) { a = a + b; }
CDEV void operator-=(V3& a, V3 b) { a = a - b; }
CDEV V3 ld3(const float4* p, int i) { const float4 q = p[i]; return v3(q.x, q.y, q.z); }
CDEV V3 normalized(V3 v) { const float n2 = gb::sqnorm(v); return n2 > 0.f ? v / sqrtf(n2) : v; }
CDEV void add_outer(float* H, float s, V3 n) {   // H(r,c) += s * n[r] * n[c], column-major
    const float v[3] = {n.x, n.y, n.z};
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int r = 0; r < 3; r++) H[r + 3 * c] += s * v[r] * v[c];
}
CDEV V3 mul3(const float* A, V3 x) {
    return v3(A[0] * x.x + A[3] * x.y + A[6] * x.z, A[1] * x.x + A[4] * x.y + A[7] * x.z, A[2] * x.x + A[5] * x.y + A[8] * x.z);
}

// CuMatrix::solve3x3_psd_stable (3P/CuMatrix/CuMatrix/MatrixOps/CuMatrix.h:235-267)
CDEV bool solve3x3(const float* m, V3 b, V3& out) {
    const float a11 = m[0], a12 = m[3], a13 = m[6], a21 = m[1], a22 = m[4], a23 = m[7], a31 = m[2], a32 = m[5], a33 = m[8];
    const float i11 = a33 * a22 - a32 * a23;
    const float i12 = -(a33 * a12 - a32 * a13);
    const float i13 = a23 * a12 - a22 * a13;
    const float det = (a11 * i11 + a21 * i12 + a31 * i13);
    if (fabsf(det) < 1e-5f * (fabsf(a11 * i11) + fabsf(a21 * i12) + fabsf(a31 * i13))) { out = b; return false; }
    const float deti = (float)(1.0 / (double)det);
    const float i21 = -(a33 * a21 - a31 * a23), i22 = a33 * a11 - a31 * a13, i23 = -(a23 * a11 - a21 * a13);
    const float i31 = a32 * a21 - a31 * a22, i32 = -(a32 * a11 - a31 * a12), i33 = a22 * a11 - a21 * a12;
    out = v3(deti * (i11 * b.x + i12 * b.y + i13 * b.z), deti * (i21 * b.x + i22 * b.y + i23 * b.z), deti * (i31 * b.x + i32 * b.y + i33 * b.z));
    return true;
}

// computeFriction (VBDClothPhysics.cpp:1500-1531)
CDEV void friction(float mu, float lambda, V3 t0, V3 t1, float u0, float u1, float epsU, V3& force, float* H) {
    const float uNorm = sqrtf(u0 * u0 + u1 * u1);
    force = v3(0, 0, 0);
#pragma unroll
    for (int k = 0; k < 9; k++) H[k] = 0.f;
    if (uNorm > 0) {
        float f1;
        if (uNorm > epsU) f1 = 1 / uNorm;
        else f1 = (-uNorm / epsU + 2) / epsU;
        force = (-mu * lambda) * ((f1 * u0) * t0 + (f1 * u1) * t1);
        add_outer(H, mu * lambda * f1, t0);
        add_outer(H, mu * lambda * f1, t1);
    }
}

// computeContactRepulsiveForce (VBDClothPhysics.cpp:1533-1579)
CDEV void repulsion(const StepP& P, float dis_, float& dEdD, float& d2E) {
    const float k = P.kContact;
    const float disAdjusted = dis_ - P.thickness;
    const float penetrationDepth = P.radius - disAdjusted;
    dEdD = 0.f; d2E = 0.f;
    if (P.energyType == 0) { dEdD = -k * penetrationDepth; d2E = k; }
    else if (P.energyType == 1) {
        const float tau = P.radius * 0.5f;
        if (disAdjusted < tau && disAdjusted > 1e-5f) {
            const float k2 = 0.5f * (tau * tau) * k;
            dEdD = -k2 / disAdjusted; d2E = k2 / (disAdjusted * disAdjusted);
        } else { dEdD = -k * penetrationDepth; d2E = k; }
    } else if (P.energyType == 2) {
        const float dr = disAdjusted - P.radius;
        dEdD = (-k * (dr * dr) / disAdjusted - 2 * k * logf(disAdjusted / P.radius) * dr);
        d2E = k * (((dr * dr) / (disAdjusted * disAdjusted)) - (4 * disAdjusted - 4 * P.radius) / disAdjusted - 2 * logf(disAdjusted / P.radius));
    }
}

// get_closest_points_between_segments (CollisionGeometry.h:160-267)
CDEV void closest_points_segments(V3 p0, V3 p1, V3 q0, V3 q1, V3& ps, V3& qt, float& s, float& t) {
    const V3 p = p1 - p0, q = q1 - q0, r = p0 - q0;
    const float a = gb::dot(p, p), b = gb::dot(p, q), c = gb::dot(q, q), d = gb::dot(p, r), e = gb::dot(q, r);
    s = 0.0f; t = 0.0f;
    const float det = a * c - b * b;
    if (det > 1e-5f) {
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

// ------------------------------------------------------------------------------------------------ LBVH
// Node index space: internal nodes [0,n), leaf at sorted position j -> n + j; root = 0 (n >= 2) or the single leaf.
struct Bvh {
    int n = 0;
    DevBuf<float4> primLo, primHi, lo, hi;
    DevBuf<uint32_t> keyIn, keyOut;
    DevBuf<int32_t> idIn, order, parent, flags;
    DevBuf<int2> child;
    DevBuf<uint8_t> tmp;
    size_t tmpBytes = 0;
    bool built = false;
    int detectionsSinceBuild = 0;
    cudaError_t alloc(int n_) {
        n = n_;
        if (!n) return cudaSuccess;
        CK(primLo.alloc(n)); CK(primHi.alloc(n)); CK(lo.alloc(2 * (size_t)n)); CK(hi.alloc(2 * (size_t)n));
        CK(keyIn.alloc(n)); CK(keyOut.alloc(n)); CK(idIn.alloc(n)); CK(order.alloc(n)); CK(parent.alloc(2 * (size_t)n));
        CK(flags.alloc(n)); CK(child.alloc(n));
        cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, keyIn.p, keyOut.p, idIn.p, order.p, n);
        CK(tmp.alloc(tmpBytes + 16));
        std::vector<int32_t> iota(n);
        for (int i = 0; i < n; i++) iota[i] = i;
        return cudaMemcpy(idIn.p, iota.data(), (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice);
    }
};
struct BvhView { int n; const int2* child; const float4 *lo, *hi; const int32_t* order; };
static BvhView view_of(const Bvh& b) { BvhView v; v.n = b.n; v.child = b.child.p; v.lo = b.lo.p; v.hi = b.hi.p; v.order = b.order.p; return v; }

CDEV int delta_key(const uint32_t* __restrict__ keys, int i, int j, int n) {
    if (j < 0 || j >= n) return -1;
    const uint32_t a = keys[i], b = keys[j];
    if (a == b) return 32 + __clz((unsigned)i ^ (unsigned)j);
    return __clz(a ^ b);
}
// Karras 2012: one thread per internal node
static __global__ void k_cloth_karras(int n, const uint32_t* __restrict__ keys, int2* __restrict__ child, int32_t* __restrict__ parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) { if (i == n - 1 && n == 1) parent[n] = -1; return; }
    if (i == 0) parent[0] = -1;
    const int d = (delta_key(keys, i, i + 1, n) - delta_key(keys, i, i - 1, n)) >= 0 ? 1 : -1;
    const int dmin = delta_key(keys, i, i - d, n);
    int lmax = 2;
    while (delta_key(keys, i, i + lmax * d, n) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta_key(keys, i, i + (l + t) * d, n) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta_key(keys, i, j, n);
    int s = 0, t = l;
    do {
        t = (t + 1) / 2;
        if (delta_key(keys, i, i + (s + t) * d, n) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int left = (min(i, j) == gamma) ? n + gamma : gamma;
    const int right = (max(i, j) == gamma + 1) ? n + gamma + 1 : gamma + 1;
    child[i] = make_int2(left, right);
    parent[left] = i;
    parent[right] = i;
}
// bottom-up refit: the second thread to reach a node merges its children
static __global__ void k_cloth_refit(int n, const int32_t* __restrict__ order, const float4* __restrict__ primLo, const float4* __restrict__ primHi,
                                     const int2* __restrict__ child, const int32_t* __restrict__ parent, float4* lo, float4* hi, int32_t* flags) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int prim = order[j];
    lo[n + j] = primLo[prim]; hi[n + j] = primHi[prim];
    int cur = parent[n + j];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(flags + cur, 1) == 0) return;
        const int2 c = child[cur];
        const float4 a0 = __ldcg(lo + c.x), a1 = __ldcg(hi + c.x), b0 = __ldcg(lo + c.y), b1 = __ldcg(hi + c.y);
        lo[cur] = make_float4(fminf(a0.x, b0.x), fminf(a0.y, b0.y), fminf(a0.z, b0.z), 0.f);
        hi[cur] = make_float4(fmaxf(a1.x, b1.x), fmaxf(a1.y, b1.y), fmaxf(a1.z, b1.z), 0.f);
        cur = parent[cur];
    }
}

CDEV int f2ord(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7FFFFFFF; }
CDEV float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }
static __global__ void k_cloth_init_bounds(int32_t* b) { if (threadIdx.x < 6) b[threadIdx.x] = threadIdx.x < 3 ? f2ord(3.0e38f) : f2ord(-3.0e38f); }
static __global__ void k_cloth_scene_bounds(int n, const float4* __restrict__ pos, int32_t* bounds) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    const float4 p = pos[min(s, n - 1)];
    float lx = p.x, ly = p.y, lz = p.z, hx = p.x, hy = p.y, hz = p.z;
    for (int o = 16; o > 0; o >>= 1) {
        lx = fminf(lx, __shfl_xor_sync(0xffffffffu, lx, o)); ly = fminf(ly, __shfl_xor_sync(0xffffffffu, ly, o)); lz = fminf(lz, __shfl_xor_sync(0xffffffffu, lz, o));
        hx = fmaxf(hx, __shfl_xor_sync(0xffffffffu, hx, o)); hy = fmaxf(hy, __shfl_xor_sync(0xffffffffu, hy, o)); hz = fmaxf(hz, __shfl_xor_sync(0xffffffffu, hz, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(bounds + 0, f2ord(lx)); atomicMin(bounds + 1, f2ord(ly)); atomicMin(bounds + 2, f2ord(lz));
        atomicMax(bounds + 3, f2ord(hx)); atomicMax(bounds + 4, f2ord(hy)); atomicMax(bounds + 5, f2ord(hz));
    }
}
CDEV uint32_t expand_bits(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
// 30-bit Morton code of a box centre normalised by the scene box (M/BVH/MortonCode.h:8-31 semantics)
CDEV uint32_t morton_of(V3 lo, V3 hi, const int32_t* __restrict__ boundsOrd) {
    const float c[3] = {0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z)};
    uint32_t e[3];
#pragma unroll
    for (int d = 0; d < 3; d++) {
        const float b0 = ord2f(boundsOrd[d]), ext = ord2f(boundsOrd[3 + d]) - b0;
        float u = ext > 0.f ? (c[d] - b0) / ext : 0.f;
        u = fminf(fmaxf(u * 1024.0f, 0.0f), 1023.0f);
        e[d] = expand_bits((uint32_t)u);
    }
    return e[0] * 4 + e[1] * 2 + e[2];
}
static const float kInflate = 1e-4f;   // conservative culling only: results equal brute force
CDEV void store_box(V3 lo, V3 hi, float4* outLo, float4* outHi) {
    const float mx = kInflate * (1.f + fmaxf(fabsf(lo.x), fabsf(hi.x)));
    const float my = kInflate * (1.f + fmaxf(fabsf(lo.y), fabsf(hi.y)));
    const float mz = kInflate * (1.f + fmaxf(fabsf(lo.z), fabsf(hi.z)));
    *outLo = make_float4(lo.x - mx, lo.y - my, lo.z - mz, 0.f);
    *outHi = make_float4(hi.x + mx, hi.y + my, hi.z + mz, 0.f);
}
CDEV void grow(V3& lo, V3& hi, V3 p) {
    lo = v3(fminf(lo.x, p.x), fminf(lo.y, p.y), fminf(lo.z, p.z));
    hi = v3(fmaxf(hi.x, p.x), fmaxf(hi.y, p.y), fmaxf(hi.z, p.z));
}
// NV = 3: faces (int4 = three sorted vertex ids, mesh); NV = 2: edges (int4 = eV1, eV2, eV12Next, eV21Next)
template <int NV>
static __global__ void k_cloth_prim_boxes(int n, const int4* __restrict__ prims, const float4* __restrict__ pos, const int32_t* __restrict__ boundsOrd,
                                          int writeKeys, float4* __restrict__ primLo, float4* __restrict__ primHi, uint32_t* __restrict__ keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int4 v = __ldg(prims + i);
    V3 lo = ld3(pos, v.x), hi = lo;
    grow(lo, hi, ld3(pos, v.y));
    if (NV == 3) grow(lo, hi, ld3(pos, v.z));
    if (writeKeys) keys[i] = morton_of(lo, hi, boundsOrd);
    store_box(lo, hi, primLo + i, primHi + i);
}

// ------------------------------------------------------------------------------------------------ device scene view
struct DevView {
    int N, nSimV, nF, nE, nSimE;
    const float4 *pos, *prev;
    const int32_t* info;          // mesh | fixed << 31
    const int32_t* s2o;           // sorted -> global original index
    const int32_t* o2s;
    const MeshDev* mt;
    const int4* faceSid;          // three sorted ids, mesh
    const int4* faceEdges;        // three global edge ids
    const int4* edgeSid;          // eV1, eV2, eV12Next, eV21Next (sorted ids, -1 = none)
    const int32_t* edgeMesh;
    const int32_t *nbrOff, *nbr;  // CSR over sorted ids: neighbour vertices (sorted ids)
    const int32_t *vfOff, *vfInc; // CSR over sorted ids: incident faces, entry = globalFace << 2 | order
    const int32_t *veOff, *veInc; // CSR over sorted ids: incident edges, entry = globalEdge << 1 | order
};

// checkEdgeFeasibleRegion (TriMeshCollisionGeometry.cpp:92-151)
CDEV bool edge_feasible(V3 p, int4 E, const float4* __restrict__ pos, float eps) {
    const V3 v1 = ld3(pos, E.x), v2 = ld3(pos, E.y);
    const V3 AP = p - v1, BP = p - v2, AB = v2 - v1;
    const float relaxed = eps + 1e-6f;
    if (gb::dot(AP, AB) < -relaxed) return false;
    if (gb::dot(BP, AB) > relaxed) return false;
    const float ABSqrNorm = gb::sqnorm(AB);
    if (ABSqrNorm < 1e-10f) return false;
#pragma unroll
    for (int side = 0; side < 2; side++) {
        const int vn = side == 0 ? E.z : E.w;
        if (vn == -1) continue;
        const V3 vNext = ld3(pos, vn);
        const float t = gb::dot(AB, vNext - v1) / ABSqrNorm;
        const V3 foot = v1 + t * AB;
        if (gb::dot(vNext - foot, p - foot) > relaxed) return false;
    }
    return true;
}
// checkVertexFeasibleRegion (:153-184)
CDEV bool vertex_feasible(V3 p, int sid, const DevView& D, float eps) {
    const V3 A = ld3(D.pos, sid);
    const V3 AP = p - A;
    const float relaxed = -eps - 1e-6f;
    for (int i = D.nbrOff[sid]; i < D.nbrOff[sid + 1]; i++) {
        const V3 BA = A - ld3(D.pos, D.nbr[i]);
        if (gb::dot(AP, BA) < relaxed) return false;
    }
    return true;
}

CDEV float box_dist2(float4 lo, float4 hi, V3 p) {
    const float dx = p.x < lo.x ? lo.x - p.x : (p.x > hi.x ? p.x - hi.x : 0.f);
    const float dy = p.y < lo.y ? lo.y - p.y : (p.y > hi.y ? p.y - hi.y : 0.f);
    const float dz = p.z < lo.z ? lo.z - p.z : (p.z > hi.z ? p.z - hi.z : 0.f);
    return dx * dx + dy * dy + dz * dz;
}

// ------------------------------------------------------------------------------------------------ VF query
// contactQueryVF + triMeshVFRadiusQueryWithTopologyFilteringAndFaceMinDisCaculatingFunc[Collider]: every vertex of every
// mesh against the faces within maxQueryDis. Key of a contact = closest primitive (type << 28 | id, mesh ignored as in the
// reference); per key the feasible face with the smallest global id is kept (rule R5), the list is sorted by face id.
static __global__ void __launch_bounds__(kB)
k_vf_query(DevView D, StepP P, BvhView bv, int32_t* __restrict__ vfCount, int2* __restrict__ vfList, float* __restrict__ vfMinDis,
           int32_t* faceMinBits, int32_t* errFlags, const int32_t* __restrict__ filtOff, const int32_t* __restrict__ filt) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= D.N) return;
    const int qm = D.info[s] & 0x00FFFFFF;
    const V3 p = ld3(D.pos, s);
    const float R = P.R, R2 = R * R * 1.0001f;
    float minDis = R;
    int gid[KV], key[KV], cnt = 0;
    int stack[kStack], sp = 0;
    int node = bv.n == 1 ? bv.n : 0;
    bool overflow = false, stackOverflow = false;
    if (bv.n > 0) for (;;) {
        if (node >= bv.n) {
            const int f = bv.order[node - bv.n];
            const int4 fv = __ldg(D.faceSid + f);
            const int fm = fv.w;
            bool skip = (fm >= P.numSim && qm >= P.numSim) || (qm == fm && (fv.x == s || fv.y == s || fv.z == s));
            // the seams' filtered pairs (filtOff: NULL unless a mesh has them; the lists hold global face ids of the vertex's own mesh)
            if (filtOff && !skip) for (int q = filtOff[s]; q < filtOff[s + 1]; q++) if (filt[q] == f) { skip = true; break; }
            if (!skip) {
                V3 bary; int type;
                const V3 cp = gb::closest_point_triangle(p, ld3(D.pos, fv.x), ld3(D.pos, fv.y), ld3(D.pos, fv.z), bary, type);
                const float d = gb::norm(p - cp);
                if (d < R) {
                    minDis = fminf(d, minDis);
                    atomicMin(faceMinBits + f, __float_as_int(d));
                    const MeshDev& M = D.mt[fm];
                    int primId, primType;
                    bool ok;
                    if (type <= gb::AtC) {
                        const int sid = type == gb::AtA ? fv.x : (type == gb::AtB ? fv.y : fv.z);
                        primType = 0; primId = D.s2o[sid] - M.vBase;
                        ok = vertex_feasible(p, sid, D, 1e-3f);
                    } else if (type <= gb::AtAC) {
                        const int4 fe = __ldg(D.faceEdges + f);
                        const int eg = type == gb::AtAB ? fe.x : (type == gb::AtBC ? fe.y : fe.z);
                        primType = 1; primId = eg - M.eBase;
                        ok = edge_feasible(p, __ldg(D.edgeSid + eg), D.pos, 1e-3f);
                    } else { primType = 2; primId = f - M.fBase; ok = true; }
                    if (ok) {
                        const int k = primType << 28 | primId;
                        int i = 0;
                        for (; i < cnt; i++) if (key[i] == k) break;
                        if (i < cnt) { if (f < gid[i]) gid[i] = f; }
                        else if (cnt < KV) { gid[cnt] = f; key[cnt] = k; cnt++; }
                        else overflow = true;
                    }
                }
            }
            if (!sp) break;
            node = stack[--sp];
        } else {
            const int2 c = bv.child[node];
            const bool hl = box_dist2(bv.lo[c.x], bv.hi[c.x], p) <= R2, hr = box_dist2(bv.lo[c.y], bv.hi[c.y], p) <= R2;
            if (hl && hr) { if (sp < kStack) stack[sp++] = c.y; else stackOverflow = true; node = c.x; }
            else if (hl) node = c.x;
            else if (hr) node = c.y;
            else { if (!sp) break; node = stack[--sp]; }
        }
    }
    for (int i = 1; i < cnt; i++) {   // ascending face id
        const int g = gid[i], k = key[i];
        int j = i - 1;
        while (j >= 0 && gid[j] > g) { gid[j + 1] = gid[j]; key[j + 1] = key[j]; j--; }
        gid[j + 1] = g; key[j + 1] = k;
    }
    vfCount[s] = cnt;
    vfMinDis[s] = minDis;
    for (int i = 0; i < cnt; i++) vfList[(size_t)s * KV + i] = make_int2(gid[i], key[i]);
    if (overflow) atomicAdd(errFlags + ERR_VF, 1);
    if (stackOverflow) atomicAdd(errFlags + ERR_STACK, 1);
}

// ------------------------------------------------------------------------------------------------ EE query
// contactQueryEE + triMeshEdgeContactQueryFunc: every edge of a simulated mesh against the edges within maxQueryDis
static __global__ void __launch_bounds__(kB)
k_ee_query(DevView D, StepP P, BvhView bv, int32_t* __restrict__ eeCount, int32_t* __restrict__ eeList, float* __restrict__ eeMinDis,
           int32_t* errFlags, const int32_t* __restrict__ filtOff, const int32_t* __restrict__ filt) {
    const int qe = blockIdx.x * blockDim.x + threadIdx.x;
    if (qe >= D.nSimE) return;
    const int4 Q = __ldg(D.edgeSid + qe);
    const V3 p1 = ld3(D.pos, Q.x), p2 = ld3(D.pos, Q.y);
    const float R = P.R, pad = R * 1.001f;
    const V3 qlo = v3(fminf(p1.x, p2.x) - pad, fminf(p1.y, p2.y) - pad, fminf(p1.z, p2.z) - pad);
    const V3 qhi = v3(fmaxf(p1.x, p2.x) + pad, fmaxf(p1.y, p2.y) + pad, fmaxf(p1.z, p2.z) + pad);
    float minDis = R;
    int lst[KE], cnt = 0;
    int stack[kStack], sp = 0;
    int node = bv.n == 1 ? bv.n : 0;
    bool overflow = false, stackOverflow = false;
    auto hit = [&](int nd) {
        const float4 lo = bv.lo[nd], hi = bv.hi[nd];
        return qlo.x <= hi.x && qhi.x >= lo.x && qlo.y <= hi.y && qhi.y >= lo.y && qlo.z <= hi.z && qhi.z >= lo.z;
    };
    if (bv.n > 0) for (;;) {
        if (node >= bv.n) {
            const int te = bv.order[node - bv.n];
            const int4 T = __ldg(D.edgeSid + te);
            // sorted vertex ids are unique over the scene: sharing an id implies the same mesh
            bool skip = te == qe || Q.x == T.x || Q.x == T.y || Q.y == T.x || Q.y == T.y;
            if (filtOff && !skip) for (int q = filtOff[qe]; q < filtOff[qe + 1]; q++) if (filt[q] == te) { skip = true; break; }
            if (!skip) {
                const V3 q1 = ld3(D.pos, T.x), q2 = ld3(D.pos, T.y);
                V3 c1, c2; float mua, mub;
                closest_points_segments(p1, p2, q1, q2, c1, c2, mua, mub);
                const float d = gb::norm(c1 - c2);
                minDis = fminf(d, minDis);
                if (d < R) {
                    bool feasible = false;
                    const V3 eVec1 = p2 - p1, eVec2 = q2 - q1;
                    if (gb::dot(c2 - p1, eVec1) >= 0 && gb::dot(p2 - c2, eVec1) >= 0) feasible = true;
                    else if (gb::dot(c1 - q1, eVec2) >= 0 && gb::dot(q2 - c1, eVec2) >= 0) feasible = true;
                    if (feasible) { if (cnt < KE) lst[cnt++] = te; else overflow = true; }
                }
            }
            if (!sp) break;
            node = stack[--sp];
        } else {
            const int2 c = bv.child[node];
            const bool hl = hit(c.x), hr = hit(c.y);
            if (hl && hr) { if (sp < kStack) stack[sp++] = c.y; else stackOverflow = true; node = c.x; }
            else if (hl) node = c.x;
            else if (hr) node = c.y;
            else { if (!sp) break; node = stack[--sp]; }
        }
    }
    for (int i = 1; i < cnt; i++) {
        const int g = lst[i];
        int j = i - 1;
        while (j >= 0 && lst[j] > g) { lst[j + 1] = lst[j]; j--; }
        lst[j + 1] = g;
    }
    eeCount[qe] = cnt;
    eeMinDis[qe] = minDis;
    for (int i = 0; i < cnt; i++) eeList[(size_t)qe * KE + i] = lst[i];
    if (overflow) atomicAdd(errFlags + ERR_EE, 1);
    if (stackOverflow) atomicAdd(errFlags + ERR_STACK, 1);
}

// ------------------------------------------------------------------------------------------------ face back-links
// FVContactInfo lists (ClothContactDetector.cpp:263-276): per face, the VF contacts that landed on it, in ascending
// (mesh, vertex, contact id) of the query vertex. Value = globalOriginalVertex * KV + contactId.
static __global__ void k_fill_i32(int n, int32_t* p, int32_t v) { const int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) p[i] = v; }
static __global__ void k_link_count(int N, const int32_t* __restrict__ vfCount, const int2* __restrict__ vfList, int32_t* linkCnt) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= N) return;
    const int c = vfCount[s];
    for (int i = 0; i < c; i++) atomicAdd(linkCnt + vfList[(size_t)s * KV + i].x, 1);
}
static __global__ void k_link_fill(int N, int64_t cap, const int32_t* __restrict__ s2o, const int32_t* __restrict__ vfCount, const int2* __restrict__ vfList,
                                   const int32_t* __restrict__ linkOff, int32_t* linkCur, int32_t* linkVal, int32_t* errFlags) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= N) return;
    const int c = vfCount[s];
    for (int i = 0; i < c; i++) {
        const int f = vfList[(size_t)s * KV + i].x;
        const int64_t at = (int64_t)linkOff[f] + atomicAdd(linkCur + f, 1);
        if (at < cap) linkVal[at] = s2o[s] * KV + i; else atomicAdd(errFlags + ERR_LINK, 1);
    }
}
static __global__ void k_link_sort(int nF, int64_t cap, const int32_t* __restrict__ linkOff, int32_t* linkVal) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nF) return;
    const int b = linkOff[f], e = linkOff[f + 1];
    if (e - b < 2 || e > cap) return;
    for (int i = b + 1; i < e; i++) {
        const int v = linkVal[i];
        int j = i - 1;
        while (j >= b && linkVal[j] > v) { linkVal[j + 1] = linkVal[j]; j--; }
        linkVal[j + 1] = v;
    }
}

// conservative bounds + active mask (applyCollisionDetection, VBDClothPhysics.cpp:775-816)
// It also flattens, per vertex with contacts, the work list of the contact pre-pass (same order as the reference walks
// it: own VF contacts, back-links of the incident faces, EE contacts of the incident edges) into one item array:
// item = ((kind | order << 1) << 29 | a, b) with VF: a = query vertex, b = face; EE: a = edge 1, b = edge 2.
static __global__ void k_bounds_active(DevView D, StepP P, const float* __restrict__ vfMinDis, const int32_t* __restrict__ vfCount,
                                       const int2* __restrict__ vfList, const int32_t* __restrict__ faceMinBits, const int32_t* __restrict__ linkOff,
                                       const int32_t* __restrict__ linkVal, const float* __restrict__ eeMinDis, const int32_t* __restrict__ eeCount,
                                       const int32_t* __restrict__ eeList, float* __restrict__ bounds, uint8_t* __restrict__ hasContact,
                                       uint8_t* __restrict__ activeMask, float4* __restrict__ posAtCD, int32_t* itemCursor, int64_t itemCap,
                                       int32_t* __restrict__ itemOff, int32_t* __restrict__ itemCnt, int2* __restrict__ items, int32_t* errFlags) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= D.nSimV) return;
    float dMin = P.R;
    dMin = fminf(dMin, vfMinDis[s]);
    const int nOwn = vfCount[s];
    int total = nOwn;
    for (int i = D.vfOff[s]; i < D.vfOff[s + 1]; i++) {
        const int f = D.vfInc[i] >> 2;
        dMin = fminf(dMin, __int_as_float(faceMinBits[f]));
        total += linkOff[f + 1] - linkOff[f];
    }
    for (int i = D.veOff[s]; i < D.veOff[s + 1]; i++) {
        const int e = D.veInc[i] >> 1;
        dMin = fminf(dMin, eeMinDis[e]);
        total += eeCount[e];
    }
    dMin = fmaxf(0.f, dMin - P.thickness);
    bounds[s] = dMin * P.relax;
    const bool active = total > 0;
    hasContact[s] = active ? 1 : 0;
    if (active) activeMask[s] = 1;
    posAtCD[s] = D.pos[s];
    int cnt = 0, off = 0;
    if (active && D.info[s] >= 0) {
        off = atomicAdd(itemCursor, total);
        if ((int64_t)off + total <= itemCap) {
            int at = off;
            for (int i = 0; i < nOwn; i++) items[at++] = make_int2((0 | 3 << 1) << 29 | s, vfList[(size_t)s * KV + i].x);
            for (int q = D.vfOff[s]; q < D.vfOff[s + 1]; q++) {
                const int ent = D.vfInc[q], f = ent >> 2, order = ent & 3;
                for (int l = linkOff[f]; l < linkOff[f + 1]; l++) items[at++] = make_int2((0 | order << 1) << 29 | D.o2s[linkVal[l] / KV], f);
            }
            for (int q = D.veOff[s]; q < D.veOff[s + 1]; q++) {
                const int ent = D.veInc[q], e = ent >> 1, order = ent & 1;
                const int ne = eeCount[e];
                for (int j = 0; j < ne; j++) items[at++] = make_int2((1 | order << 1) << 29 | e, eeList[(size_t)e * KE + j]);
            }
            cnt = total;
        } else atomicAdd(errFlags + ERR_LINK, 1);
    }
    itemOff[s] = off; itemCnt[s] = cnt;
}
static __global__ void k_no_collision_bounds(int n, float* bounds, uint8_t* hasContact) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n) { bounds[s] = FLT_MAX; hasContact[s] = 0; }
}

// ------------------------------------------------------------------------------------------------ contact terms
// accumulateVFContactForceAndHessian (VBDClothPhysics.cpp:1248-1321), contact refreshed from the current positions
// (updateVFContactPointInfo, ClothContactDetector.cpp:1197-1228); vertexOrder 0..2 = face corner, 3 = the vertex side
CDEV void vf_contact(const DevView& D, const StepP& P, int qs, int f, int vertexOrder, V3& force, float* H) {
    const int4 fv = __ldg(D.faceSid + f);
    const V3 x = ld3(D.pos, qs);
    const V3 a = ld3(D.pos, fv.x), b_ = ld3(D.pos, fv.y), c_ = ld3(D.pos, fv.z);
    V3 bary; int type;
    const V3 cp = gb::closest_point_triangle(x, a, b_, c_, bary, type);
    const V3 diff = x - cp;
    const V3 n = normalized(diff);
    const float dis = gb::dot(diff, n);
    if (dis < P.radius + P.thickness) {
        const float b = vertexOrder == 0 ? -bary.x : (vertexOrder == 1 ? -bary.y : (vertexOrder == 2 ? -bary.z : 1.f));
        float dEdD, d2E;
        repulsion(P, dis, dEdD, d2E);
        const float lambda = -dEdD;
        force += (b * lambda) * n;
        add_outer(H, d2E * b * b, n);
        const V3 dx3 = x - ld3(D.prev, qs);
        const V3 cpPrev = bary.x * ld3(D.prev, fv.x) + bary.y * ld3(D.prev, fv.y) + bary.z * ld3(D.prev, fv.z);
        const V3 dx = dx3 - (cp - cpPrev);
        const V3 e0 = normalized(b_ - a);
        const V3 e1 = normalized(gb::cross(e0, n));
        const MeshDev& mf = D.mt[fv.w];
        const MeshDev& mv = D.mt[D.info[qs] & 0x00FFFFFF];
        const float mu = (mf.frictionDynamic + mv.frictionDynamic) * 0.5f;
        const float epsV = (mf.frictionEpsV + mv.frictionEpsV) * 0.5f;
        V3 ff; float fh[9];
        friction(mu, lambda, e0, e1, gb::dot(e0, dx), gb::dot(e1, dx), epsV * P.dt, ff, fh);
        force += b * ff;
#pragma unroll
        for (int k = 0; k < 9; k++) H[k] += b * b * fh[k];
    }
}
// accumulateEEContactForceAndHessian (:1323-1399) with updateEEContactPointInfo (ClothContactDetector.cpp:1230-1251)
CDEV void ee_contact(const DevView& D, const StepP& P, int e1, int e2, int vertexOrder, V3& force, float* H) {
    const int4 E1 = __ldg(D.edgeSid + e1), E2 = __ldg(D.edgeSid + e2);
    const V3 a1 = ld3(D.pos, E1.x), b1 = ld3(D.pos, E1.y), a2 = ld3(D.pos, E2.x), b2 = ld3(D.pos, E2.y);
    V3 c1, c2; float mu1, mu2;
    closest_points_segments(a1, b1, a2, b2, c1, c2, mu1, mu2);
    const V3 diff = c1 - c2;
    const float dis = gb::norm(diff);
    const V3 v1 = b1 - a1, v2 = b2 - a2;
    const float parallelEps = gb::sqnorm(gb::cross(v1, v2)) / (gb::norm(v1) * gb::norm(v2));
    if (dis < P.radius + P.thickness && parallelEps > 1e-5f) {
        const V3 n = diff / dis;
        const float b = vertexOrder == 0 ? 1.f - mu1 : (vertexOrder == 1 ? mu1 : (vertexOrder == 2 ? -1.f + mu2 : -mu2));
        float dEdD, d2E;
        repulsion(P, dis, dEdD, d2E);
        const float lambda = -dEdD;
        force += (b * lambda) * n;
        add_outer(H, d2E * b * b, n);
        const V3 c1Prev = (1.f - mu1) * ld3(D.prev, E1.x) + mu1 * ld3(D.prev, E1.y);
        const V3 c2Prev = (1.f - mu2) * ld3(D.prev, E2.x) + mu2 * ld3(D.prev, E2.y);
        const V3 dx = (c1 - c1Prev) - (c2 - c2Prev);
        const V3 t0 = normalized(v1);
        const V3 t1 = normalized(gb::cross(t0, n));
        const MeshDev& m1 = D.mt[D.edgeMesh[e1]];
        const MeshDev& m2 = D.mt[D.edgeMesh[e2]];
        const float mu = (m1.frictionDynamic + m2.frictionDynamic) * 0.5f;
        const float epsV = (m1.frictionEpsV + m2.frictionEpsV) * 0.5f;
        V3 ff; float fh[9];
        friction(mu, lambda, t0, t1, gb::dot(t0, dx), gb::dot(t1, dx), epsV * P.dt, ff, fh);
        force += b * ff;
#pragma unroll
        for (int k = 0; k < 9; k++) H[k] += b * b * fh[k];
    }
}

// per colour, before its sweep: the contact part of evaluateVertexForceAndHessian for every vertex with contacts
static __global__ void __launch_bounds__(kB)
k_contact_terms(int gs, int n, DevView D, StepP P, const uint8_t* __restrict__ hasContact, const int32_t* __restrict__ vfCount,
                const int2* __restrict__ vfList, const int32_t* __restrict__ linkOff, const int32_t* __restrict__ linkVal,
                const int32_t* __restrict__ eeCount, const int32_t* __restrict__ eeList, float4* __restrict__ cTerm) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int s = gs + t;
    if (!hasContact[s] || D.info[s] < 0) return;
    V3 force = v3(0, 0, 0);
    float H[9];
#pragma unroll
    for (int k = 0; k < 9; k++) H[k] = 0.f;
    const int nOwn = vfCount[s];
    for (int i = 0; i < nOwn; i++) vf_contact(D, P, s, vfList[(size_t)s * KV + i].x, 3, force, H);
    for (int q = D.vfOff[s]; q < D.vfOff[s + 1]; q++) {
        const int ent = D.vfInc[q], f = ent >> 2, order = ent & 3;
        for (int l = linkOff[f]; l < linkOff[f + 1]; l++) {
            const int val = linkVal[l];
            vf_contact(D, P, D.o2s[val / KV], f, order, force, H);
        }
    }
    for (int q = D.veOff[s]; q < D.veOff[s + 1]; q++) {
        const int ent = D.veInc[q], e = ent >> 1, order = ent & 1;
        const int ne = eeCount[e];
        for (int j = 0; j < ne; j++) ee_contact(D, P, e, eeList[(size_t)e * KE + j], order, force, H);
    }
    cTerm[3 * (size_t)s + 0] = make_float4(force.x, force.y, force.z, H[0]);
    cTerm[3 * (size_t)s + 1] = make_float4(H[1], H[2], H[3], H[4]);
    cTerm[3 * (size_t)s + 2] = make_float4(H[5], H[6], H[7], H[8]);
}

// ------------------------------------------------------------------------------------------------ colour sweep
// VBDStepWithExistingCollisions (VBDClothPhysics.cpp:1026-1154): one thread per vertex of the colour.
static __global__ void __launch_bounds__(kB)
k_solve_colour(int gs, int n, int ellF, int ellH, StepP P, float4* pos, const float4* __restrict__ prev, const float4* __restrict__ inertia,
               const int32_t* __restrict__ info, const MeshDev* __restrict__ mt, const float4* __restrict__ recFA, const int4* __restrict__ recFB,
               const float4* __restrict__ recHA, const int4* __restrict__ recHB, const uint8_t* __restrict__ hasContact,
               const float4* __restrict__ cTerm, const float4* __restrict__ posAtCD, const float* __restrict__ bounds, float4* next,
               int32_t* errFlags) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int s = gs + t;
    const int inf = info[s];
    if (inf < 0) return;                       // fixed
    const MeshDev M = mt[inf & 0x00FFFFFF];
    const float4 p4 = pos[s];
    const V3 x = v3(p4.x, p4.y, p4.z);
    const V3 xPrev = ld3(prev, s);
    const float4 in4 = inertia[s];
    const float mass = in4.w;
    V3 force = v3(0, 0, 0);
    float H[9];
#pragma unroll
    for (int k = 0; k < 9; k++) H[k] = 0.f;
    if (hasContact[s]) {
        const float4 c0 = cTerm[3 * (size_t)s], c1 = cTerm[3 * (size_t)s + 1], c2 = cTerm[3 * (size_t)s + 2];
        force = v3(c0.x, c0.y, c0.z);
        H[0] = c0.w; H[1] = c1.x; H[2] = c1.y; H[3] = c1.z; H[4] = c1.w; H[5] = c2.x; H[6] = c2.y; H[7] = c2.z; H[8] = c2.w;
    }
    // accumlateInertiaForceAndHessian (VBDBaseTriMesh.cpp:211-218)
    force += mass * (v3(in4.x, in4.y, in4.z) - x) * P.dtSqrRec;
    {
        const float hi = mass * P.dtSqrRec;
        H[0] += hi; H[4] += hi; H[8] += hi;
    }
    // ---- StVK membrane (VBDTriMeshStVK.cpp:721-911)
    float old[9];
#pragma unroll
    for (int k = 0; k < 9; k++) old[k] = H[k];
    const float lmbd = M.lambda, miu = M.miu;
    for (int k = 0; k < ellF; k++) {
        const int4 fb = __ldg(recFB + (size_t)k * n + t);
        if (fb.x < 0) break;
        const float4 I = __ldg(recFA + (size_t)k * n + t);
        const int order = fb.x == s ? 0 : (fb.y == s ? 1 : 2);
        const V3 x0 = order == 0 ? x : ld3(pos, fb.x), x1 = order == 1 ? x : ld3(pos, fb.y), x2 = order == 2 ? x : ld3(pos, fb.z);
        const V3 v1 = x1 - x0, v2 = x2 - x0;
        const float A = __int_as_float(fb.w);
        const float Dm_inv1_1 = I.x, Dm_inv2_1 = I.y, Dm_inv1_2 = I.z, Dm_inv2_2 = I.w;
        const float F1_1 = v1.x * I.x + v2.x * I.y, F2_1 = v1.y * I.x + v2.y * I.y, F3_1 = v1.z * I.x + v2.z * I.y;
        const float F1_2 = v1.x * I.z + v2.x * I.w, F2_2 = v1.y * I.z + v2.y * I.w, F3_2 = v1.z * I.z + v2.z * I.w;
        const float G00 = 0.5f * ((F1_1 * F1_1 + F2_1 * F2_1 + F3_1 * F3_1) - 1.f), G11 = 0.5f * ((F1_2 * F1_2 + F2_2 * F2_2 + F3_2 * F3_2) - 1.f);
        const float G01 = 0.5f * (F1_1 * F1_2 + F2_1 * F2_2 + F3_1 * F3_2);
        const float tr = G00 + G11;
        const float S00 = 2.f * miu * G00 + lmbd * tr, S11 = 2.f * miu * G11 + lmbd * tr, S01 = 2.f * miu * G01;
        const float Fc[6] = {F1_1, F2_1, F3_1, F1_2, F2_2, F3_2};
        float F12[6];
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const float fs0 = Fc[r] * S00 + Fc[3 + r] * S01, fs1 = Fc[r] * S01 + Fc[3 + r] * S11;
            F12[r] = -A * (fs0 * Dm_inv1_1 + fs1 * Dm_inv1_2);
            F12[3 + r] = -A * (fs0 * Dm_inv2_1 + fs1 * Dm_inv2_2);
        }
        const float e_uu = G00, e_vv = G11, e_uv = G01, e_sum = e_uu + e_vv;
        float Dm[6][6];
        Dm[0][0] = F1_1 * (F1_1 * lmbd + 2 * F1_1 * miu) + 2 * miu * e_uu + lmbd * e_sum + F1_2 * F1_2 * miu;
        Dm[0][1] = Dm[1][0] = F1_1 * (F2_1 * lmbd + 2 * F2_1 * miu) + F1_2 * F2_2 * miu;
        Dm[0][2] = Dm[2][0] = F1_1 * (F3_1 * lmbd + 2 * F3_1 * miu) + F1_2 * F3_2 * miu;
        Dm[0][3] = Dm[3][0] = 2 * miu * e_uv + F1_1 * F1_2 * lmbd + F1_1 * F1_2 * miu;
        Dm[0][4] = Dm[4][0] = F1_1 * F2_2 * lmbd + F1_2 * F2_1 * miu;
        Dm[0][5] = Dm[5][0] = F1_1 * F3_2 * lmbd + F1_2 * F3_1 * miu;
        Dm[1][1] = F2_1 * (F2_1 * lmbd + 2 * F2_1 * miu) + 2 * miu * e_uu + lmbd * e_sum + F2_2 * F2_2 * miu;
        Dm[1][2] = Dm[2][1] = F2_1 * (F3_1 * lmbd + 2 * F3_1 * miu) + F2_2 * F3_2 * miu;
        Dm[1][3] = Dm[3][1] = F1_2 * F2_1 * lmbd + F1_1 * F2_2 * miu;
        Dm[1][4] = Dm[4][1] = 2 * miu * e_uv + F2_1 * F2_2 * lmbd + F2_1 * F2_2 * miu;
        Dm[1][5] = Dm[5][1] = F2_1 * F3_2 * lmbd + F2_2 * F3_1 * miu;
        Dm[2][2] = F3_1 * (F3_1 * lmbd + 2 * F3_1 * miu) + 2 * miu * e_uu + lmbd * e_sum + F3_2 * F3_2 * miu;
        Dm[2][3] = Dm[3][2] = F1_2 * F3_1 * lmbd + F1_1 * F3_2 * miu;
        Dm[2][4] = Dm[4][2] = F2_2 * F3_1 * lmbd + F2_1 * F3_2 * miu;
        Dm[2][5] = Dm[5][2] = 2 * miu * e_uv + F3_1 * F3_2 * lmbd + F3_1 * F3_2 * miu;
        Dm[3][3] = F1_2 * (F1_2 * lmbd + 2 * F1_2 * miu) + 2 * miu * e_vv + lmbd * e_sum + F1_1 * F1_1 * miu;
        Dm[3][4] = Dm[4][3] = F1_2 * (F2_2 * lmbd + 2 * F2_2 * miu) + F1_1 * F2_1 * miu;
        Dm[3][5] = Dm[5][3] = F1_2 * (F3_2 * lmbd + 2 * F3_2 * miu) + F1_1 * F3_1 * miu;
        Dm[4][4] = F2_2 * (F2_2 * lmbd + 2 * F2_2 * miu) + 2 * miu * e_vv + lmbd * e_sum + F2_1 * F2_1 * miu;
        Dm[4][5] = Dm[5][4] = F2_2 * (F3_2 * lmbd + 2 * F3_2 * miu) + F2_1 * F3_1 * miu;
        Dm[5][5] = F3_2 * (F3_2 * lmbd + 2 * F3_2 * miu) + 2 * miu * e_vv + lmbd * e_sum + F3_1 * F3_1 * miu;
#pragma unroll
        for (int a = 0; a < 6; a++)
#pragma unroll
            for (int b = 0; b < 6; b++) Dm[a][b] *= A;
        float m1, m2;
        if (order == 0) { m1 = -Dm_inv1_1 - Dm_inv2_1; m2 = -Dm_inv1_2 - Dm_inv2_2; force += v3(-F12[0] - F12[3], -F12[1] - F12[4], -F12[2] - F12[5]); }
        else if (order == 1) { m1 = Dm_inv1_1; m2 = Dm_inv1_2; force += v3(F12[0], F12[1], F12[2]); }
        else { m1 = Dm_inv2_1; m2 = Dm_inv2_2; force += v3(F12[3], F12[4], F12[5]); }
        // assembleMembraneHessian (VBDBaseTriMesh.h:149-160)
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int r = 0; r < 3; r++) {
                const float hl0 = Dm[r][c] * m1 + Dm[r + 3][c] * m2, hl1 = Dm[r][c + 3] * m1 + Dm[r + 3][c + 3] * m2;
                H[r + 3 * c] += hl0 * m1 + hl1 * m2;
            }
    }
    {
        float dampH[9];
        const float ds = M.dampStVK / P.dt;
#pragma unroll
        for (int k = 0; k < 9; k++) dampH[k] = (H[k] - old[k]) * ds;
        force -= mul3(dampH, x - xPrev);
#pragma unroll
        for (int k = 0; k < 9; k++) H[k] += dampH[k];
    }
    // ---- quadratic bending (VBDTriMeshStVK.cpp:913-982)
    {
        const float kBd = M.kBend, damping = M.dampBend;
        for (int k = 0; k < ellH; k++) {
            const int4 hb = __ldg(recHB + (size_t)k * n + t);
            if (hb.x < 0) break;
            const float4 q = __ldg(recHA + (size_t)k * n + t);
            const int order = hb.x == s ? 0 : (hb.y == s ? 1 : (hb.z == s ? 2 : 3));
            const V3 X0 = order == 0 ? x : ld3(pos, hb.x), X1 = order == 1 ? x : ld3(pos, hb.y);
            const V3 X2 = order == 2 ? x : ld3(pos, hb.z), X3 = order == 3 ? x : ld3(pos, hb.w);
            const V3 n1 = gb::cross(X1 - X0, X2 - X0), n2 = gb::cross(X2 - X3, X1 - X3);
            if (gb::sqnorm(n1) < P.degThres2 || gb::sqnorm(n2) < P.degThres2) continue;
            V3 dE = v3(0, 0, 0);
            dE += (kBd * q.x) * X0; dE += (kBd * q.y) * X1; dE += (kBd * q.z) * X2; dE += (kBd * q.w) * X3;
            const float qd = order == 0 ? q.x : (order == 1 ? q.y : (order == 2 ? q.z : q.w));
            const float h = kBd * qd;
            force -= dE + (h * (x - xPrev)) * (damping / P.dt);
            const float hh = h * (1.f + damping / P.dt);
            H[0] += hh; H[4] += hh; H[8] += hh;
        }
    }
    // ---- world box (accumlateBoundaryForceAndHessian, VBDClothPhysics.cpp:1401-1463)
#pragma unroll
    for (int d = 0; d < 3; d++) {
        const float xd = gb::comp(x, d);
        float depth, sign;
        if (xd < P.bmin[d]) { depth = P.bmin[d] - xd; sign = 1.f; }
        else if (xd > P.bmax[d]) { depth = xd - P.bmax[d]; sign = -1.f; }
        else continue;
        const float fd = sign * (depth * P.kBoundary);
        if (d == 0) force.x += fd; else if (d == 1) force.y += fd; else force.z += fd;
        H[d + 3 * d] += P.kBoundary;
        const V3 dx = x - xPrev;
        const V3 t0 = v3((d + 1) % 3 == 0 ? 1.f : 0.f, (d + 1) % 3 == 1 ? 1.f : 0.f, (d + 1) % 3 == 2 ? 1.f : 0.f);
        const V3 t1 = v3((d + 2) % 3 == 0 ? 1.f : 0.f, (d + 2) % 3 == 1 ? 1.f : 0.f, (d + 2) % 3 == 2 ? 1.f : 0.f);
        V3 ff; float fh[9];
        friction(P.muBoundary, depth * P.kBoundary, t0, t1, gb::dot(t0, dx), gb::dot(t1, dx), P.epsVBoundary * P.dt, ff, fh);
        force += ff;
#pragma unroll
        for (int k = 0; k < 9; k++) H[k] += fh[k];
    }
    // ---- 3x3 solve, conservative-bound truncation
    V3 dxs;
    V3 out;
    if (solve3x3(H, force, dxs)) {
        V3 posNew = x + P.stepSize * dxs;
        const V3 c = ld3(posAtCD, s);
        V3 acc = posNew - c;
        const float size = gb::norm(acc), bound = bounds[s];
        if (size > bound) {
            acc = acc * (bound / size);
            posNew = c + acc;
            errFlags[FLAG_TRUNC] = 1;
        }
        out = posNew;
        next[s] = make_float4(out.x, out.y, out.z, 0.f);
    } else {
        atomicAdd(errFlags + ERR_SOLVER, 1);      // "Solver failed": the vertex takes vertexPositionNext as it was
        const float4 nx = next[s];
        out = v3(nx.x, nx.y, nx.z);
    }
    if (!(out.x == out.x) || !(out.y == out.y) || !(out.z == out.z)) atomicAdd(errFlags + ERR_NAN, 1);
    pos[s] = make_float4(out.x, out.y, out.z, p4.w);
}

// ------------------------------------------------------------------------------------------------ fast kernels
// The two kernels above walk the lists in the CPU path's order with one thread per vertex, which makes every sum
// bit-identical to the oracle's but leaves one long dependent chain per thread. The default path below spreads each
// vertex over a sub-warp (same terms, pairwise-tree summation instead of left-to-right): results agree to rounding
// (tests: 1e-5 relative after one iteration). GAIA_B200_CLOTH_EXACT_ORDER=1 selects the exact-order kernels.
static const int LC = 8;   // lanes per vertex in the contact pre-pass (the colour sweep takes 1, 4 or 8: see iteration())

static __global__ void __launch_bounds__(kB)
k_contact_terms_fast(int gs, int n, DevView D, StepP P, const int32_t* __restrict__ itemOff, const int32_t* __restrict__ itemCnt,
                     const int2* __restrict__ items, float4* __restrict__ cTerm) {
    const int T = blockIdx.x * blockDim.x + threadIdx.x;
    const int t = T / LC, j = T % LC;
    const int s = gs + min(t, n - 1);
    const int cnt = t < n ? itemCnt[s] : 0;
    if (!__any_sync(0xffffffffu, cnt > 0)) return;
    V3 force = v3(0, 0, 0);
    float H[9];
#pragma unroll
    for (int k = 0; k < 9; k++) H[k] = 0.f;
    const int off = itemOff[s];
    for (int i = j; i < cnt; i += LC) {
        const int2 it = items[off + i];
        const int head = (unsigned)it.x >> 29, a = it.x & 0x1FFFFFFF;
        if (head & 1) ee_contact(D, P, a, it.y, head >> 1, force, H);
        else vf_contact(D, P, a, it.y, head >> 1, force, H);
    }
#pragma unroll
    for (int o = LC / 2; o > 0; o >>= 1) {
        force.x += __shfl_xor_sync(0xffffffffu, force.x, o); force.y += __shfl_xor_sync(0xffffffffu, force.y, o);
        force.z += __shfl_xor_sync(0xffffffffu, force.z, o);
#pragma unroll
        for (int k = 0; k < 9; k++) H[k] += __shfl_xor_sync(0xffffffffu, H[k], o);
    }
    if (j == 0 && cnt > 0) {
        cTerm[3 * (size_t)s + 0] = make_float4(force.x, force.y, force.z, H[0]);
        cTerm[3 * (size_t)s + 1] = make_float4(H[1], H[2], H[3], H[4]);
        cTerm[3 * (size_t)s + 2] = make_float4(H[5], H[6], H[7], H[8]);
    }
}

// ------------------------------------------------------------------------------------------------ sewing / attachments
// VBDBaseTriMesh::accumlateSewingForceAndHessian (M/VBDCloth/VBDBaseTriMesh.cpp:320-344) and VBDClothSimulationFramework::
// accumulateConstraintForceAndHessian (VBDClothPhysics.cpp:1464-1498) are both "a stiffness times a fixed linear combination
// of up to four vertex positions (plus a constant)": entry e of vertex s holds r = sum_i w_i x(id_i) + c and contributes
// force -= km r, Hessian += kmm I (km = m k, kmm = m^2 k with m the coefficient of s's own role in the seam / attachment).
// They ride on the contact pre-pass slot (cTerm, hasContact) the colour sweep already reads, so the sweep kernels are
// untouched and scenes without such constraints launch nothing extra: per colour, k_constraint_clear zeroes the slots of
// the constrained vertices BEFORE the contact pre-pass (which only writes where it has contacts), k_constraint_terms adds
// the terms AFTER it and raises bit 1 of hasContact (detection rewrites that byte; this kernel runs before every sweep).
struct ConRec { int32_t id[4]; float w[4]; float c[3]; float km; float kmm; float pad[3]; };

static __global__ void k_constraint_clear(int gs, int n, const int32_t* __restrict__ conOff, float4* __restrict__ cTerm) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int s = gs + t;
    if (conOff[s + 1] == conOff[s]) return;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    cTerm[3 * (size_t)s + 0] = z; cTerm[3 * (size_t)s + 1] = z; cTerm[3 * (size_t)s + 2] = z;
}

static __global__ void k_constraint_terms(int gs, int n, const int32_t* __restrict__ conOff, const ConRec* __restrict__ con,
                                          const float4* __restrict__ pos, const int32_t* __restrict__ info, float4* __restrict__ cTerm,
                                          uint8_t* __restrict__ hasContact) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int s = gs + t;
    const int b = conOff[s], e = conOff[s + 1];
    if (e == b || info[s] < 0) return;
    V3 force = v3(0, 0, 0);
    float hd = 0.f;
    for (int q = b; q < e; q++) {
        const ConRec R = con[q];
        V3 r = v3(0, 0, 0);
#pragma unroll
        for (int i = 0; i < 4; i++) if (R.id[i] >= 0) r += R.w[i] * ld3(pos, R.id[i]);
        r += v3(R.c[0], R.c[1], R.c[2]);
        force -= R.km * r;
        hd += R.kmm;
    }
    float4 c0 = cTerm[3 * (size_t)s], c1 = cTerm[3 * (size_t)s + 1], c2 = cTerm[3 * (size_t)s + 2];
    c0.x += force.x; c0.y += force.y; c0.z += force.z;
    c0.w += hd; c1.w += hd; c2.w += hd;          // H[0], H[4], H[8]
    cTerm[3 * (size_t)s] = c0; cTerm[3 * (size_t)s + 1] = c1; cTerm[3 * (size_t)s + 2] = c2;
    hasContact[s] |= 2;
}

// Same update as k_solve_colour with LS lanes per vertex: lane j takes the incident faces / hinges j, j + LS, ... of the
// per-colour ELL streams; the StVK Hessian is contracted in closed form,
//   H = A [ a11 f1 f1^T + a22 f2 f2^T + a12 (f1 f2^T + f2 f1^T) + c I ],  f = -A [ f1 (S00 m1 + S01 m2) + f2 (S01 m1 + S11 m2) ],
//   a11 = m1^2 (l + 2u) + m2^2 u, a22 = m2^2 (l + 2u) + m1^2 u, a12 = m1 m2 (l + u),
//   c = m1^2 (2u Guu + l tr) + m2^2 (2u Gvv + l tr) + 4 m1 m2 u Guv
// (the 6x6 d2W/dF2 of VBDTriMeshStVK.cpp:760-870 contracted with (m1 I, m2 I), VBDBaseTriMesh.h:149-160).
template <int LS, int UF, int UH>
static __global__ void __launch_bounds__(kB, 6)
k_solve_colour_fast(int gs, int n, int ellF, int ellH, StepP P, float4* pos, const float4* __restrict__ prev, const float4* __restrict__ inertia,
                    const int32_t* __restrict__ info, const MeshDev* __restrict__ mt, const float4* __restrict__ recFA, const int4* __restrict__ recFB,
                    const float4* __restrict__ recHA, const int4* __restrict__ recHB, const uint8_t* __restrict__ hasContact,
                    const float4* __restrict__ cTerm, const float4* __restrict__ posAtCD, const float* __restrict__ bounds, float4* next,
                    int32_t* errFlags) {
    const int T = blockIdx.x * blockDim.x + threadIdx.x;
    const int tt = T / LS, j = T % LS;
    const bool valid = tt < n;
    const int t = min(tt, n - 1);
    const int s = gs + t;
    const int inf = info[s];
    const bool solve = valid && inf >= 0;
    const MeshDev& M = mt[inf & 0x00FFFFFF];
    const float4 p4 = pos[s];
    const V3 x = v3(p4.x, p4.y, p4.z);
    const float lmbd = M.lambda, miu = M.miu, kBd = M.kBend;
    V3 f = v3(0, 0, 0);
    float h00 = 0.f, h01 = 0.f, h02 = 0.f, h11 = 0.f, h12 = 0.f, h22 = 0.f, hb = 0.f;
    if (solve) {
        // UF (UH) rounds of records are loaded, then all their neighbour positions gathered, before the first use: the sweep is
        // bound by the latency of these dependent loads, so they are kept in flight together
        for (int k0 = j; k0 < ellF; k0 += UF * LS) {
            int4 fb[UF]; float4 I[UF]; V3 xa[UF], xb[UF];
#pragma unroll
            for (int u = 0; u < UF; u++) {
                const int k = k0 + u * LS;
                fb[u] = k < ellF ? __ldg(recFB + (size_t)k * n + t) : make_int4(-1, -1, -1, 0);
                I[u] = k < ellF ? __ldg(recFA + (size_t)k * n + t) : make_float4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < UF; u++) {
                const int ia = fb[u].x == s ? fb[u].y : fb[u].x, ib = fb[u].z == s ? fb[u].y : fb[u].z;
                xa[u] = fb[u].x >= 0 ? ld3(pos, ia) : x; xb[u] = fb[u].x >= 0 ? ld3(pos, ib) : x;
            }
#pragma unroll
            for (int u = 0; u < UF; u++) {
                if (fb[u].x < 0) continue;
                const int order = fb[u].x == s ? 0 : (fb[u].y == s ? 1 : 2);
                const V3 x0 = order == 0 ? x : xa[u], x1 = order == 0 ? xa[u] : (order == 1 ? x : xb[u]), x2 = order == 2 ? x : xb[u];
                const V3 v1 = x1 - x0, v2 = x2 - x0;
                const float A = __int_as_float(fb[u].w);
                const float4 Iu = I[u];
                const V3 f1 = Iu.x * v1 + Iu.y * v2, f2 = Iu.z * v1 + Iu.w * v2;
                const float G00 = 0.5f * (gb::dot(f1, f1) - 1.f), G11 = 0.5f * (gb::dot(f2, f2) - 1.f), G01 = 0.5f * gb::dot(f1, f2);
                const float tr = G00 + G11;
                const float S00 = 2.f * miu * G00 + lmbd * tr, S11 = 2.f * miu * G11 + lmbd * tr, S01 = 2.f * miu * G01;
                float m1, m2;
                if (order == 0) { m1 = -Iu.x - Iu.y; m2 = -Iu.z - Iu.w; }
                else if (order == 1) { m1 = Iu.x; m2 = Iu.z; }
                else { m1 = Iu.y; m2 = Iu.w; }
                f += (-A) * ((S00 * m1 + S01 * m2) * f1 + (S01 * m1 + S11 * m2) * f2);
                const float l2u = lmbd + 2.f * miu;
                const float a11 = A * (m1 * m1 * l2u + m2 * m2 * miu), a22 = A * (m2 * m2 * l2u + m1 * m1 * miu), a12 = A * (m1 * m2 * (lmbd + miu));
                const float c = A * (m1 * m1 * (2.f * miu * G00 + lmbd * tr) + m2 * m2 * (2.f * miu * G11 + lmbd * tr) + 4.f * m1 * m2 * miu * G01);
                h00 += a11 * f1.x * f1.x + a22 * f2.x * f2.x + 2.f * a12 * f1.x * f2.x + c;
                h11 += a11 * f1.y * f1.y + a22 * f2.y * f2.y + 2.f * a12 * f1.y * f2.y + c;
                h22 += a11 * f1.z * f1.z + a22 * f2.z * f2.z + 2.f * a12 * f1.z * f2.z + c;
                h01 += a11 * f1.x * f1.y + a22 * f2.x * f2.y + a12 * (f1.x * f2.y + f2.x * f1.y);
                h02 += a11 * f1.x * f1.z + a22 * f2.x * f2.z + a12 * (f1.x * f2.z + f2.x * f1.z);
                h12 += a11 * f1.y * f1.z + a22 * f2.y * f2.z + a12 * (f1.y * f2.z + f2.y * f1.z);
            }
        }
        for (int k0 = j; k0 < ellH; k0 += UH * LS) {
            int4 hbi[UH]; float4 q[UH]; V3 Xa[UH], Xb[UH], Xc[UH];
#pragma unroll
            for (int u = 0; u < UH; u++) {
                const int k = k0 + u * LS;
                hbi[u] = k < ellH ? __ldg(recHB + (size_t)k * n + t) : make_int4(-1, -1, -1, -1);
                q[u] = k < ellH ? __ldg(recHA + (size_t)k * n + t) : make_float4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < UH; u++) {
                // the three vertices of the hinge other than s, in slot order
                const int4 h4 = hbi[u];
                const int order = h4.x == s ? 0 : (h4.y == s ? 1 : (h4.z == s ? 2 : 3));
                const int ia = order == 0 ? h4.y : h4.x, ib = order <= 1 ? h4.z : h4.y, ic = order == 3 ? h4.z : h4.w;
                const bool on = h4.x >= 0;
                Xa[u] = on ? ld3(pos, ia) : x; Xb[u] = on ? ld3(pos, ib) : x; Xc[u] = on ? ld3(pos, ic) : x;
            }
#pragma unroll
            for (int u = 0; u < UH; u++) {
                const int4 h4 = hbi[u];
                if (h4.x < 0) continue;
                const int order = h4.x == s ? 0 : (h4.y == s ? 1 : (h4.z == s ? 2 : 3));
                const V3 X0 = order == 0 ? x : Xa[u], X1 = order == 0 ? Xa[u] : (order == 1 ? x : Xb[u]);
                const V3 X2 = order <= 1 ? Xb[u] : (order == 2 ? x : Xc[u]), X3 = order == 3 ? x : Xc[u];
                const V3 n1 = gb::cross(X1 - X0, X2 - X0), n2 = gb::cross(X2 - X3, X1 - X3);
                if (gb::sqnorm(n1) < P.degThres2 || gb::sqnorm(n2) < P.degThres2) continue;
                const float4 qu = q[u];
                f -= kBd * (qu.x * X0 + qu.y * X1 + qu.z * X2 + qu.w * X3);
                hb += kBd * (order == 0 ? qu.x : (order == 1 ? qu.y : (order == 2 ? qu.z : qu.w)));
            }
        }
    }
#pragma unroll
    for (int o = LS / 2; o > 0; o >>= 1) {
        f.x += __shfl_xor_sync(0xffffffffu, f.x, o); f.y += __shfl_xor_sync(0xffffffffu, f.y, o); f.z += __shfl_xor_sync(0xffffffffu, f.z, o);
        h00 += __shfl_xor_sync(0xffffffffu, h00, o); h01 += __shfl_xor_sync(0xffffffffu, h01, o); h02 += __shfl_xor_sync(0xffffffffu, h02, o);
        h11 += __shfl_xor_sync(0xffffffffu, h11, o); h12 += __shfl_xor_sync(0xffffffffu, h12, o); h22 += __shfl_xor_sync(0xffffffffu, h22, o);
        hb += __shfl_xor_sync(0xffffffffu, hb, o);
    }
    if (!solve || j != 0) return;
    // ---- tail (one lane per vertex): contacts, inertia, damping, world box, 3x3 solve, truncation
    const V3 xPrev = ld3(prev, s);
    const float4 in4 = inertia[s];
    const float mass = in4.w;
    V3 force = v3(0, 0, 0);
    float H[9];
#pragma unroll
    for (int k = 0; k < 9; k++) H[k] = 0.f;
    if (hasContact[s]) {
        const float4 c0 = cTerm[3 * (size_t)s], c1 = cTerm[3 * (size_t)s + 1], c2 = cTerm[3 * (size_t)s + 2];
        force = v3(c0.x, c0.y, c0.z);
        H[0] = c0.w; H[1] = c1.x; H[2] = c1.y; H[3] = c1.z; H[4] = c1.w; H[5] = c2.x; H[6] = c2.y; H[7] = c2.z; H[8] = c2.w;
    }
    force += mass * (v3(in4.x, in4.y, in4.z) - x) * P.dtSqrRec;
    {
        const float hi = mass * P.dtSqrRec;
        H[0] += hi; H[4] += hi; H[8] += hi;
    }
    const V3 disp = x - xPrev;
    {
        const float ds = M.dampStVK / P.dt, g1 = 1.f + ds;
        force += f - ds * v3(h00 * disp.x + h01 * disp.y + h02 * disp.z, h01 * disp.x + h11 * disp.y + h12 * disp.z, h02 * disp.x + h12 * disp.y + h22 * disp.z);
        H[0] += g1 * h00; H[4] += g1 * h11; H[8] += g1 * h22;
        H[1] += g1 * h01; H[3] += g1 * h01; H[2] += g1 * h02; H[6] += g1 * h02; H[5] += g1 * h12; H[7] += g1 * h12;
        const float db = M.dampBend / P.dt;
        force -= (hb * db) * disp;
        const float hh = hb * (1.f + db);
        H[0] += hh; H[4] += hh; H[8] += hh;
    }
#pragma unroll
    for (int d = 0; d < 3; d++) {
        const float xd = gb::comp(x, d);
        float depth, sign;
        if (xd < P.bmin[d]) { depth = P.bmin[d] - xd; sign = 1.f; }
        else if (xd > P.bmax[d]) { depth = xd - P.bmax[d]; sign = -1.f; }
        else continue;
        const float fd = sign * (depth * P.kBoundary);
        if (d == 0) force.x += fd; else if (d == 1) force.y += fd; else force.z += fd;
        H[d + 3 * d] += P.kBoundary;
        const V3 t0 = v3((d + 1) % 3 == 0 ? 1.f : 0.f, (d + 1) % 3 == 1 ? 1.f : 0.f, (d + 1) % 3 == 2 ? 1.f : 0.f);
        const V3 t1 = v3((d + 2) % 3 == 0 ? 1.f : 0.f, (d + 2) % 3 == 1 ? 1.f : 0.f, (d + 2) % 3 == 2 ? 1.f : 0.f);
        V3 ff; float fh[9];
        friction(P.muBoundary, depth * P.kBoundary, t0, t1, gb::dot(t0, disp), gb::dot(t1, disp), P.epsVBoundary * P.dt, ff, fh);
        force += ff;
#pragma unroll
        for (int k = 0; k < 9; k++) H[k] += fh[k];
    }
    V3 dxs, out;
    if (solve3x3(H, force, dxs)) {
        V3 posNew = x + P.stepSize * dxs;
        const V3 c = ld3(posAtCD, s);
        V3 acc = posNew - c;
        const float size = gb::norm(acc), bound = bounds[s];
        if (size > bound) {
            acc = acc * (bound / size);
            posNew = c + acc;
            errFlags[FLAG_TRUNC] = 1;
        }
        out = posNew;
        next[s] = make_float4(out.x, out.y, out.z, 0.f);
    } else {
        atomicAdd(errFlags + ERR_SOLVER, 1);
        const float4 nx = next[s];
        out = v3(nx.x, nx.y, nx.z);
    }
    if (!(out.x == out.x) || !(out.y == out.y) || !(out.z == out.z)) atomicAdd(errFlags + ERR_NAN, 1);
    pos[s] = make_float4(out.x, out.y, out.z, p4.w);
}

// ------------------------------------------------------------------------------------------------ elementwise passes
// VBDBaseTriMesh::applyInitialStep + forwardStep_intertia (VBDBaseTriMesh.cpp:60-182)
static __global__ void k_initial_step(int n, StepP P, int hasVelPrev, float4* pos, float4* prev, float4* vel, float4* velPrev, float4* inertia,
                                      const int32_t* __restrict__ info, const MeshDev* __restrict__ mt, const float* __restrict__ bounds) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const int inf = info[s];
    const MeshDev& M = mt[inf & 0x00FFFFFF];
    const float4 p4 = pos[s];
    const V3 x = v3(p4.x, p4.y, p4.z);
    prev[s] = p4;
    V3 v = ld3(vel, s);
    const V3 accel = (v - ld3(velPrev, s)) / P.dt;   // only meaningful when hasVelPrev
    const V3 vBefore = v;                           // the new velocitiesPrev
    velPrev[s] = make_float4(v.x, v.y, v.z, 0.f);
    const V3 g = v3(P.g[0], P.g[1], P.g[2]);
    if (P.assocGravity) v += P.dt * g;
    if (inf < 0) v = v3(0, 0, 0);
    vel[s] = make_float4(v.x, v.y, v.z, 0.f);
    const V3 y = x + P.dt * v;
    inertia[s] = make_float4(y.x, y.y, y.z, inertia[s].w);
    V3 xn = x;
    switch (M.initType) {
        case 0: xn = y; break;
        case 1:
            if (hasVelPrev) {
                const float gravNorm = gb::norm(g);
                const V3 gravDir = g / gravNorm;
                float ac = gb::dot(accel, gravDir);
                ac = ac < gravNorm ? ac : gravNorm;
                ac = ac > 1e-5f ? ac : 0.f;
                xn = x + P.dt * vBefore + ((P.dt * P.dt) * gravDir) * ac;
            } else xn = y;
            break;
        case 2: xn = x; break;
        case 3: {
            const float bound = bounds[s];
            V3 step = y - x;
            const float size = gb::norm(step);
            if (P.handleCollision && size > bound && size) step = step * (bound / size);
            xn = x + step;
        } break;
        default: break;
    }
    pos[s] = make_float4(xn.x, xn.y, xn.z, p4.w);
}

// updateVelocity (VBDClothPhysics.cpp:2318-2359)
// ------------------------------------------------------------------------------------------------ deformers
// ClothDeformerRotator::operator() (M/VBDCloth/VBDClothDeformer.h:71-117): p = root + R * RootP, R built on the host in
// double from theta = curTime * angularVelocity and rounded to float like Eigen's Matrix3f; row sums left to right.
struct Rot9 { float m[9]; };
static __global__ void k_deform_rotate(int n, const int32_t* __restrict__ sid, const float4* __restrict__ root, const float4* __restrict__ rootP,
                                       Rot9 R, float4* pos) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 r = root[i], q = rootP[i];
    const int s = sid[i];
    float4 p = pos[s];
    p.x = r.x + ((R.m[0] * q.x + R.m[1] * q.y) + R.m[2] * q.z);
    p.y = r.y + ((R.m[3] * q.x + R.m[4] * q.y) + R.m[5] * q.z);
    p.z = r.z + ((R.m[6] * q.x + R.m[7] * q.y) + R.m[8] * q.z);
    pos[s] = p;
}
// ClothDeformerTranslater::operator() (:176-190): p += speed * dt while the deformation is active
static __global__ void k_deform_translate(int n, const int32_t* __restrict__ sid, float3 speed, float dt, float4* pos) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int s = sid[i];
    float4 p = pos[s];
    p.x += speed.x * dt; p.y += speed.y * dt; p.z += speed.z * dt;
    pos[s] = p;
}

// ColliderTrimeshSequence::update / set_state (M/SpatialQuery/ColiiderTriMeshSequence.h:59-129): positions = prev * (1 - t) + next * t
static __global__ void k_collider_lerp(int n, const float4* __restrict__ a, const float4* __restrict__ b, float t, float4* pos) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 A = a[i], B = b[i];
    const float w = 1 - t;
    float4 p = pos[i];
    p.x = A.x * w + B.x * t; p.y = A.y * w + B.y * t; p.z = A.z * w + B.z * t;
    pos[i] = p;
}

static __global__ void k_update_velocity(int n, StepP P, const float4* __restrict__ pos, const float4* __restrict__ prev, float4* vel,
                                         const int32_t* __restrict__ info, const MeshDev* __restrict__ mt) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const int inf = info[s];
    const MeshDev& M = mt[inf & 0x00FFFFFF];
    V3 v = (ld3(pos, s) - ld3(prev, s)) / P.dt;
    const float vMag = gb::norm(v);
    if (M.hasNoGravZone && M.maxVel > 0 && v.y > M.noGravThr) {
        if (vMag > M.maxVel) v = v * (M.maxVel / vMag);
    } else if (vMag > 1e-6f) {
        float vNew = vMag * M.expDamp - M.constDamp;
        vNew = vNew > 1e-6f ? vNew : 0.f;
        v = v * (vNew / vMag);
    } else v = v * 0.f;
    if (inf < 0) v = v3(0, 0, 0);
    vel[s] = make_float4(v.x, v.y, v.z, 0.f);
}

// applyAcceleration (VBDClothPhysics.cpp:893-915)
static __global__ void k_accelerate(int n, float omega, float4* pos, const float4* __restrict__ prevIter, const uint8_t* __restrict__ activeMask,
                                    const float4* __restrict__ posAtCD, const float* __restrict__ bounds, int32_t* errFlags) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n || activeMask[s]) return;
    const float4 p4 = pos[s];
    const V3 pi = ld3(prevIter, s);
    V3 posNew = omega * (v3(p4.x, p4.y, p4.z) - pi) + pi;
    const V3 c = ld3(posAtCD, s);
    V3 acc = posNew - c;
    const float size = gb::norm(acc), bound = bounds[s];
    if (size > bound) { acc = acc * (bound / size); posNew = c + acc; errFlags[FLAG_TRUNC] = 1; }
    pos[s] = make_float4(posNew.x, posNew.y, posNew.z, p4.w);
}

static __global__ void k_gather_xyz(int count, int origBegin, const int32_t* __restrict__ o2s, const float4* __restrict__ src, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const float4 p = src[o2s[origBegin + i]];
    out[3 * (size_t)i] = p.x; out[3 * (size_t)i + 1] = p.y; out[3 * (size_t)i + 2] = p.z;
}
static __global__ void k_scatter_xyz(int count, int origBegin, const int32_t* __restrict__ o2s, const float* __restrict__ in, float4* dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const int s = o2s[origBegin + i];
    const float w = dst[s].w;
    dst[s] = make_float4(in[3 * (size_t)i], in[3 * (size_t)i + 1], in[3 * (size_t)i + 2], w);
}
static __global__ void k_gather_f(int count, int origBegin, const int32_t* __restrict__ o2s, const float* __restrict__ src, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) out[i] = src[o2s[origBegin + i]];
}

// ------------------------------------------------------------------------------------------------ scene
struct ClothScene {
    ClothHost h;
    int device = 0;
    std::string err;
    bool finalized = false;
    cudaStream_t stream = nullptr;
    int G = 0;
    // state (sorted order)
    DevBuf<float4> pos, prev, vel, velPrev, inertia, next, posAtCD, prevIter, cTerm;
    DevBuf<float> bounds;
    DevBuf<uint8_t> hasContact, activeMask;
    DevBuf<int32_t> info, s2o, o2s;
    DevBuf<MeshDev> meshTab;
    // per-colour ELL records
    DevBuf<float4> recFA, recHA;
    DevBuf<int4> recFB, recHB;
    // topology for detection / contact terms
    DevBuf<int4> faceSid, faceEdges, edgeSid;
    DevBuf<int32_t> edgeMesh, nbrOff, nbr, vfOff, vfInc, veOff, veInc;
    Bvh faceBvh, edgeBvh;
    DevBuf<int32_t> boundsOrd;
    // detection results
    DevBuf<int32_t> vfCount, eeCount, eeList, faceMinBits, linkCnt, linkOff, linkCur, linkVal;
    DevBuf<int2> vfList;
    DevBuf<float> vfMinDis, eeMinDis;
    DevBuf<uint8_t> scanTmp;
    size_t scanTmpBytes = 0;
    int64_t linkCap = 0;
    DevBuf<int32_t> itemOff, itemCnt, itemCursor;
    DevBuf<int2> items;
    int64_t itemCap = 0;
    DevBuf<int32_t> vfFiltOff, vfFilt, eeFiltOff, eeFilt;   // filtered contact pairs of seams (CSR over sorted vertices / simulated edges); empty: none
    DevBuf<int32_t> conOff;             // sewing / attachment entries per simulated sorted vertex (CSR); empty: none
    DevBuf<ConRec> con;
    bool hasConstraints = false;
    bool exactOrder = false;            // GAIA_B200_CLOTH_EXACT_ORDER: one thread per vertex, the CPU path's summation order
    int variant = 0;
    bool exactContact = false, exactSolve = false;   // (2 / 3 select one of the two kernels only: debugging aid)
    DevBuf<int32_t> errFlags;
    int32_t* h_flags = nullptr;         // pinned
    DevBuf<float> ioStage;
    // step state
    StepP P;
    bool collisionDetectionRequired = false;
    int64_t substepsTaken = 0;
    int frameId = 0, substep = 0;
    float curTime = 0.f;                // += dt every substep, like the framework's curTime (float accumulation)
    struct DeformerDev { DevBuf<int32_t> sid; DevBuf<float4> root, rootP; };
    std::vector<DeformerDev> deformerDev;   // one per h.deformers entry
    std::vector<DevBuf<float4>> colliderKeyDev;   // one per h.colliderSeqs entry: nKeys x nV positions
    bool detectedOnce = false;
    gaia_cloth_stats stats;
    bool profiling = false;
    std::vector<cudaEvent_t> ev;
    int64_t nSolved = 0, iterAlgBytes = 0;
    ~ClothScene() {
        for (auto e : ev) cudaEventDestroy(e);
        if (h_flags) cudaFreeHost(h_flags);
        if (stream) cudaStreamDestroy(stream);
    }
    DevView view() const {
        DevView D;
        D.N = (int)h.N; D.nSimV = (int)h.nSimV; D.nF = (int)h.nF; D.nE = (int)h.nE; D.nSimE = (int)h.nSimE;
        D.pos = pos.p; D.prev = prev.p; D.info = info.p; D.s2o = s2o.p; D.o2s = o2s.p; D.mt = meshTab.p;
        D.faceSid = faceSid.p; D.faceEdges = faceEdges.p; D.edgeSid = edgeSid.p; D.edgeMesh = edgeMesh.p;
        D.nbrOff = nbrOff.p; D.nbr = nbr.p; D.vfOff = vfOff.p; D.vfInc = vfInc.p; D.veOff = veOff.p; D.veInc = veInc.p;
        return D;
    }
};

template <class T>
static cudaError_t upload(DevBuf<T>& buf, const std::vector<T>& v) {
    CK(buf.alloc(v.size()));
    if (v.empty()) return cudaSuccess;
    return cudaMemcpy(buf.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
}

static cudaError_t upload_deformers(ClothScene& s) {
    s.deformerDev.clear();
    s.deformerDev.resize(s.h.deformers.size());
    for (size_t i = 0; i < s.h.deformers.size(); i++) {
        const ClothDeformer& d = s.h.deformers[i];
        const ClothMesh& m = s.h.meshes[d.mesh];
        std::vector<int32_t> sid(d.verts.size());
        std::vector<float4> root(d.verts.size()), rootP(d.verts.size());
        for (size_t k = 0; k < d.verts.size(); k++) {
            sid[k] = m.sortedId[d.verts[k]];
            if (d.type == 0) {
                root[k] = make_float4(d.roots[3 * k], d.roots[3 * k + 1], d.roots[3 * k + 2], 0.f);
                rootP[k] = make_float4(d.rootPs[3 * k], d.rootPs[3 * k + 1], d.rootPs[3 * k + 2], 0.f);
            }
        }
        CK(upload(s.deformerDev[i].sid, sid));
        if (d.type == 0) { CK(upload(s.deformerDev[i].root, root)); CK(upload(s.deformerDev[i].rootP, rootP)); }
    }
    s.colliderKeyDev.clear();
    s.colliderKeyDev.resize(s.h.colliderSeqs.size());
    for (size_t q = 0; q < s.h.colliderSeqs.size(); q++) {
        const ColliderSequence& c = s.h.colliderSeqs[q];
        const ClothMesh& m = s.h.meshes[c.mesh];
        std::vector<float4> keys(c.keyFrames.size() * (size_t)m.nV);
        for (size_t k = 0; k < c.keyFrames.size(); k++)
            for (int v = 0; v < m.nV; v++) {          // colliders keep their vertex order in the sorted arrays
                const float* p = &c.keyPos[(k * (size_t)m.nV + v) * 3];
                keys[k * (size_t)m.nV + (m.sortedId[v] - m.sortedId[0])] = make_float4(p[0], p[1], p[2], 0.f);
            }
        CK(upload(s.colliderKeyDev[q], keys));
    }
    return cudaSuccess;
}

static void make_step_params(ClothScene& s) {
    const gaia_cloth_physics_params& p = s.h.phys;
    StepP& P = s.P;
    P.dt = s.h.dt; P.dtSqrRec = s.h.dtSqrReciprocal;
    for (int d = 0; d < 3; d++) { P.g[d] = p.gravity[d]; P.bmin[d] = p.worldBoundsMin[d]; P.bmax[d] = p.worldBoundsMax[d]; }
    P.kBoundary = p.boundaryContactStiffness; P.muBoundary = p.boundaryFrictionDynamic; P.epsVBoundary = p.boundaryFrictionEpsValue;
    P.stepSize = p.stepSize; P.kContact = p.contactStiffness; P.radius = p.contactRadius; P.thickness = p.thickness;
    P.relax = p.conservativeStepRelaxation; P.R = p.maxQueryDis; P.degThres2 = p.degenerateTriangleThres * p.degenerateTriangleThres;
    P.energyType = p.contactEnergyTypes; P.handleCollision = p.handleCollision; P.assocGravity = p.associateGravityWithInertia;
    P.numSim = s.h.numSim;
}

static cudaError_t device_upload(ClothScene& s) {
    ClothHost& h = s.h;
    const int64_t N = h.N;
    s.G = (int)h.groups.size();
    CK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    make_step_params(s);
    std::vector<float4> pos(N), zero(N, make_float4(0, 0, 0, 0)), inertia(N);
    std::vector<int32_t> info(N), o2s(N);
    std::vector<MeshDev> mt(h.meshes.size());
    for (auto& m : h.meshes) {
        MeshDev& M = mt[m.id];
        M.vBase = (int32_t)m.vBase; M.fBase = (int32_t)m.fBase; M.eBase = (int32_t)m.eBase; M.simulated = m.simulated;
        M.miu = m.obj.miu; M.lambda = m.obj.lambda; M.dampStVK = m.obj.dampingStVK; M.kBend = m.obj.bendingStiffness; M.dampBend = m.obj.dampingBending;
        M.frictionDynamic = m.obj.frictionDynamic; M.frictionEpsV = m.obj.frictionEpsV; M.expDamp = m.obj.exponentialVelDamping;
        M.constDamp = m.obj.constantVelDamping; M.maxVel = m.obj.maxVelocityMagnitude; M.noGravThr = m.obj.noGravZoneThreshold;
        M.hasNoGravZone = m.obj.hasNoGravZone; M.initType = m.obj.initializationType;
        for (int v = 0; v < m.nV; v++) {
            const int32_t sd = m.sortedId[v];
            pos[sd] = make_float4(m.pos0[3 * (size_t)v], m.pos0[3 * (size_t)v + 1], m.pos0[3 * (size_t)v + 2], 0.f);
            inertia[sd] = make_float4(0, 0, 0, m.mass[v]);
            info[sd] = m.id | (m.fixedMask[v] ? (int32_t)0x80000000 : 0);
            o2s[m.vBase + v] = sd;
        }
    }
    CK(upload(s.pos, pos)); CK(upload(s.prev, zero)); CK(upload(s.vel, zero)); CK(upload(s.velPrev, zero)); CK(upload(s.inertia, inertia));
    CK(upload(s.next, pos)); CK(upload(s.posAtCD, zero)); CK(upload(s.prevIter, pos));
    CK(s.cTerm.alloc(3 * (size_t)h.nSimV + 3));
    CK(s.bounds.alloc(N)); CK(cudaMemset(s.bounds.p, 0, N * sizeof(float)));
    CK(s.hasContact.alloc(N)); CK(cudaMemset(s.hasContact.p, 0, N));
    CK(s.activeMask.alloc(N)); CK(cudaMemset(s.activeMask.p, 0, N));
    CK(upload(s.info, info)); CK(upload(s.s2o, h.sortedToOrig)); CK(upload(s.o2s, o2s)); CK(upload(s.meshTab, mt));
    // ---- ELL records per colour
    {
        std::vector<float4> fa(h.nRecF, make_float4(0, 0, 0, 0)), ha(h.nRecH, make_float4(0, 0, 0, 0));
        std::vector<int4> fb(h.nRecF, make_int4(-1, -1, -1, 0)), hb(h.nRecH, make_int4(-1, -1, -1, -1));
        for (int g = 0; g < s.G; g++) {
            const int64_t n = (int64_t)h.groups[g].size() / 2;
            for (int64_t i = 0; i < n; i++) {
                const ClothMesh& m = h.meshes[h.groups[g][2 * i]];
                const int32_t v = h.groups[g][2 * i + 1];
                int k = 0;
                for (int q = m.vNbrFOff[v]; q < m.vNbrFOff[v + 1]; q++, k++) {
                    const int32_t f = m.vNbrF[q];
                    const int64_t at = h.ellFBase[g] + (int64_t)k * n + i;
                    const float* I = &m.DmInv[4 * (size_t)f];
                    fa[at] = make_float4(I[0], I[1], I[2], I[3]);
                    int32_t ab;
                    std::memcpy(&ab, &m.restArea[f], 4);
                    fb[at] = make_int4(m.sortedId[m.faces[3 * (size_t)f]], m.sortedId[m.faces[3 * (size_t)f + 1]], m.sortedId[m.faces[3 * (size_t)f + 2]], ab);
                }
                k = 0;
                for (int q = m.vBendOff[v]; q < m.vBendOff[v + 1]; q++, k++) {
                    const int32_t e = m.vBend[q], order = m.vBendOrder[q];
                    const int64_t at = h.ellHBase[g] + (int64_t)k * n + i;
                    const float* Q = &m.Q[16 * (size_t)e + 4 * order];
                    ha[at] = make_float4(Q[0], Q[1], Q[2], Q[3]);
                    const int32_t* E = &m.edges[6 * (size_t)e];
                    hb[at] = make_int4(m.sortedId[E[0]], m.sortedId[E[1]], m.sortedId[E[2]], m.sortedId[E[3]]);
                }
            }
        }
        CK(upload(s.recFA, fa)); CK(upload(s.recFB, fb)); CK(upload(s.recHA, ha)); CK(upload(s.recHB, hb));
    }
    // ---- detection topology
    {
        std::vector<int4> faceSid(h.nF), faceEdges(h.nF), edgeSid(h.nE);
        std::vector<int32_t> edgeMesh(h.nE), nbrOff(N + 1, 0), nbr, vfOff(N + 1, 0), vfInc, veOff(N + 1, 0), veInc;
        for (auto& m : h.meshes) {
            for (int f = 0; f < m.nF; f++) {
                const int32_t* fv = &m.faces[3 * (size_t)f];
                faceSid[m.fBase + f] = make_int4(m.sortedId[fv[0]], m.sortedId[fv[1]], m.sortedId[fv[2]], m.id);
                const int32_t* fe = &m.face3Edges[3 * (size_t)f];
                faceEdges[m.fBase + f] = make_int4((int32_t)(m.eBase + fe[0]), (int32_t)(m.eBase + fe[1]), (int32_t)(m.eBase + fe[2]), 0);
            }
            for (int e = 0; e < m.nE; e++) {
                const int32_t* E = &m.edges[6 * (size_t)e];
                edgeSid[m.eBase + e] = make_int4(m.sortedId[E[0]], m.sortedId[E[1]], E[2] < 0 ? -1 : m.sortedId[E[2]], E[3] < 0 ? -1 : m.sortedId[E[3]]);
                edgeMesh[m.eBase + e] = m.id;
            }
        }
        // CSR over sorted ids: fill counts, then lists, walking sorted ids in order
        for (int64_t sd = 0; sd < N; sd++) {
            const int64_t o = h.sortedToOrig[sd];
            int mi = 0;
            while (mi + 1 < (int)h.meshes.size() && o >= h.meshes[mi + 1].vBase) mi++;
            const ClothMesh& m = h.meshes[mi];
            const int32_t v = (int32_t)(o - m.vBase);
            for (int q = m.vNbrVOff[v]; q < m.vNbrVOff[v + 1]; q++) nbr.push_back(m.sortedId[m.vNbrV[q]]);
            for (int q = m.vNbrFOff[v]; q < m.vNbrFOff[v + 1]; q++) vfInc.push_back((int32_t)((m.fBase + m.vNbrF[q]) << 2 | m.vNbrFOrder[q]));
            for (int q = m.vNbrEOff[v]; q < m.vNbrEOff[v + 1]; q++) veInc.push_back((int32_t)((m.eBase + m.vNbrE[q]) << 1 | m.vNbrEOrder[q]));
            nbrOff[sd + 1] = (int32_t)nbr.size(); vfOff[sd + 1] = (int32_t)vfInc.size(); veOff[sd + 1] = (int32_t)veInc.size();
        }
        CK(upload(s.faceSid, faceSid)); CK(upload(s.faceEdges, faceEdges)); CK(upload(s.edgeSid, edgeSid)); CK(upload(s.edgeMesh, edgeMesh));
        CK(upload(s.nbrOff, nbrOff)); CK(upload(s.nbr, nbr)); CK(upload(s.vfOff, vfOff)); CK(upload(s.vfInc, vfInc));
        CK(upload(s.veOff, veOff)); CK(upload(s.veInc, veInc));
    }
    CK(s.faceBvh.alloc((int)h.nF)); CK(s.edgeBvh.alloc((int)h.nE));
    CK(s.boundsOrd.alloc(8));
    CK(s.vfCount.alloc(N)); CK(cudaMemset(s.vfCount.p, 0, N * 4));
    CK(s.vfList.alloc((size_t)N * KV)); CK(s.vfMinDis.alloc(N));
    CK(s.eeCount.alloc(h.nSimE + 1)); CK(cudaMemset(s.eeCount.p, 0, (h.nSimE + 1) * 4));
    CK(s.eeList.alloc((size_t)h.nSimE * KE + 1)); CK(s.eeMinDis.alloc(h.nSimE + 1));
    CK(s.faceMinBits.alloc(h.nF + 1)); CK(s.linkCnt.alloc(h.nF + 1)); CK(s.linkOff.alloc(h.nF + 1)); CK(s.linkCur.alloc(h.nF + 1));
    CK(cudaMemset(s.linkOff.p, 0, (h.nF + 1) * 4)); CK(cudaMemset(s.linkCnt.p, 0, (h.nF + 1) * 4));
    s.linkCap = std::min<int64_t>((int64_t)N * KV, 8 * N + 4096);
    CK(s.linkVal.alloc(s.linkCap));
    cub::DeviceScan::ExclusiveSum(nullptr, s.scanTmpBytes, s.linkCnt.p, s.linkOff.p, (int)(h.nF + 1));
    CK(s.scanTmp.alloc(s.scanTmpBytes + 16));
    s.itemCap = std::min<int64_t>(32 * h.nSimV + 4096, (int64_t)1 << 30);
    CK(s.items.alloc(s.itemCap)); CK(s.itemOff.alloc(h.nSimV + 1)); CK(s.itemCnt.alloc(h.nSimV + 1)); CK(s.itemCursor.alloc(1));
    CK(cudaMemset(s.itemCnt.p, 0, (h.nSimV + 1) * 4)); CK(cudaMemset(s.itemOff.p, 0, (h.nSimV + 1) * 4));
    // ---- filtered contact pairs of seams (TriMeshFEM::generate_sewing_filters): per sorted vertex the global ids of the faces it
    //      never pairs with, per global (simulated) edge the global ids of the edges it never pairs with. Only for meshes that
    //      have seams, as the reference attaches its ...WithSewing callbacks (ClothContactDetector.cpp:969-974, 1011-1016).
    {
        std::vector<std::vector<int32_t>> vf(N), ee(h.nSimE);
        bool any = false;
        for (auto& m : h.meshes) {
            if (!m.simulated || m.sewRatios.empty()) continue;
            for (size_t i = 0; i + 1 < m.filtVF.size(); i += 2) { vf[m.sortedId[m.filtVF[i]]].push_back((int32_t)(m.fBase + m.filtVF[i + 1])); any = true; }
            for (size_t i = 0; i + 1 < m.filtEE.size(); i += 2) {
                ee[m.eBase + m.filtEE[i]].push_back((int32_t)(m.eBase + m.filtEE[i + 1]));
                ee[m.eBase + m.filtEE[i + 1]].push_back((int32_t)(m.eBase + m.filtEE[i]));
                any = true;
            }
        }
        if (any) {
            auto flatten = [](const std::vector<std::vector<int32_t>>& per, std::vector<int32_t>& off, std::vector<int32_t>& flat) {
                off.assign(per.size() + 1, 0);
                for (size_t i = 0; i < per.size(); i++) { off[i + 1] = off[i] + (int32_t)per[i].size(); flat.insert(flat.end(), per[i].begin(), per[i].end()); }
                if (flat.empty()) flat.push_back(-1);
            };
            std::vector<int32_t> o1, f1, o2, f2;
            flatten(vf, o1, f1); flatten(ee, o2, f2);
            CK(upload(s.vfFiltOff, o1)); CK(upload(s.vfFilt, f1)); CK(upload(s.eeFiltOff, o2)); CK(upload(s.eeFilt, f2));
        }
    }
    // ---- sewing / attachment entries per simulated sorted vertex (ConRec; the order per vertex is the reference's: seams as
    //      sewingVMap lists them, then attachments in the order the set-up loop appends them, VBDClothPhysics.cpp:161-219)
    {
        std::vector<std::vector<ConRec>> per(h.nSimV);
        auto blank = [] { ConRec r; std::memset(&r, 0, sizeof(r)); for (int i = 0; i < 4; i++) r.id[i] = -1; return r; };
        for (auto& m : h.meshes) {
            const float k = m.sewStiffness;
            for (size_t i = 0; i < m.sewRatios.size(); i++) {
                const float lbda = m.sewRatios[i];
                for (int o = 0; o < 3; o++) {
                    ConRec r = blank();
                    for (int q = 0; q < 3; q++) r.id[q] = m.sortedId[m.sewVerts[3 * i + q]];
                    r.w[0] = 1.f; r.w[1] = -(1 - lbda); r.w[2] = -lbda;
                    const float w = o == 0 ? 1.f : (o == 1 ? -(1 - lbda) : -lbda);
                    r.km = w * k; r.kmm = w * w * k;
                    per[m.sortedId[m.sewVerts[3 * i + o]]].push_back(r);
                }
            }
        }
        for (auto& a : h.attachments) {
            const ClothMesh& mv = h.meshes[a.meshV];
            auto entry = [&](int solvedMesh, int vId, float b) {
                ConRec r = blank();
                r.id[0] = mv.sortedId[vId]; r.w[0] = 1.f;          // x(meshV, vId): the V-side mesh read with the solved vertex's id (:1472)
                if (a.meshF >= 0) for (int q = 0; q < 3; q++) { r.id[1 + q] = h.meshes[a.meshF].sortedId[a.face[q]]; r.w[1 + q] = -a.pos[q]; }
                else for (int q = 0; q < 3; q++) r.c[q] = -a.pos[q];
                r.km = a.stiffness * b; r.kmm = a.stiffness * b * b;
                per[h.meshes[solvedMesh].sortedId[vId]].push_back(r);
            };
            entry(a.meshV, a.vertex, 1.f);
            if (a.meshF >= 0 && h.meshes[a.meshF].simulated)
                for (int q = 0; q < 3; q++) if (a.pos[q] != 0.f) entry(a.meshF, a.face[q], -a.pos[q]);
        }
        std::vector<int32_t> off(h.nSimV + 1, 0);
        std::vector<ConRec> flat;
        for (int64_t v = 0; v < h.nSimV; v++) { off[v + 1] = off[v] + (int32_t)per[v].size(); flat.insert(flat.end(), per[v].begin(), per[v].end()); }
        s.hasConstraints = !flat.empty();
        if (s.hasConstraints) { CK(upload(s.conOff, off)); CK(upload(s.con, flat)); }
    }
    { const char* ex = getenv("GAIA_B200_CLOTH_EXACT_ORDER"); s.exactOrder = ex && ex[0] == '1'; s.exactContact = s.exactOrder || (ex && ex[0] == '2'); s.exactSolve = s.exactOrder || (ex && ex[0] == '3'); }
    { const char* va = getenv("GAIA_B200_CLOTH_VARIANT"); s.variant = va ? atoi(va) : 0; }
    CK(s.errFlags.alloc(N_FLAGS)); CK(cudaMemset(s.errFlags.p, 0, N_FLAGS * 4));
    CK(cudaMallocHost((void**)&s.h_flags, N_FLAGS * sizeof(int32_t)));
    std::memset(s.h_flags, 0, N_FLAGS * sizeof(int32_t));
    int64_t maxV = 1;
    for (auto& m : h.meshes) maxV = std::max<int64_t>(maxV, m.nV);
    CK(s.ioStage.alloc(3 * (size_t)N));
    s.ev.resize(2 * (size_t)std::max(1, s.G));
    for (auto& e : s.ev) CK(cudaEventCreate(&e));
    s.nSolved = 0; s.iterAlgBytes = 0;
    for (int g = 0; g < s.G; g++) { s.nSolved += h.groupSolved[g]; s.iterAlgBytes += h.groupAlgBytes[g]; }
    return cudaDeviceSynchronize();
}

// ------------------------------------------------------------------------------------------------ detection
template <int NV>
static cudaError_t update_bvh(ClothScene& s, Bvh& b, const int4* prims, bool rebuild) {
    if (!b.n) return cudaSuccess;
    const bool doBuild = !b.built || rebuild;
    k_cloth_prim_boxes<NV><<<nblk(b.n), kB, 0, s.stream>>>(b.n, prims, s.pos.p, s.boundsOrd.p, doBuild ? 1 : 0, b.primLo.p, b.primHi.p, b.keyIn.p);
    s.stats.kernelLaunches++;
    if (doBuild) {
        CK(cub::DeviceRadixSort::SortPairs(b.tmp.p, b.tmpBytes, b.keyIn.p, b.keyOut.p, b.idIn.p, b.order.p, b.n, 0, 30, s.stream));
        k_cloth_karras<<<nblk(b.n), kB, 0, s.stream>>>(b.n, b.keyOut.p, b.child.p, b.parent.p);
        s.stats.kernelLaunches += 3;
        b.built = true;
        b.detectionsSinceBuild = 0;
    }
    CK(cudaMemsetAsync(b.flags.p, 0, (size_t)b.n * sizeof(int32_t), s.stream));
    k_cloth_refit<<<nblk(b.n), kB, 0, s.stream>>>(b.n, b.order.p, b.primLo.p, b.primHi.p, b.child.p, b.parent.p, b.lo.p, b.hi.p, b.flags.p);
    s.stats.kernelLaunches++;
    b.detectionsSinceBuild++;
    return cudaGetLastError();
}

// applyCollisionDetection (VBDClothPhysics.cpp:735-820)
static cudaError_t detect(ClothScene& s) {
    ClothHost& h = s.h;
    const int N = (int)h.N;
    const int rebuildEvery = std::max(1, h.phys.contactBVHReconstructionSteps);
    const bool rebuild = !s.faceBvh.built || s.faceBvh.detectionsSinceBuild >= rebuildEvery;
    if (rebuild) {
        k_cloth_init_bounds<<<1, 32, 0, s.stream>>>(s.boundsOrd.p);
        k_cloth_scene_bounds<<<nblk(N, 256), 256, 0, s.stream>>>(N, s.pos.p, s.boundsOrd.p);
        s.stats.kernelLaunches += 2;
    }
    CK(update_bvh<3>(s, s.faceBvh, s.faceSid.p, rebuild));
    CK(update_bvh<2>(s, s.edgeBvh, s.edgeSid.p, rebuild));
    const DevView D = s.view();
    int32_t rBits;
    std::memcpy(&rBits, &s.P.R, 4);
    k_fill_i32<<<nblk(h.nF + 1), kB, 0, s.stream>>>((int)h.nF + 1, s.faceMinBits.p, rBits);   // resetFaceContactInfo
    CK(cudaMemsetAsync(s.linkCnt.p, 0, (h.nF + 1) * sizeof(int32_t), s.stream));
    CK(cudaMemsetAsync(s.linkCur.p, 0, (h.nF + 1) * sizeof(int32_t), s.stream));
    k_vf_query<<<nblk(N), kB, 0, s.stream>>>(D, s.P, view_of(s.faceBvh), s.vfCount.p, s.vfList.p, s.vfMinDis.p, s.faceMinBits.p, s.errFlags.p,
                                             s.vfFiltOff.p, s.vfFilt.p);
    if (h.nSimE) k_ee_query<<<nblk(h.nSimE), kB, 0, s.stream>>>(D, s.P, view_of(s.edgeBvh), s.eeCount.p, s.eeList.p, s.eeMinDis.p, s.errFlags.p,
                                                                s.eeFiltOff.p, s.eeFilt.p);
    k_link_count<<<nblk(N), kB, 0, s.stream>>>(N, s.vfCount.p, s.vfList.p, s.linkCnt.p);
    CK(cub::DeviceScan::ExclusiveSum(s.scanTmp.p, s.scanTmpBytes, s.linkCnt.p, s.linkOff.p, (int)(h.nF + 1), s.stream));
    k_link_fill<<<nblk(N), kB, 0, s.stream>>>(N, s.linkCap, s.s2o.p, s.vfCount.p, s.vfList.p, s.linkOff.p, s.linkCur.p, s.linkVal.p, s.errFlags.p);
    k_link_sort<<<nblk(h.nF), kB, 0, s.stream>>>((int)h.nF, s.linkCap, s.linkOff.p, s.linkVal.p);
    CK(cudaMemsetAsync(s.itemCursor.p, 0, sizeof(int32_t), s.stream));
    if (h.nSimV)
        k_bounds_active<<<nblk(h.nSimV), kB, 0, s.stream>>>(D, s.P, s.vfMinDis.p, s.vfCount.p, s.vfList.p, s.faceMinBits.p, s.linkOff.p, s.linkVal.p,
                                                           s.eeMinDis.p, s.eeCount.p, s.eeList.p, s.bounds.p, s.hasContact.p, s.activeMask.p,
                                                           s.posAtCD.p, s.itemCursor.p, s.itemCap, s.itemOff.p, s.itemCnt.p, s.items.p, s.errFlags.p);
    s.stats.kernelLaunches += 8;
    s.collisionDetectionRequired = false;
    s.stats.numberCollisionDetection++;
    s.detectedOnce = true;
    return cudaGetLastError();
}

static int check_flags(ClothScene& s) {
    const int32_t* f = s.h_flags;
    if (f[ERR_VF]) { s.err = "more than 24 vertex-face contacts on one vertex"; return GAIA_ERR_OVERFLOW; }
    if (f[ERR_EE]) { s.err = "more than 64 edge-edge contacts on one edge"; return GAIA_ERR_OVERFLOW; }
    if (f[ERR_STACK]) { s.err = "BVH traversal stack overflow"; return GAIA_ERR_OVERFLOW; }
    if (f[ERR_LINK]) { s.err = "face back-link storage exhausted"; return GAIA_ERR_OVERFLOW; }
    if (f[ERR_NAN]) { s.err = "NaN position"; return GAIA_ERR_NAN; }
    return GAIA_OK;
}

// ChebysevAccelerator::getAcceleratorOmega
static float accelerator_omega(float rho, int order, float prevOmega) {
    switch (order) {
        case 1: return 1.f;
        case 2: return 2.f / (2.f - rho * rho);
        default: return 4.f / (4.f - rho * rho * prevOmega);
    }
}

// one iteration of runStep_CDEveryIter (:399-509); returns a GAIA status
static void apply_deformers(ClothScene& s) {
    for (size_t i = 0; i < s.h.deformers.size(); i++) {
        const ClothDeformer& d = s.h.deformers[i];
        const int n = (int)d.verts.size();
        if (!n) continue;
        auto& dev = s.deformerDev[i];
        if (d.type == 0) {
            float curTime = s.curTime;
            if (curTime >= d.rotationEnd && d.rotationEnd > 0) curTime = (float)d.rotationEnd;
            const double theta = curTime * d.angularVelocity;
            const double ux = d.axis[0], uy = d.axis[1], uz = d.axis[2], c = std::cos(theta), sn = std::sin(theta);
            Rot9 R;
            R.m[0] = (float)(c + ux * ux * (1 - c));      R.m[1] = (float)(ux * uy * (1 - c) - uz * sn); R.m[2] = (float)(ux * uz * (1 - c) + uy * sn);
            R.m[3] = (float)(uy * ux * (1 - c) + uz * sn); R.m[4] = (float)(c + uy * uy * (1 - c));      R.m[5] = (float)(uy * uz * (1 - c) - ux * sn);
            R.m[6] = (float)(uz * ux * (1 - c) - uy * sn); R.m[7] = (float)(uz * uy * (1 - c) + ux * sn); R.m[8] = (float)(c + uz * uz * (1 - c));
            k_deform_rotate<<<nblk(n), kB, 0, s.stream>>>(n, dev.sid.p, dev.root.p, dev.rootP.p, R, s.pos.p);
        } else {
            if (!(s.curTime >= d.deformationStartTime && (s.curTime < d.deformationEndTime || d.deformationEndTime < 0))) continue;
            k_deform_translate<<<nblk(n), kB, 0, s.stream>>>(n, dev.sid.p, make_float3(d.speed[0], d.speed[1], d.speed[2]), s.h.dt, s.pos.p);
        }
        s.stats.kernelLaunches++;
    }
}

// DynamicCollider::updateColliderMeshes (M/SpatialQuery/DynamicCollider.h:39-50) over the key-frame sequences; returns
// true when a collider moved (collision detection is then required, VBDClothPhysics.cpp:415-418)
static bool update_colliders(ClothScene& s, int substep, int iIter) {
    bool any = false;
    const int numSubsteps = s.h.phys.numSubsteps;
    for (size_t q = 0; q < s.h.colliderSeqs.size(); q++) {
        ColliderSequence& c = s.h.colliderSeqs[q];
        const ClothMesh& m = s.h.meshes[c.mesh];
        const int nV = m.nV, frameId = s.frameId;
        float4* dst = s.pos.p + m.sortedId[0];
        const float4* keys = s.colliderKeyDev[q].p;
        auto key = [&](size_t k) { return keys + k * (size_t)nV; };
        auto lerp = [&](size_t k0, size_t k1, float t) {
            k_collider_lerp<<<nblk(nV), kB, 0, s.stream>>>(nV, key(k0), key(k1), t, dst);
            s.stats.kernelLaunches++;
        };
        const std::vector<int32_t>& kf = c.keyFrames;
        bool updated = false;
        if (c.firstTime) {            // set_state(frameId)
            c.firstTime = false;
            if (frameId < kf.front()) lerp(0, 0, 0.f);
            else if (frameId >= kf.back()) lerp(kf.size() - 1, kf.size() - 1, 0.f);
            else {
                const size_t hi = std::upper_bound(kf.begin(), kf.end(), frameId) - kf.begin();
                lerp(hi - 1, hi, float(frameId - kf[hi - 1]) / (kf[hi] - kf[hi - 1]));
            }
            updated = true;
        } else if (!(frameId < kf.front() || frameId >= kf.back())) {
            const size_t hi = std::upper_bound(kf.begin(), kf.end(), frameId) - kf.begin();
            const int prevF = kf[hi - 1], numFrames = kf[hi] - prevF;
            if (c.interpolateSubstep) {
                if (iIter < c.interpolateIter) {
                    float t = float(frameId - prevF) / numFrames;
                    t += float(substep) / ((size_t)numSubsteps * numFrames);
                    t += float(iIter) / ((size_t)c.interpolateIter * numSubsteps * numFrames);
                    lerp(hi - 1, hi, t);
                    updated = true;
                }
            } else if (substep == 0 && iIter == 0) {
                lerp(hi - 1, hi, float(frameId - prevF) / numFrames);
                updated = true;
            }
        }
        any = any || updated;
    }
    return any;
}

static int iteration(ClothScene& s, int iIter, float& omega) {
    ClothHost& h = s.h;
    const int nSim = (int)h.nSimV;
    cudaError_t e;
    auto cuda = [&](cudaError_t x) { if (x != cudaSuccess) { s.err = std::string("CUDA: ") + cudaGetErrorString(x); return false; } return true; };
    if (iIter == 0) apply_deformers(s);       // applyDeformers(), VBDClothPhysics.cpp:408-412
    if (iIter == 0 && nSim) if (!cuda(cudaMemsetAsync(s.activeMask.p, 0, nSim, s.stream))) return GAIA_ERR_CUDA;
    if (!h.colliderSeqs.empty() && update_colliders(s, s.substep, iIter)) s.collisionDetectionRequired = true;
    cudaEvent_t evD0 = nullptr;
    if (s.collisionDetectionRequired || iIter == 0) {
        if (s.profiling) { evD0 = s.ev[0]; cudaEventRecord(s.ev[0], s.stream); }
        if (h.phys.handleCollision) { if (!cuda(detect(s))) return GAIA_ERR_CUDA; }
        else if (nSim) { k_no_collision_bounds<<<nblk(nSim), kB, 0, s.stream>>>(nSim, s.bounds.p, s.hasContact.p); s.stats.kernelLaunches++; }
        if (s.profiling) {
            cudaEventRecord(s.ev[1], s.stream);
            cudaEventSynchronize(s.ev[1]);
            float ms = 0;
            cudaEventElapsedTime(&ms, s.ev[0], s.ev[1]);
            s.stats.timeCsmpUpdatingCollisionInfoDCD += ms;
        }
    }
    (void)evD0;
    if (h.phys.applyAcceleration && nSim)
        if (!cuda(cudaMemcpyAsync(s.prevIter.p, s.pos.p, (size_t)nSim * sizeof(float4), cudaMemcpyDeviceToDevice, s.stream))) return GAIA_ERR_CUDA;
    if (iIter == 0 && nSim) {
        k_initial_step<<<nblk(nSim), kB, 0, s.stream>>>(nSim, s.P, s.substepsTaken > 0 ? 1 : 0, s.pos.p, s.prev.p, s.vel.p, s.velPrev.p, s.inertia.p,
                                                         s.info.p, s.meshTab.p, s.bounds.p);
        s.stats.kernelLaunches++;
    }
    const DevView D = s.view();
    for (int g = 0; g < s.G; g++) {
        const int gs = h.groupStart[g], n = h.groupStart[g + 1] - gs;
        if (!n) continue;
        if (s.hasConstraints) { k_constraint_clear<<<nblk(n), kB, 0, s.stream>>>(gs, n, s.conOff.p, s.cTerm.p); s.stats.kernelLaunches++; }
        if (h.phys.handleCollision) {
            if (s.exactContact)
                k_contact_terms<<<nblk(n), kB, 0, s.stream>>>(gs, n, D, s.P, s.hasContact.p, s.vfCount.p, s.vfList.p, s.linkOff.p, s.linkVal.p,
                                                              s.eeCount.p, s.eeList.p, s.cTerm.p);
            else
                k_contact_terms_fast<<<nblk((int64_t)n * LC), kB, 0, s.stream>>>(gs, n, D, s.P, s.itemOff.p, s.itemCnt.p, s.items.p, s.cTerm.p);
            s.stats.kernelLaunches++;
        }
        if (s.hasConstraints) {
            k_constraint_terms<<<nblk(n), kB, 0, s.stream>>>(gs, n, s.conOff.p, s.con.p, s.pos.p, s.info.p, s.cTerm.p, s.hasContact.p);
            s.stats.kernelLaunches++;
        }
        if (s.profiling) cudaEventRecord(s.ev[2 * g], s.stream);
        if (s.exactSolve)
            k_solve_colour<<<nblk(n), kB, 0, s.stream>>>(gs, n, h.ellF[g], h.ellH[g], s.P, s.pos.p, s.prev.p, s.inertia.p, s.info.p, s.meshTab.p,
                                                         s.recFA.p + h.ellFBase[g], s.recFB.p + h.ellFBase[g], s.recHA.p + h.ellHBase[g],
                                                         s.recHB.p + h.ellHBase[g], s.hasContact.p, s.cTerm.p, s.posAtCD.p, s.bounds.p, s.next.p,
                                                         s.errFlags.p);
        else {
            auto launch = [&](auto kern, int lanes) {
                kern<<<nblk((int64_t)n * lanes), kB, 0, s.stream>>>(gs, n, h.ellF[g], h.ellH[g], s.P, s.pos.p, s.prev.p, s.inertia.p, s.info.p, s.meshTab.p,
                                                                   s.recFA.p + h.ellFBase[g], s.recFB.p + h.ellFBase[g], s.recHA.p + h.ellHBase[g],
                                                                   s.recHB.p + h.ellHBase[g], s.hasContact.p, s.cTerm.p, s.posAtCD.p, s.bounds.p,
                                                                   s.next.p, s.errFlags.p);
            };
            // lanes per vertex / rounds of face and hinge records in flight, by colour size (measured on B200, scripts/
            // cloth_variants.py): big colours fill the GPU with one thread per vertex and want loads in flight; small ones
            // need the extra lanes to have enough warps. GAIA_B200_CLOTH_VARIANT = 1 / 2 / 3 forces one (tests).
            const int v = s.variant ? s.variant : (n >= 80000 ? 3 : (n >= 8000 ? 1 : 2));
            switch (v) {
                case 1: launch(k_solve_colour_fast<4, 1, 1>, 4); break;
                case 2: launch(k_solve_colour_fast<8, 1, 2>, 8); break;
                default: launch(k_solve_colour_fast<1, 2, 2>, 1); break;
            }
        }
        if (s.profiling) cudaEventRecord(s.ev[2 * g + 1], s.stream);
        s.stats.kernelLaunches++;
        s.stats.colourSweeps++;
    }
    s.stats.vertexSolves += s.nSolved;
    s.stats.algorithmicBytes += s.iterAlgBytes;
    if (h.phys.applyAcceleration) {
        omega = accelerator_omega(h.phys.acceleratorRho, iIter + 1, omega);
        if (omega > 1.f && nSim) {
            k_accelerate<<<nblk(nSim), kB, 0, s.stream>>>(nSim, omega, s.pos.p, s.prevIter.p, s.activeMask.p, s.posAtCD.p, s.bounds.p, s.errFlags.p);
            s.stats.kernelLaunches++;
        }
    }
    // truncation flag (rule R6) + error flags: one small readback per iteration decides whether the next one re-detects
    if (!cuda(cudaMemcpyAsync(s.h_flags, s.errFlags.p, N_FLAGS * sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream))) return GAIA_ERR_CUDA;
    if (!cuda(cudaMemsetAsync(s.errFlags.p + FLAG_TRUNC, 0, sizeof(int32_t), s.stream))) return GAIA_ERR_CUDA;
    e = cudaStreamSynchronize(s.stream);
    if (!cuda(e)) return GAIA_ERR_CUDA;
    if (s.profiling)
        for (int g = 0; g < s.G; g++) {
            if (h.groupStart[g + 1] == h.groupStart[g]) continue;
            float ms = 0;
            cudaEventElapsedTime(&ms, s.ev[2 * g], s.ev[2 * g + 1]);
            s.stats.solveKernelMs += ms;
        }
    if (s.h_flags[FLAG_TRUNC]) s.collisionDetectionRequired = true;
    return check_flags(s);
}

static int run_substeps(ClothScene& s, int nSubsteps) {
    const int nSim = (int)s.h.nSimV;
    for (int sub = 0; sub < nSubsteps; sub++) {
        s.substep = sub;
        s.curTime += s.h.dt;
        float omega = 1.f;
        for (int it = 0; it < s.h.phys.iterations; it++) { const int rc = iteration(s, it, omega); if (rc < 0) return rc; }
        if (nSim) {
            k_update_velocity<<<nblk(nSim), kB, 0, s.stream>>>(nSim, s.P, s.pos.p, s.prev.p, s.vel.p, s.info.p, s.meshTab.p);
            s.stats.kernelLaunches++;
        }
        s.substepsTaken++;
        s.stats.substeps++;
    }
    s.stats.frames++;
    s.frameId++;
    return GAIA_OK;
}

}  // namespace gbc

// ================================================================================================== C ABI
using namespace gbc;
struct gaia_cloth_handle { ClothScene s; };
static thread_local std::string g_clothCreateErr;

static int cfail(ClothScene& s, int code, const std::string& msg) { s.err = msg; return code; }
static int ccuda(ClothScene& s, cudaError_t e) { s.err = std::string("CUDA: ") + cudaGetErrorString(e); return GAIA_ERR_CUDA; }
#define CREQ(h) if (!(h)) return GAIA_ERR_INVALID_ARG
#define CFIN(h) if (!(h)->s.finalized) return cfail((h)->s, GAIA_ERR_STATE, "scene not finalized")
#define CCUDA(h, x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return ccuda((h)->s, _e); } while (0)

extern "C" {

void gaia_b200_cloth_default_physics_params(gaia_cloth_physics_params* p) {
    std::memset(p, 0, sizeof(*p));
    p->gravity[1] = -10.f; p->timeStep = 0.01666666666666666666f; p->numSubsteps = 20; p->iterations = 1;
    for (int d = 0; d < 3; d++) { p->worldBoundsMin[d] = -1e7f; p->worldBoundsMax[d] = 1e7f; }
    p->boundaryFrictionDynamic = 0.1f; p->stepSize = 1.0f; p->associateGravityWithInertia = 1; p->applyAcceleration = 0;
    p->acceleratorRho = 0.95f; p->handleCollision = 1; p->contactStiffness = 1e4f; p->contactRadius = 1.0f; p->thickness = 0.f;
    p->conservativeStepRelaxation = 0.45f; p->boundaryContactStiffness = 1e4f; p->boundaryFrictionEpsValue = 1e-2f;
    p->contactEnergyTypes = 1; p->contactBVHReconstructionSteps = 32; p->degenerateTriangleThres = 1e-6f; p->maxQueryDis = 0.4f;
}
void gaia_b200_cloth_default_object_params(gaia_cloth_object_params* o) {
    std::memset(o, 0, sizeof(*o));
    for (int d = 0; d < 3; d++) o->scale[d] = 1.f;
    o->density = 1.f; o->maxVelocityMagnitude = -1.f; o->frictionDynamic = 0.1f; o->frictionEpsV = 0.1f;
    o->exponentialVelDamping = 1.0f; o->miu = 1e4f; o->lambda = 1e4f;
}
const char* gaia_b200_cloth_last_error(const gaia_cloth_handle* h) { return h ? h->s.err.c_str() : g_clothCreateErr.c_str(); }

int gaia_b200_cloth_create(const gaia_cloth_physics_params* phys, int device, gaia_cloth_handle** out) {
    if (!phys || !out) { g_clothCreateErr = "null argument"; return GAIA_ERR_INVALID_ARG; }
    *out = nullptr;
    int nDev = 0;
    cudaError_t e = cudaGetDeviceCount(&nDev);
    if (e != cudaSuccess || nDev <= 0) {
        g_clothCreateErr = std::string("no CUDA device available (") + cudaGetErrorString(e) + "); gaia_b200 has no CPU path";
        return GAIA_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= nDev) { g_clothCreateErr = "device ordinal out of range"; return GAIA_ERR_INVALID_ARG; }
    if (phys->numSubsteps <= 0 || phys->iterations < 0 || !(phys->timeStep > 0)) { g_clothCreateErr = "bad step parameters"; return GAIA_ERR_INVALID_ARG; }
    if (phys->contactEnergyTypes < 0 || phys->contactEnergyTypes > 2) { g_clothCreateErr = "bad contactEnergyTypes"; return GAIA_ERR_INVALID_ARG; }
    if ((e = cudaSetDevice(device)) != cudaSuccess) { g_clothCreateErr = cudaGetErrorString(e); return GAIA_ERR_CUDA; }
    gaia_cloth_handle* h = new gaia_cloth_handle();
    ClothScene& s = h->s;
    s.h.phys = *phys; s.device = device;
    s.h.dt = phys->timeStep / phys->numSubsteps;             // BasePhysicFramework::timeStepEqualize
    s.h.dtSqrReciprocal = 1.f / (s.h.dt * s.h.dt);
    std::memset(&s.stats, 0, sizeof(s.stats));
    *out = h;
    return GAIA_OK;
}
void gaia_b200_cloth_destroy(gaia_cloth_handle* h) {
    if (!h) return;
    cudaSetDevice(h->s.device);
    if (h->s.stream) cudaStreamSynchronize(h->s.stream);
    delete h;
}

int gaia_b200_cloth_add_tri_mesh(gaia_cloth_handle* h, int32_t simulated, int32_t nV, int32_t nF, const float* xyz, const int32_t* faces,
                                 const gaia_cloth_object_params* obj, int32_t nColors, const int32_t* color_offsets, const int32_t* color_verts,
                                 int32_t nFixed, const int32_t* fixedPoints) {
    CREQ(h);
    ClothScene& s = h->s;
    if (s.finalized) return cfail(s, GAIA_ERR_STATE, "scene already finalized");
    if (!xyz || !faces || !obj || nV <= 0 || nF <= 0) return cfail(s, GAIA_ERR_INVALID_ARG, "bad mesh arguments");
    if (simulated && !s.h.meshes.empty() && !s.h.meshes.back().simulated) return cfail(s, GAIA_ERR_STATE, "colliders must come after all simulated meshes");
    if (simulated && (nColors <= 0 || !color_offsets || !color_verts)) return cfail(s, GAIA_ERR_INVALID_ARG, "a simulated mesh needs a vertex colouring");
    for (int64_t i = 0; i < 3 * (int64_t)nF; i++) if (faces[i] < 0 || faces[i] >= nV) return cfail(s, GAIA_ERR_INVALID_ARG, "face index out of range");
    for (int i = 0; i < nFixed; i++) if (fixedPoints[i] < 0 || fixedPoints[i] >= nV) return cfail(s, GAIA_ERR_INVALID_ARG, "fixed point out of range");
    s.h.meshes.emplace_back();
    ClothMesh& m = s.h.meshes.back();
    m.id = (int)s.h.meshes.size() - 1;
    cloth_mesh_init(m, simulated != 0, nV, nF, xyz, faces, *obj, nColors, color_offsets, color_verts, nFixed, fixedPoints);
    return m.id;
}

int gaia_b200_cloth_finalize(gaia_cloth_handle* h) {
    CREQ(h);
    ClothScene& s = h->s;
    if (s.finalized) return cfail(s, GAIA_ERR_STATE, "scene already finalized");
    if (s.h.meshes.empty()) return cfail(s, GAIA_ERR_STATE, "no meshes");
    std::string err;
    if (!cloth_host_finalize(s.h, err)) return cfail(s, GAIA_ERR_INVALID_ARG, err);
    CCUDA(h, cudaSetDevice(s.device));
    CCUDA(h, device_upload(s));
    CCUDA(h, upload_deformers(s));
    s.finalized = true;
    return GAIA_OK;
}

// loadClothDeformers (M/VBDCloth/VBDClothDeformer.h:202-242): after the mesh is added, before finalize. The selected vertices
// join the mesh's fixed points (setFixedPointCloth, :5-21); a rotator keeps root / RootP of their positions at creation.
static int add_deformer_common(gaia_cloth_handle* h, int mesh, const int32_t* verts, int32_t n, ClothDeformer& d) {
    ClothScene& s = h->s;
    if (s.finalized) return cfail(s, GAIA_ERR_STATE, "deformers are added before finalize");
    if (mesh < 0 || mesh >= (int)s.h.meshes.size()) return cfail(s, GAIA_ERR_INVALID_ARG, "mesh id out of range");
    ClothMesh& m = s.h.meshes[mesh];
    if (!m.simulated) return cfail(s, GAIA_ERR_INVALID_ARG, "deformers select vertices of simulated meshes");
    if (n < 0 || (n > 0 && !verts)) return cfail(s, GAIA_ERR_INVALID_ARG, "bad vertex selection");
    for (int i = 0; i < n; i++) if (verts[i] < 0 || verts[i] >= m.nV) return cfail(s, GAIA_ERR_INVALID_ARG, "selected vertex out of range");
    d.mesh = mesh;
    d.verts.assign(verts, verts + n);
    for (int i = 0; i < n; i++) m.fixedMask[verts[i]] = 1;
    return GAIA_OK;
}
int gaia_b200_cloth_add_rotator(gaia_cloth_handle* h, int mesh, const int32_t* verts, int32_t n, const float center[3], const float axis[3],
                                double angularVelocity, double rotationEndTime) {
    CREQ(h);
    if (!center || !axis) return cfail(h->s, GAIA_ERR_INVALID_ARG, "null argument");
    ClothDeformer d;
    d.type = 0;
    const int rc = add_deformer_common(h, mesh, verts, n, d);
    if (rc < 0) return rc;
    for (int k = 0; k < 3; k++) { d.o[k] = center[k]; d.axis[k] = axis[k]; }
    d.angularVelocity = angularVelocity; d.rotationEnd = rotationEndTime;
    const ClothMesh& m = h->s.h.meshes[mesh];
    for (int i = 0; i < n; i++) {
        const float* P = &m.pos0[3 * (size_t)verts[i]];
        const float op[3] = {P[0] - d.o[0], P[1] - d.o[1], P[2] - d.o[2]};
        const float t = (op[0] * d.axis[0] + op[1] * d.axis[1]) + op[2] * d.axis[2];
        for (int k = 0; k < 3; k++) { const float r = t * d.axis[k]; d.roots.push_back(r); d.rootPs.push_back(P[k] - r); }
    }
    h->s.h.deformers.push_back(std::move(d));
    return (int)h->s.h.deformers.size() - 1;
}
// ColliderTrimeshSequenceParams::fromJson + ColliderTrimeshSequence::initialize (M/SpatialQuery/ColiiderTriMeshSequence.h:13-41,
// 131-144): key-frame meshes of a collider added with simulated = 0 (same topology), transformed like the mesh itself
int gaia_b200_cloth_set_collider_sequence(gaia_cloth_handle* h, int mesh, int32_t nKeys, const int32_t* keyFrames, const float* xyz,
                                          int32_t interpolateSubstep, int32_t interpolateIter) {
    CREQ(h);
    ClothScene& s = h->s;
    if (s.finalized) return cfail(s, GAIA_ERR_STATE, "collider sequences are set before finalize");
    if (mesh < 0 || mesh >= (int)s.h.meshes.size()) return cfail(s, GAIA_ERR_INVALID_ARG, "mesh id out of range");
    const ClothMesh& m = s.h.meshes[mesh];
    if (m.simulated) return cfail(s, GAIA_ERR_INVALID_ARG, "a key-frame sequence drives a collider (simulated = 0)");
    if (nKeys < 1 || !xyz) return cfail(s, GAIA_ERR_INVALID_ARG, "bad key frames");
    if (interpolateIter < 1) return cfail(s, GAIA_ERR_INVALID_ARG, "interpolateIter must be >= 1");
    for (auto& c : s.h.colliderSeqs) if (c.mesh == mesh) return cfail(s, GAIA_ERR_STATE, "the collider already has a sequence");
    ColliderSequence c;
    c.mesh = mesh;
    c.keyFrames.resize(nKeys);
    for (int i = 0; i < nKeys; i++) c.keyFrames[i] = keyFrames ? keyFrames[i] : i;      // default: one key per frame
    for (int i = 1; i < nKeys; i++) if (c.keyFrames[i] <= c.keyFrames[i - 1]) return cfail(s, GAIA_ERR_INVALID_ARG, "keyFrames should be in increasing order");
    c.interpolateIter = interpolateIter;
    c.interpolateSubstep = interpolateSubstep != 0 || interpolateIter > 1;
    if (c.interpolateIter > std::max(1, s.h.phys.iterations)) return cfail(s, GAIA_ERR_INVALID_ARG, "interpolateIter exceeds the iterations of a substep");
    for (int k = 0; k < nKeys; k++) {
        std::vector<float> p(xyz + (size_t)k * 3 * m.nV, xyz + (size_t)(k + 1) * 3 * m.nV);
        cloth_transform_points(m.obj, m.nV, p);
        c.keyPos.insert(c.keyPos.end(), p.begin(), p.end());
    }
    s.h.colliderSeqs.push_back(std::move(c));
    return GAIA_OK;
}
int gaia_b200_cloth_add_translator(gaia_cloth_handle* h, int mesh, const int32_t* verts, int32_t n, const float speed[3],
                                   float deformationStartTime, float deformationEndTime) {
    CREQ(h);
    if (!speed) return cfail(h->s, GAIA_ERR_INVALID_ARG, "null argument");
    ClothDeformer d;
    d.type = 1;
    const int rc = add_deformer_common(h, mesh, verts, n, d);
    if (rc < 0) return rc;
    for (int k = 0; k < 3; k++) d.speed[k] = speed[k];
    d.deformationStartTime = deformationStartTime; d.deformationEndTime = deformationEndTime;
    h->s.h.deformers.push_back(std::move(d));
    return (int)h->s.h.deformers.size() - 1;
}

// ObjectParams::sewingPath / ::sewingStiffness (M/TriMesh/TriMesh.cpp:95-110, 458); the force is k_constraint_terms
int gaia_b200_cloth_set_sewing(gaia_cloth_handle* h, int mesh, int32_t nSeams, const int32_t* verts3, const float* ratios, float sewingStiffness) {
    CREQ(h);
    ClothScene& s = h->s;
    if (s.finalized) return cfail(s, GAIA_ERR_STATE, "sewing is set before finalize");
    if (mesh < 0 || mesh >= (int)s.h.meshes.size()) return cfail(s, GAIA_ERR_INVALID_ARG, "mesh id out of range");
    ClothMesh& m = s.h.meshes[mesh];
    if (!m.simulated) return cfail(s, GAIA_ERR_INVALID_ARG, "sewing belongs to a simulated mesh");
    if (nSeams < 0 || (nSeams > 0 && (!verts3 || !ratios))) return cfail(s, GAIA_ERR_INVALID_ARG, "bad seam list");
    for (int i = 0; i < 3 * nSeams; i++) if (verts3[i] < 0 || verts3[i] >= m.nV) return cfail(s, GAIA_ERR_INVALID_ARG, "seam vertex out of range");
    m.sewVerts.assign(verts3, verts3 + 3 * (size_t)nSeams);
    m.sewRatios.assign(ratios, ratios + nSeams);
    m.sewStiffness = sewingStiffness;
    return GAIA_OK;
}
// ObjectParams::filteredPairsPath (TriMeshFEM::generate_sewing_filters, M/TriMesh/TriMesh.h:596-622); applied in k_vf_query / k_ee_query
int gaia_b200_cloth_set_filtered_pairs(gaia_cloth_handle* h, int mesh, int32_t nVF, const int32_t* vfPairs, int32_t nEE, const int32_t* eePairs) {
    CREQ(h);
    ClothScene& s = h->s;
    if (s.finalized) return cfail(s, GAIA_ERR_STATE, "filtered pairs are set before finalize");
    if (mesh < 0 || mesh >= (int)s.h.meshes.size()) return cfail(s, GAIA_ERR_INVALID_ARG, "mesh id out of range");
    ClothMesh& m = s.h.meshes[mesh];
    if (!m.simulated) return cfail(s, GAIA_ERR_INVALID_ARG, "filtered pairs belong to a simulated mesh");
    if (nVF < 0 || nEE < 0 || (nVF > 0 && !vfPairs) || (nEE > 0 && !eePairs)) return cfail(s, GAIA_ERR_INVALID_ARG, "bad pair list");
    for (int i = 0; i < nVF; i++)
        if (vfPairs[2 * i] < 0 || vfPairs[2 * i] >= m.nV || vfPairs[2 * i + 1] < 0 || vfPairs[2 * i + 1] >= m.nF) return cfail(s, GAIA_ERR_INVALID_ARG, "filtered (vertex, face) out of range");
    for (int i = 0; i < 2 * nEE; i++) if (eePairs[i] < 0 || eePairs[i] >= m.nE) return cfail(s, GAIA_ERR_INVALID_ARG, "filtered edge out of range");
    m.filtVF.assign(vfPairs, vfPairs + 2 * (size_t)nVF);
    m.filtEE.assign(eePairs, eePairs + 2 * (size_t)nEE);
    return GAIA_OK;
}
// one entry of PhysicsParams.VertexFaceAttachmentConstraintParams (set-up: VBDClothPhysics.cpp:152-220)
int gaia_b200_cloth_add_attachment(gaia_cloth_handle* h, int meshV, int32_t vertex, int meshF, const int32_t faceVerts[3], const float baryOrPos[3],
                                   float stiffness) {
    CREQ(h);
    ClothScene& s = h->s;
    if (s.finalized) return cfail(s, GAIA_ERR_STATE, "attachments are added before finalize");
    const int nM = (int)s.h.meshes.size();
    if (meshV < 0 || meshV >= nM || !s.h.meshes[meshV].simulated) return cfail(s, GAIA_ERR_INVALID_ARG, "the vertex side of an attachment is a simulated mesh");
    if (vertex < 0 || vertex >= s.h.meshes[meshV].nV) return cfail(s, GAIA_ERR_INVALID_ARG, "attached vertex out of range");
    if (meshF < -1 || meshF >= nM || !baryOrPos) return cfail(s, GAIA_ERR_INVALID_ARG, "bad attachment target");
    ClothAttachment a;
    a.meshV = meshV; a.vertex = vertex; a.meshF = meshF; a.stiffness = stiffness;
    for (int k = 0; k < 3; k++) a.pos[k] = baryOrPos[k];
    if (meshF >= 0) {
        if (!faceVerts) return cfail(s, GAIA_ERR_INVALID_ARG, "an attachment to a mesh names three face vertices");
        for (int k = 0; k < 3; k++) {
            if (faceVerts[k] < 0 || faceVerts[k] >= s.h.meshes[meshF].nV) return cfail(s, GAIA_ERR_INVALID_ARG, "attachment face vertex out of range");
            // the reaction on a simulated face vertex is computed from x(meshV, that vertex's id) (the reference, :1472)
            if (s.h.meshes[meshF].simulated && baryOrPos[k] != 0.f && faceVerts[k] >= s.h.meshes[meshV].nV)
                return cfail(s, GAIA_ERR_INVALID_ARG, "attachment face vertex id does not exist in the vertex-side mesh (the reference reads it there)");
            a.face[k] = faceVerts[k];
        }
    }
    s.h.attachments.push_back(a);
    return GAIA_OK;
}
int gaia_b200_cloth_run_frames(gaia_cloth_handle* h, int32_t nFrames) {
    CREQ(h); CFIN(h);
    CCUDA(h, cudaSetDevice(h->s.device));
    for (int i = 0; i < nFrames; i++) { const int rc = run_substeps(h->s, h->s.h.phys.numSubsteps); if (rc < 0) return rc; }
    return GAIA_OK;
}
int gaia_b200_cloth_run_iterations(gaia_cloth_handle* h, int32_t nIters) {
    CREQ(h); CFIN(h);
    CCUDA(h, cudaSetDevice(h->s.device));
    h->s.curTime += h->s.h.dt;
    h->s.substep = 0;
    float omega = 1.f;
    for (int it = 0; it < nIters; it++) { const int rc = iteration(h->s, it, omega); if (rc < 0) return rc; }
    return GAIA_OK;
}
int gaia_b200_cloth_frame_id(const gaia_cloth_handle* h) { return h ? h->s.frameId : GAIA_ERR_INVALID_ARG; }
double gaia_b200_cloth_cur_time(const gaia_cloth_handle* h) { return h ? (double)h->s.curTime : 0.0; }
int gaia_b200_cloth_set_time(gaia_cloth_handle* h, int32_t frameId, double curTime) {
    CREQ(h);
    h->s.frameId = frameId;
    h->s.curTime = (float)curTime;
    return GAIA_OK;
}
int gaia_b200_cloth_synchronize(gaia_cloth_handle* h) {
    CREQ(h); CFIN(h);
    CCUDA(h, cudaStreamSynchronize(h->s.stream));
    return GAIA_OK;
}

int gaia_b200_cloth_num_meshes(const gaia_cloth_handle* h) { return h ? (int)h->s.h.meshes.size() : GAIA_ERR_INVALID_ARG; }
int gaia_b200_cloth_mesh_sizes(const gaia_cloth_handle* h, int mesh, int32_t* nV, int32_t* nF, int32_t* nE) {
    CREQ(h);
    if (mesh < 0 || mesh >= (int)h->s.h.meshes.size()) return GAIA_ERR_INVALID_ARG;
    const ClothMesh& m = h->s.h.meshes[mesh];
    if (nV) *nV = m.nV;
    if (nF) *nF = m.nF;
    if (nE) *nE = m.nE;
    return GAIA_OK;
}

static int cloth_get(gaia_cloth_handle* h, const float4* src, int64_t origBegin, int64_t count, float* out) {
    ClothScene& s = h->s;
    if (!out) return cfail(s, GAIA_ERR_INVALID_ARG, "null output");
    CCUDA(h, cudaSetDevice(s.device));
    k_gather_xyz<<<nblk(count), kB, 0, s.stream>>>((int)count, (int)origBegin, s.o2s.p, src, s.ioStage.p);
    CCUDA(h, cudaMemcpyAsync(out, s.ioStage.p, 3 * (size_t)count * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
    CCUDA(h, cudaStreamSynchronize(s.stream));
    return GAIA_OK;
}
static int cloth_set(gaia_cloth_handle* h, float4* dst, int64_t origBegin, int64_t count, const float* in) {
    ClothScene& s = h->s;
    CCUDA(h, cudaSetDevice(s.device));
    CCUDA(h, cudaMemcpyAsync(s.ioStage.p, in, 3 * (size_t)count * sizeof(float), cudaMemcpyHostToDevice, s.stream));
    k_scatter_xyz<<<nblk(count), kB, 0, s.stream>>>((int)count, (int)origBegin, s.o2s.p, s.ioStage.p, dst);
    CCUDA(h, cudaStreamSynchronize(s.stream));
    return GAIA_OK;
}
#define CMESH(h, mesh) if ((mesh) < 0 || (mesh) >= (int)(h)->s.h.meshes.size()) return cfail((h)->s, GAIA_ERR_INVALID_ARG, "mesh id out of range")

int gaia_b200_cloth_get_positions(gaia_cloth_handle* h, int mesh, float* xyz) {
    CREQ(h); CFIN(h); CMESH(h, mesh);
    const ClothMesh& m = h->s.h.meshes[mesh];
    return cloth_get(h, h->s.pos.p, m.vBase, m.nV, xyz);
}
int gaia_b200_cloth_get_velocities(gaia_cloth_handle* h, int mesh, float* xyz) {
    CREQ(h); CFIN(h); CMESH(h, mesh);
    const ClothMesh& m = h->s.h.meshes[mesh];
    return cloth_get(h, h->s.vel.p, m.vBase, m.nV, xyz);
}
int gaia_b200_cloth_set_state(gaia_cloth_handle* h, int mesh, const float* pos, const float* vel) {
    CREQ(h); CFIN(h); CMESH(h, mesh);
    const ClothMesh& m = h->s.h.meshes[mesh];
    if (pos) { const int rc = cloth_set(h, h->s.pos.p, m.vBase, m.nV, pos); if (rc < 0) return rc; }
    if (vel) { const int rc = cloth_set(h, h->s.vel.p, m.vBase, m.nV, vel); if (rc < 0) return rc; }
    return GAIA_OK;
}
int64_t gaia_b200_cloth_total_vertices(const gaia_cloth_handle* h) {
    if (!h) return GAIA_ERR_INVALID_ARG;
    int64_t n = 0;
    for (auto& m : h->s.h.meshes) n += m.nV;
    return n;
}
int gaia_b200_cloth_get_all_positions(gaia_cloth_handle* h, float* xyz) {
    CREQ(h); CFIN(h);
    return cloth_get(h, h->s.pos.p, 0, h->s.h.N, xyz);
}
int gaia_b200_cloth_set_all_state(gaia_cloth_handle* h, const float* pos, const float* vel) {
    CREQ(h); CFIN(h);
    if (pos) { const int rc = cloth_set(h, h->s.pos.p, 0, h->s.h.N, pos); if (rc < 0) return rc; }
    if (vel) { const int rc = cloth_set(h, h->s.vel.p, 0, h->s.h.N, vel); if (rc < 0) return rc; }
    return GAIA_OK;
}

int gaia_b200_cloth_get_edges(const gaia_cloth_handle* h, int mesh, int32_t* out) {
    CREQ(h);
    if (mesh < 0 || mesh >= (int)h->s.h.meshes.size() || !out) return GAIA_ERR_INVALID_ARG;
    const ClothMesh& m = h->s.h.meshes[mesh];
    std::memcpy(out, m.edges.data(), m.edges.size() * sizeof(int32_t));
    return GAIA_OK;
}
int gaia_b200_cloth_num_parallel_groups(const gaia_cloth_handle* h) {
    CREQ(h);
    if (!h->s.finalized) return GAIA_ERR_STATE;
    return (int)h->s.h.groups.size();
}
int gaia_b200_cloth_get_parallel_group(const gaia_cloth_handle* h, int group, int32_t* out, int32_t* nPairs) {
    CREQ(h);
    if (!h->s.finalized) return GAIA_ERR_STATE;
    if (group < 0 || group >= (int)h->s.h.groups.size()) return GAIA_ERR_INVALID_ARG;
    const auto& g = h->s.h.groups[group];
    if (nPairs) *nPairs = (int32_t)(g.size() / 2);
    if (out) std::memcpy(out, g.data(), g.size() * sizeof(int32_t));
    return GAIA_OK;
}
int gaia_b200_cloth_get_vertex_masses(const gaia_cloth_handle* h, int mesh, float* out) {
    CREQ(h);
    if (mesh < 0 || mesh >= (int)h->s.h.meshes.size() || !out) return GAIA_ERR_INVALID_ARG;
    const ClothMesh& m = h->s.h.meshes[mesh];
    std::memcpy(out, m.mass.data(), m.mass.size() * sizeof(float));
    return GAIA_OK;
}

int gaia_b200_cloth_detect_contacts(gaia_cloth_handle* h) {
    CREQ(h); CFIN(h);
    ClothScene& s = h->s;
    CCUDA(h, cudaSetDevice(s.device));
    CCUDA(h, detect(s));
    CCUDA(h, cudaMemcpyAsync(s.h_flags, s.errFlags.p, N_FLAGS * sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream));
    CCUDA(h, cudaStreamSynchronize(s.stream));
    return check_flags(s);
}

int gaia_b200_cloth_get_vf_contacts(gaia_cloth_handle* h, gaia_cloth_vf_contact* out, int64_t capacity, int64_t* n) {
    CREQ(h); CFIN(h);
    ClothScene& s = h->s;
    CCUDA(h, cudaSetDevice(s.device));
    CCUDA(h, cudaStreamSynchronize(s.stream));
    const int64_t N = s.h.N;
    std::vector<int32_t> cnt(N);
    std::vector<int2> lst((size_t)N * KV);
    CCUDA(h, cudaMemcpy(cnt.data(), s.vfCount.p, N * sizeof(int32_t), cudaMemcpyDeviceToHost));
    CCUDA(h, cudaMemcpy(lst.data(), s.vfList.p, lst.size() * sizeof(int2), cudaMemcpyDeviceToHost));
    int64_t k = 0;
    for (auto& m : s.h.meshes)
        for (int v = 0; v < m.nV; v++) {
            const int32_t sd = m.sortedId[v];
            for (int i = 0; i < cnt[sd]; i++) {
                const int2 c = lst[(size_t)sd * KV + i];
                if (out && k < capacity) {
                    int fm = 0;
                    while (fm + 1 < (int)s.h.meshes.size() && c.x >= s.h.meshes[fm + 1].fBase) fm++;
                    out[k] = {m.id, v, fm, (int32_t)(c.x - s.h.meshes[fm].fBase), c.y >> 28, c.y & 0x0FFFFFFF};
                }
                k++;
            }
        }
    if (n) *n = k;
    return GAIA_OK;
}
int gaia_b200_cloth_get_ee_contacts(gaia_cloth_handle* h, gaia_cloth_ee_contact* out, int64_t capacity, int64_t* n) {
    CREQ(h); CFIN(h);
    ClothScene& s = h->s;
    CCUDA(h, cudaSetDevice(s.device));
    CCUDA(h, cudaStreamSynchronize(s.stream));
    const int64_t nE = s.h.nSimE;
    std::vector<int32_t> cnt(nE), lst((size_t)nE * KE);
    if (nE) {
        CCUDA(h, cudaMemcpy(cnt.data(), s.eeCount.p, nE * sizeof(int32_t), cudaMemcpyDeviceToHost));
        CCUDA(h, cudaMemcpy(lst.data(), s.eeList.p, lst.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
    }
    int64_t k = 0;
    for (int mi = 0; mi < s.h.numSim; mi++) {
        const ClothMesh& m = s.h.meshes[mi];
        for (int e = 0; e < m.nE; e++) {
            const int64_t eg = m.eBase + e;
            for (int i = 0; i < cnt[eg]; i++) {
                const int32_t te = lst[(size_t)eg * KE + i];
                if (out && k < capacity) {
                    int tm = 0;
                    while (tm + 1 < (int)s.h.meshes.size() && te >= s.h.meshes[tm + 1].eBase) tm++;
                    out[k] = {mi, e, tm, (int32_t)(te - s.h.meshes[tm].eBase)};
                }
                k++;
            }
        }
    }
    if (n) *n = k;
    return GAIA_OK;
}
int gaia_b200_cloth_get_conservative_bounds(gaia_cloth_handle* h, int mesh, float* out) {
    CREQ(h); CFIN(h); CMESH(h, mesh);
    ClothScene& s = h->s;
    if (!out) return cfail(s, GAIA_ERR_INVALID_ARG, "null output");
    const ClothMesh& m = s.h.meshes[mesh];
    CCUDA(h, cudaSetDevice(s.device));
    k_gather_f<<<nblk(m.nV), kB, 0, s.stream>>>(m.nV, (int)m.vBase, s.o2s.p, s.bounds.p, s.ioStage.p);
    CCUDA(h, cudaMemcpyAsync(out, s.ioStage.p, (size_t)m.nV * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
    CCUDA(h, cudaStreamSynchronize(s.stream));
    return GAIA_OK;
}

int gaia_b200_cloth_get_stats(gaia_cloth_handle* h, gaia_cloth_stats* out) {
    CREQ(h);
    if (!out) return GAIA_ERR_INVALID_ARG;
    ClothScene& s = h->s;
    if (s.finalized && s.detectedOnce) {
        CCUDA(h, cudaSetDevice(s.device));
        CCUDA(h, cudaStreamSynchronize(s.stream));
        std::vector<int32_t> a(s.h.N), b(s.h.nSimE);
        CCUDA(h, cudaMemcpy(a.data(), s.vfCount.p, a.size() * 4, cudaMemcpyDeviceToHost));
        if (!b.empty()) CCUDA(h, cudaMemcpy(b.data(), s.eeCount.p, b.size() * 4, cudaMemcpyDeviceToHost));
        int64_t nv = 0, ne = 0;
        for (int32_t x : a) nv += x;
        for (int32_t x : b) ne += x;
        s.stats.numContactsVF = nv; s.stats.numContactsEE = ne;
    }
    *out = s.stats;
    return GAIA_OK;
}
int gaia_b200_cloth_reset_stats(gaia_cloth_handle* h) {
    CREQ(h);
    const int64_t a = h->s.stats.numContactsVF, b = h->s.stats.numContactsEE;
    std::memset(&h->s.stats, 0, sizeof(h->s.stats));
    h->s.stats.numContactsVF = a; h->s.stats.numContactsEE = b;
    return GAIA_OK;
}
int gaia_b200_cloth_set_profiling(gaia_cloth_handle* h, int enabled) { CREQ(h); h->s.profiling = enabled != 0; return GAIA_OK; }
void* gaia_b200_cloth_cuda_stream(gaia_cloth_handle* h) { return h ? (void*)h->s.stream : nullptr; }

}  // extern "C"
