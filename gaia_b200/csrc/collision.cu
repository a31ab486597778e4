// Collision detection of the VBD step on the GPU: LBVH (Morton sort, Karras build, per-iteration refit) with DCD / CCD
// queries, contact-list preparation and the per-colour contact force/Hessian pre-pass. Replaces the reference's
// Embree/TBB detectors:
//   VBDPhysics::dcd / intermediateDCD / ccd / intermediateCCD        M/VBD/VBDPhysics.cpp:1958-2156
//   DiscreteCollisionDetector::vertexCollisionDetection / closestPointQuery (+ callbacks)
//                                                                     M/CollisionDetector/DiscreteCollisionDetector.cpp:12-319, 517-662
//   ContinuousCollisionDetector::vertexContinuousCollisionDetection  M/CollisionDetector/ContinuousCollisionDetector.cpp:70-274, 352-381
//   updateDCDBVH / updateCCDBVH (refit every detection, scheduled rebuilds)  M/VBD/VBDPhysics.cpp:2174-2187
//   prepareCollisionDataCPU + ActiveCollisionList                     M/VBD/VBDPhysics.cpp:1709-1810, M/VBD/ActiveCollisionList.cpp:43-110
//   updateCollisionInfo                                               M/VBD/VBDPhysics.cpp:2205-2311
//   Morton code                                                       M/BVH/MortonCode.h:8-31
//
// Embree's traversal order decides some outcomes in the reference ("first tet found", candidate order of the closest
// point query). This implementation applies the canonical rules of DESIGN.md §6 (R1-R4), the same ones the CPU oracle
// applies, so contact lists are well-defined and bit-identical between the two:
//   R1 containing tet with the smallest (meshId, tetId);  R2 closest acceptable face, ties -> smallest face id;
//   R3 CCD: smallest t*|p1-p0|, ties -> smallest (meshId, faceId);  R4 traversal stacks never drop entries (overflow is
//   reported through errFlags[3]).
// The BVHs only cull conservatively (boxes inflated by 1e-4 relative), so every result equals brute force.
//
// Compiled with -fmad=false: the predicates in geom.cuh round exactly like the CPU path.
#include "scene_impl.h"
#include "geom.cuh"
#include <cub/cub.cuh>
#include <algorithm>
#include <cstring>

namespace gb {

#define CK(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return _e; } while (0)
static const int kB = 128;
static const int kRejCap = 1536;   // DCD phase 2: rejected-face distances kept per query (rule R2')
static const int kCand = 64;        // DCD candidate cache: tets kept per surface vertex between refits (k_dcd_phase1)
static inline int nblk(int64_t n, int per = kB) { return (int)((n + per - 1) / per); }
// grid of a grid-stride kernel: at most one full wave of 148 SMs x 16 blocks (a launch that decides on the device to do nothing costs one wave)
static inline int capped(int64_t n, int per = kB) { return (int)std::min<int64_t>((n + per - 1) / per, 148 * 16); }

struct ColParamsDev {
    int32_t allowCCD, checkFeasibleRegion, checkTetTraverse, handleSelfCollision, stopAfterPassing, traverseNonSelf, shiftQueryToCenter;
    int32_t maxBVHQuery;
    float centerShiftLevel, feasibleEps, rayEps, maxSearchMult, offHeight, kCollision, dt;
};

// ------------------------------------------------------------------------------------------------ LBVH
// Node index space: internal node i in [0,n), leaf at sorted position j -> n + j. For a segmented scene (one tree per
// mesh, keys carry the segment id in their high word) the internal nodes of segment [lo,hi) are lo..hi-2, root = lo.
struct Lbvh {
    int n = 0;
    DevBuf<float4> primLo, primHi;       // by primitive id (inflated; lo.x > hi.x = empty)
    DevBuf<uint64_t> keyIn, keyOut;      // segment << 32 | morton30 (inactive primitives: 1 << 30)
    DevBuf<int32_t> idIn, order;         // order[j] = primitive at sorted position j
    DevBuf<int2> child;
    DevBuf<int32_t> parent;              // [2n]
    DevBuf<float4> box;                  // [2n][2]: lo, hi of a node side by side (one 32-byte sector per node visit)
    DevBuf<int32_t> flags;
    DevBuf<int2> segRange;               // per segment: [lo,hi) in sorted positions
    DevBuf<uint8_t> tmp;
    size_t tmpBytes = 0;
    bool built = false;
    std::vector<uint8_t> activeAtBuild;
    cudaError_t alloc(int n_, const std::vector<int2>& segs) {
        n = n_;
        if (!n) return cudaSuccess;
        CK(primLo.alloc(n)); CK(primHi.alloc(n)); CK(keyIn.alloc(n)); CK(keyOut.alloc(n)); CK(idIn.alloc(n)); CK(order.alloc(n));
        CK(child.alloc(n)); CK(parent.alloc(2 * (size_t)n)); CK(box.alloc(4 * (size_t)n)); CK(flags.alloc(n));
        CK(segRange.alloc(segs.size()));
        CK(cudaMemcpy(segRange.p, segs.data(), segs.size() * sizeof(int2), cudaMemcpyHostToDevice));
        cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, keyIn.p, keyOut.p, idIn.p, order.p, n);
        CK(tmp.alloc(tmpBytes + 16));
        std::vector<int32_t> iota(n);
        for (int i = 0; i < n; i++) iota[i] = i;
        CK(cudaMemcpy(idIn.p, iota.data(), (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice));
        return cudaSuccess;
    }
};

__device__ __forceinline__ int delta_key(const uint64_t* __restrict__ keys, int i, int j, int lo, int hi) {
    if (j < lo || j >= hi) return -1;
    const uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((unsigned)i ^ (unsigned)j);
    return __clzll((long long)(a ^ b));
}

// Karras 2012, one thread per internal node
__global__ void k_karras(int n, const uint64_t* __restrict__ keys, const int2* __restrict__ segRange, int2* __restrict__ child,
                         int32_t* __restrict__ parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int seg = (int)(keys[i] >> 32);
    const int2 r = segRange[seg];
    if (i == r.x) parent[i] = -1;                      // root of the segment (internal, or unused when the segment has one leaf)
    if (r.y - r.x == 1) { parent[n + i] = -1; return; }
    if (i >= r.y - 1) return;                          // the last position of a segment has no internal node
    const int lo = r.x, hi = r.y;
    const int d = (delta_key(keys, i, i + 1, lo, hi) - delta_key(keys, i, i - 1, lo, hi)) >= 0 ? 1 : -1;
    const int dmin = delta_key(keys, i, i - d, lo, hi);
    int lmax = 2;
    while (delta_key(keys, i, i + lmax * d, lo, hi) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta_key(keys, i, i + (l + t) * d, lo, hi) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta_key(keys, i, j, lo, hi);
    int s = 0, t = l;
    do {
        t = (t + 1) / 2;
        if (delta_key(keys, i, i + (s + t) * d, lo, hi) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int left = (min(i, j) == gamma) ? n + gamma : gamma;
    const int right = (max(i, j) == gamma + 1) ? n + gamma + 1 : gamma + 1;
    child[i] = make_int2(left, right);
    parent[left] = i;
    parent[right] = i;
}

__device__ __forceinline__ float4 ldcg4(const float4* p) { return __ldcg(p); }

// bottom-up refit: the second thread to reach a node merges its children (and leaves the node's counter at zero for the
// next refit). `need` (optional): device flag of the lazy refit - when it reads 0 the boxes are still conservative and the
// whole launch is a no-op.
__global__ void k_refit(int n, const int32_t* __restrict__ order, const float4* __restrict__ primLo, const float4* __restrict__ primHi,
                        const int2* __restrict__ child, const int32_t* __restrict__ parent, float4* box, int32_t* flags,
                        const int32_t* __restrict__ need) {
    if (need && !*need) return;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {     // capped grid: a skipped refit costs one wave
        const int prim = order[j];
        box[2 * (n + j)] = primLo[prim]; box[2 * (n + j) + 1] = primHi[prim];
        int cur = parent[n + j];
        while (cur >= 0) {
            __threadfence();
            if (atomicAdd(flags + cur, 1) == 0) break;
            flags[cur] = 0;
            const int2 c = child[cur];
            const float4 a0 = ldcg4(box + 2 * c.x), a1 = ldcg4(box + 2 * c.x + 1), b0 = ldcg4(box + 2 * c.y), b1 = ldcg4(box + 2 * c.y + 1);
            box[2 * cur] = make_float4(fminf(a0.x, b0.x), fminf(a0.y, b0.y), fminf(a0.z, b0.z), 0.f);
            box[2 * cur + 1] = make_float4(fmaxf(a1.x, b1.x), fmaxf(a1.y, b1.y), fmaxf(a1.z, b1.z), 0.f);
            cur = parent[cur];
        }
    }
}

// ------------------------------------------------------------------------------------------------ boxes and Morton keys
__device__ __forceinline__ int f2ord(float f) { const int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7FFFFFFF; }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }

// per-mesh bounds of all vertices (ordered-int atomics; warp-aggregated when a warp holds a single mesh)
__global__ void k_mesh_bounds(int n, const float4* __restrict__ pos, const int32_t* __restrict__ info, int32_t* bounds /*6 per mesh*/) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = s < n;
    const int mesh = in ? (__ldg(info + s) & 0x00FFFFFF) : -1;
    float4 p = in ? pos[s] : make_float4(0, 0, 0, 0);
    const unsigned full = 0xffffffffu;
    const int m0 = __shfl_sync(full, mesh, 0);
    if (__all_sync(full, mesh == m0) && m0 >= 0) {
        float lx = p.x, ly = p.y, lz = p.z, hx = p.x, hy = p.y, hz = p.z;
        for (int o = 16; o > 0; o >>= 1) {
            lx = fminf(lx, __shfl_xor_sync(full, lx, o)); ly = fminf(ly, __shfl_xor_sync(full, ly, o)); lz = fminf(lz, __shfl_xor_sync(full, lz, o));
            hx = fmaxf(hx, __shfl_xor_sync(full, hx, o)); hy = fmaxf(hy, __shfl_xor_sync(full, hy, o)); hz = fmaxf(hz, __shfl_xor_sync(full, hz, o));
        }
        if ((threadIdx.x & 31) == 0) {
            int32_t* b = bounds + 6 * m0;
            atomicMin(b + 0, f2ord(lx)); atomicMin(b + 1, f2ord(ly)); atomicMin(b + 2, f2ord(lz));
            atomicMax(b + 3, f2ord(hx)); atomicMax(b + 4, f2ord(hy)); atomicMax(b + 5, f2ord(hz));
        }
    } else if (in) {
        int32_t* b = bounds + 6 * mesh;
        atomicMin(b + 0, f2ord(p.x)); atomicMin(b + 1, f2ord(p.y)); atomicMin(b + 2, f2ord(p.z));
        atomicMax(b + 3, f2ord(p.x)); atomicMax(b + 4, f2ord(p.y)); atomicMax(b + 5, f2ord(p.z));
    }
}
__global__ void k_init_bounds(int nMeshes, int32_t* bounds) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 6 * (nMeshes + 1)) return;
    bounds[i] = (i % 6 < 3) ? f2ord(3.0e38f) : f2ord(-3.0e38f);
}
// entry nMeshes = union over meshes active for collision; output as floats (lo xyz, hi xyz)
__global__ void k_finish_bounds(int nMeshes, const int32_t* __restrict__ bounds, const MeshParamsDev* __restrict__ mp, float* __restrict__ normBox) {
    if (blockIdx.x || threadIdx.x) return;
    float u[6] = {3.0e38f, 3.0e38f, 3.0e38f, -3.0e38f, -3.0e38f, -3.0e38f};
    for (int m = 0; m < nMeshes; m++) {
        for (int k = 0; k < 6; k++) {
            const float v = ord2f(bounds[6 * m + k]);
            normBox[6 * m + k] = v;
            if (mp[m].activeCollide) u[k] = k < 3 ? fminf(u[k], v) : fmaxf(u[k], v);
        }
    }
    for (int k = 0; k < 6; k++) normBox[6 * nMeshes + k] = u[k];
}

__device__ __forceinline__ uint32_t expand_bits(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
// lbvh::morton_code (MortonCode.h:22-31) of the box centre normalised by the box `nb` (lo xyz, hi xyz)
__device__ __forceinline__ uint32_t morton_of(float3 lo, float3 hi, const float* __restrict__ nb) {
    float c[3] = {0.5f * (lo.x + hi.x), 0.5f * (lo.y + hi.y), 0.5f * (lo.z + hi.z)};
    uint32_t e[3];
#pragma unroll
    for (int d = 0; d < 3; d++) {
        const float ext = nb[3 + d] - nb[d];
        float u = ext > 0.f ? (c[d] - nb[d]) / ext : 0.f;
        u = fminf(fmaxf(u * 1024.0f, 0.0f), 1023.0f);
        e[d] = expand_bits((uint32_t)u);
    }
    return e[0] * 4 + e[1] * 2 + e[2];
}
static const float kInflate = 1e-4f;
__device__ __forceinline__ void store_box(float3 lo, float3 hi, float4* outLo, float4* outHi, float margin) {
    const float mx = kInflate * (1.f + fmaxf(fabsf(lo.x), fabsf(hi.x))) + margin;
    const float my = kInflate * (1.f + fmaxf(fabsf(lo.y), fabsf(hi.y))) + margin;
    const float mz = kInflate * (1.f + fmaxf(fabsf(lo.z), fabsf(hi.z))) + margin;
    *outLo = make_float4(lo.x - mx, lo.y - my, lo.z - mz, 0.f);
    *outHi = make_float4(hi.x + mx, hi.y + my, hi.z + mz, 0.f);
}
__device__ __forceinline__ void grow(float3& lo, float3& hi, float4 p) {
    lo.x = fminf(lo.x, p.x); lo.y = fminf(lo.y, p.y); lo.z = fminf(lo.z, p.z);
    hi.x = fmaxf(hi.x, p.x); hi.y = fmaxf(hi.y, p.y); hi.z = fmaxf(hi.z, p.z);
}

// which: 0 tets (4 vertices), 1 faces (3 vertices), 2 swept faces (3 vertices at pos and prev)
// segmented: key high word = mesh id and the code is normalised by the mesh's own box; else by the union box
template <int WHICH>
__global__ void k_prim_boxes(int n, const int4* __restrict__ prims, const int32_t* __restrict__ primMesh, const float4* __restrict__ pos,
                             const float4* __restrict__ prev, const MeshParamsDev* __restrict__ mp, const float* __restrict__ normBox, int nMeshes,
                             int segmented, int writeKeys, float4* __restrict__ primLo, float4* __restrict__ primHi, uint64_t* __restrict__ keys,
                             float margin, const int32_t* __restrict__ need) {
    if (need && !*need) return;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int4 v = __ldg(prims + i);
        const int mesh = WHICH == 0 ? __ldg(primMesh + i) : v.w;
        const uint64_t segHi = segmented ? ((uint64_t)(uint32_t)mesh << 32) : 0ull;
        if (!mp[mesh].activeCollide) {
            primLo[i] = make_float4(3e38f, 3e38f, 3e38f, 0.f); primHi[i] = make_float4(-3e38f, -3e38f, -3e38f, 0.f);
            if (writeKeys) keys[i] = segHi | (1ull << 30);
            continue;
        }
        float3 lo = make_float3(3e38f, 3e38f, 3e38f), hi = make_float3(-3e38f, -3e38f, -3e38f);
        grow(lo, hi, pos[v.x]); grow(lo, hi, pos[v.y]); grow(lo, hi, pos[v.z]);
        if (WHICH == 0) grow(lo, hi, pos[v.w]);
        if (WHICH == 2) { grow(lo, hi, prev[v.x]); grow(lo, hi, prev[v.y]); grow(lo, hi, prev[v.z]); }
        if (writeKeys) keys[i] = segHi | morton_of(lo, hi, normBox + 6 * (segmented ? mesh : nMeshes));
        store_box(lo, hi, primLo + i, primHi + i, margin);
    }
}

// Lazy refit of the DCD trees. Their boxes are inflated by `margin` when they are computed, around the positions kept in
// refPos; as long as no vertex has moved further than `margin` from refPos every box still contains its primitive, the
// culling stays conservative (queries test the primitives themselves, so results do not change), and the refit - the
// expensive part of a detection on a 10 M-tet mesh, repeated every few iterations by intermediateCollisionIterations - is
// skipped. The decision is taken on the device: one flag, read by the box / refit kernels.
__global__ void k_disp_check(int n, const float4* __restrict__ pos, const float4* __restrict__ ref, float margin2, int32_t* need) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = pos[i], b = __ldg(ref + i);
    const float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z;
    if (!(dx * dx + dy * dy + dz * dz <= margin2)) *need = 1;      // (NaN counts as moved)
}
__global__ void k_copy_ref(int n, const float4* __restrict__ pos, float4* __restrict__ ref, const int32_t* __restrict__ need) {
    if (!*need) return;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) ref[i] = pos[i];
}

// ------------------------------------------------------------------------------------------------ collision state
struct CollisionState {
    // static topology, global ids (tets / faces / surface-vertex slots of all meshes concatenated in mesh order)
    DevBuf<int4> tetSid;          // 4 sorted vertex ids
    DevBuf<int32_t> tetMesh;
    DevBuf<int4> tetNbr;          // neighbour global tet ids (-1 = boundary), face i opposite vertex i
    DevBuf<int4> faceSid;         // 3 sorted vertex ids, w = mesh id
    DevBuf<int4> faceInfo;        // x = belonging tet (global), y = id of that face in its tet
    DevBuf<int4> faceNbr;         // 3 neighbour global faces (-1 = none)
    DevBuf<int32_t> surfSid, surfMesh, surfNbrOff, surfNbr, slotOfSorted;
    DevBuf<int32_t> groupStart;   // G + 1
    int G = 0;
    Lbvh tetBvh, surfBvh, ccdBvh;
    DevBuf<float4> refPos;        // lazy refit of the DCD trees: positions their boxes were computed at
    DevBuf<int32_t> needRefit;    // device flag: some vertex left its margin (or a rebuild forces the refit)
    DevBuf<int32_t> candCount, cand;   // per surface vertex: tets it can enter before the next refit (k_dcd_phase1), -1 = too many
    float lazyMargin = 0.f;       // margin the NEXT refit will use
    // adaptive margin (Scene::lazyRefitMargin < 0): grown while refits come more often than every 6 detections, up to just under
    // half an edge (beyond it the one-ring of candidate tets doubles and the candidate cache overflows), taken back for good when
    // it does not help (bodies that move further than any margin between two detections) or when the cache overflows
    float marginBuilt = 0.f, marginBase = 0.f, marginCap = 0.f;   // margin of the CURRENT boxes; start value; upper bound
    int detSinceRefit = 0, shortAtCap = 0;
    bool adaptive = false, adaptFrozen = false;
    DevBuf<int32_t> boundsOrd;    // 6 * (nMeshes + 1)
    DevBuf<float> normBox;        // 6 * (nMeshes + 1)
    // dense results per surface-vertex slot
    DevBuf<int32_t> cNum, cHitTet, cFace, cFlags;   // flags: bit0 fromCCD, bit1 shortestPathFound, bits 8.. closest point type
    DevBuf<uint8_t> penetrated;
    DevBuf<float4> cF, cH0, cH1;
    DevBuf<float> cB2;
    DevBuf<int32_t> hitList;      // compacted slots for phase 2
    // prepared data
    DevBuf<int32_t> contactOfVertex, relCount, relOff;
    DevBuf<int2> rel;
    DevBuf<uint8_t> colliding;
    DevBuf<int32_t> groupCount, groupOff, groupList;   // groupCount: G + 4 (…, nDCD, nCCD, nHits, nActive)
    DevBuf<uint8_t> scanTmp;
    size_t scanTmpBytes = 0;
    std::vector<int32_t> h_groupCount, h_groupOff;
    int64_t nActive = 0;
    bool prepared = false;
    ColParamsDev cp;
};

static ColParamsDev make_col_params(const Scene& s) {
    ColParamsDev c;
    c.allowCCD = s.col.allowCCD; c.checkFeasibleRegion = s.col.checkFeasibleRegion; c.checkTetTraverse = s.col.checkTetTraverse;
    c.handleSelfCollision = s.col.handleSelfCollision; c.stopAfterPassing = s.col.stopTraversingAfterPassingQueryPoint;
    c.traverseNonSelf = s.col.tetrahedralTraverseForNonSelfIntersection; c.shiftQueryToCenter = s.col.shiftQueryPointToCenter;
    c.maxBVHQuery = s.col.maxNumberOfBVHQuery;
    c.centerShiftLevel = s.col.centerShiftLevel; c.feasibleEps = s.col.feasibleRegionEpsilon; c.rayEps = s.col.rayTriIntersectionEPSILON;
    c.maxSearchMult = s.col.maxSearchDistanceMultiplier; c.offHeight = s.phys.collisionOffHeight; c.kCollision = s.phys.collisionStiffness;
    c.dt = s.dt;
    return c;
}

cudaError_t collision_init(Scene& s) {
    if (!s.col.allowDCD && !s.col.allowCCD) return cudaSuccess;
    CollisionState* c = new CollisionState();
    s.cs = c;
    const int64_t nT = s.nTets, nF = s.nSurfFaces, nS = s.nSurfVerts, N = s.nVerts;
    std::vector<int4> tetSid(nT), tetNbr(nT), faceSid(nF), faceInfo(nF), faceNbr(nF);
    std::vector<int32_t> tetMesh(nT), surfSid(nS), surfMesh(nS), surfNbrOff(nS + 1, 0), surfNbr, slotOfSorted(N, -1);
    std::vector<int2> faceSegs, oneSegT(1), oneSegF(1);
    for (auto& m : s.meshes) {
        for (int t = 0; t < m.nT; t++) {
            const int32_t* tv = &m.tets[4 * (size_t)t];
            tetSid[m.tetBase + t] = make_int4(m.sortedId[tv[0]], m.sortedId[tv[1]], m.sortedId[tv[2]], m.sortedId[tv[3]]);
            tetMesh[m.tetBase + t] = m.id;
            const int32_t* nb = &m.tetNbrTets[4 * (size_t)t];
            auto g = [&](int32_t x) { return x < 0 ? -1 : (int32_t)(m.tetBase + x); };
            tetNbr[m.tetBase + t] = make_int4(g(nb[0]), g(nb[1]), g(nb[2]), g(nb[3]));
        }
        for (int f = 0; f < m.nSurfF(); f++) {
            const int32_t* fv = &m.surfFaces[3 * (size_t)f];
            faceSid[m.surfFBase + f] = make_int4(m.sortedId[fv[0]], m.sortedId[fv[1]], m.sortedId[fv[2]], m.id);
            faceInfo[m.surfFBase + f] = make_int4((int32_t)(m.tetBase + m.surfFaceTet[f]), m.surfFaceIdAtTet[f], 0, 0);
            const int32_t* nb = &m.surfFace3Nbr[3 * (size_t)f];
            auto g = [&](int32_t x) { return x < 0 ? -1 : (int32_t)(m.surfFBase + x); };
            faceNbr[m.surfFBase + f] = make_int4(g(nb[0]), g(nb[1]), g(nb[2]), 0);
        }
        for (int sv = 0; sv < m.nSurfV(); sv++) {
            const int64_t slot = m.surfVBase + sv;
            surfSid[slot] = m.sortedId[m.surfaceVIds[sv]];
            surfMesh[slot] = m.id;
            slotOfSorted[surfSid[slot]] = (int32_t)slot;
            for (int q = m.surfNbrOff[sv]; q < m.surfNbrOff[sv + 1]; q++) surfNbr.push_back(m.sortedId[m.surfNbrVerts[q]]);
            surfNbrOff[slot + 1] = (int32_t)surfNbr.size();
        }
        faceSegs.push_back(make_int2((int)m.surfFBase, (int)(m.surfFBase + m.nSurfF())));
    }
    oneSegT[0] = make_int2(0, (int)nT); oneSegF[0] = make_int2(0, (int)nF);
    auto up = [](auto& buf, const auto& v) -> cudaError_t {
        CK(buf.alloc(v.size()));
        if (v.empty()) return cudaSuccess;
        return cudaMemcpy(buf.p, v.data(), v.size() * sizeof(v[0]), cudaMemcpyHostToDevice);
    };
    CK(up(c->tetSid, tetSid)); CK(up(c->tetMesh, tetMesh)); CK(up(c->tetNbr, tetNbr)); CK(up(c->faceSid, faceSid));
    CK(up(c->faceInfo, faceInfo)); CK(up(c->faceNbr, faceNbr)); CK(up(c->surfSid, surfSid)); CK(up(c->surfMesh, surfMesh));
    CK(up(c->surfNbrOff, surfNbrOff)); CK(up(c->surfNbr, surfNbr)); CK(up(c->slotOfSorted, slotOfSorted));
    c->G = (int)s.groupStart.size() - 1;
    CK(up(c->groupStart, s.groupStart));
    if (s.col.allowDCD) {
        CK(c->tetBvh.alloc((int)nT, oneSegT)); CK(c->surfBvh.alloc((int)nF, faceSegs));
        CK(c->refPos.alloc(N + 1)); CK(c->needRefit.alloc(2));      // [0] refit needed, [1] vertices whose candidate list overflowed
        CK(c->candCount.alloc(std::max<int64_t>(nS, 1))); CK(c->cand.alloc((size_t)std::max<int64_t>(nS, 1) * kCand));
        CK(cudaMemset(c->candCount.p, 0xff, std::max<int64_t>(nS, 1) * sizeof(int32_t)));
        // margin = a fifth of the typical tet edge (cube root of 6 x the mean rest volume): refits become rare for slowly
        // deforming bodies; the larger boxes cost little because between refits queries run from the candidate cache
        double vol = 0.0; int64_t cnt = 0;
        for (auto& m : s.meshes) for (float v : m.restVol) { vol += std::fabs((double)v); cnt++; }
        const float edge = cnt ? (float)std::cbrt(6.0 * vol / (double)cnt) : 0.f;
        c->adaptive = s.lazyRefitMargin < 0.f;
        c->marginBase = (c->adaptive ? 0.2f : s.lazyRefitMargin) * edge;
        c->marginCap = 0.48f * edge;
        c->lazyMargin = c->marginBuilt = c->marginBase;
    }
    if (s.col.allowCCD) CK(c->ccdBvh.alloc((int)nF, oneSegF));
    const size_t nM = s.meshes.size();
    CK(c->boundsOrd.alloc(6 * (nM + 1))); CK(c->normBox.alloc(6 * (nM + 1)));
    CK(c->cNum.alloc(nS)); CK(c->cHitTet.alloc(nS)); CK(c->cFace.alloc(nS)); CK(c->cFlags.alloc(nS)); CK(c->penetrated.alloc(nS));
    CK(c->cF.alloc(nS)); CK(c->cH0.alloc(nS)); CK(c->cH1.alloc(nS)); CK(c->cB2.alloc(nS)); CK(c->hitList.alloc(nS));
    CK(cudaMemset(c->cNum.p, 0, nS * 4)); CK(cudaMemset(c->cFlags.p, 0, nS * 4)); CK(cudaMemset(c->penetrated.p, 0, nS));
    CK(cudaMemset(c->cHitTet.p, 0xff, nS * 4)); CK(cudaMemset(c->cFace.p, 0xff, nS * 4));
    CK(cudaMemset(c->cF.p, 0, nS * 16)); CK(cudaMemset(c->cH0.p, 0, nS * 16)); CK(cudaMemset(c->cH1.p, 0, nS * 16)); CK(cudaMemset(c->cB2.p, 0, nS * 4));
    CK(c->contactOfVertex.alloc(N)); CK(c->relCount.alloc(N + 1)); CK(c->relOff.alloc(N + 1)); CK(c->rel.alloc(3 * (size_t)nS + 1));
    CK(c->colliding.alloc(N));
    CK(c->groupCount.alloc(c->G + 4)); CK(c->groupOff.alloc(c->G + 1)); CK(c->groupList.alloc(4 * (size_t)nS + 1));
    cub::DeviceScan::ExclusiveSum(nullptr, c->scanTmpBytes, c->relCount.p, c->relOff.p, (int)(N + 1));
    CK(c->scanTmp.alloc(c->scanTmpBytes + 16));
    c->h_groupCount.assign(c->G + 4, 0); c->h_groupOff.assign(c->G + 1, 0);
    c->cp = make_col_params(s);
    return cudaSuccess;
}

void collision_destroy(Scene& s) { delete s.cs; s.cs = nullptr; }
void collision_invalidate(Scene& s) {
    if (!s.cs) return;
    s.cs->tetBvh.built = s.cs->surfBvh.built = s.cs->ccdBvh.built = false;
}

// ------------------------------------------------------------------------------------------------ BVH maintenance
static bool activity_changed(const Scene& s, std::vector<uint8_t>& snap) {
    bool ch = snap.size() != s.meshes.size();
    snap.resize(s.meshes.size());
    for (size_t i = 0; i < s.meshes.size(); i++) { const uint8_t a = s.meshes[i].activeForCollision; if (snap[i] != a) ch = true; snap[i] = a; }
    return ch;
}

static cudaError_t compute_norm_boxes(Scene& s) {
    CollisionState* c = s.cs;
    const int nM = (int)s.meshes.size();
    k_init_bounds<<<nblk(6 * (nM + 1)), kB, 0, s.stream>>>(nM, c->boundsOrd.p);
    k_mesh_bounds<<<nblk(s.nVerts, 256), 256, 0, s.stream>>>((int)s.nVerts, s.pos.p, s.info.p, c->boundsOrd.p);
    k_finish_bounds<<<1, 32, 0, s.stream>>>(nM, c->boundsOrd.p, s.meshParams.p, c->normBox.p);
    s.stats.kernelLaunches += 3;
    return cudaGetLastError();
}

static bool activity_differs(const Scene& s, const std::vector<uint8_t>& snap) {
    if (snap.size() != s.meshes.size()) return true;
    for (size_t i = 0; i < s.meshes.size(); i++) if (snap[i] != (uint8_t)s.meshes[i].activeForCollision) return true;
    return false;
}

// margin / need: lazy refit (DCD trees); need == nullptr: boxes and refit always run
template <int WHICH>
static cudaError_t update_bvh(Scene& s, Lbvh& b, bool rebuild, bool segmented, float margin = 0.f, const int32_t* need = nullptr) {
    CollisionState* c = s.cs;
    if (!b.n) return cudaSuccess;
    const int nM = (int)s.meshes.size();
    const bool ch = activity_changed(s, b.activeAtBuild);
    const bool doBuild = !b.built || rebuild || ch;
    if (doBuild) { CK(compute_norm_boxes(s)); need = nullptr; }
    const int4* prims = WHICH == 0 ? c->tetSid.p : c->faceSid.p;
    k_prim_boxes<WHICH><<<capped(b.n), kB, 0, s.stream>>>(b.n, prims, c->tetMesh.p, s.pos.p, s.prev.p, s.meshParams.p, c->normBox.p, nM,
                                                      segmented ? 1 : 0, doBuild ? 1 : 0, b.primLo.p, b.primHi.p, b.keyIn.p, margin, need);
    s.stats.kernelLaunches++;
    if (doBuild) {
        CK(cub::DeviceRadixSort::SortPairs(b.tmp.p, b.tmpBytes, b.keyIn.p, b.keyOut.p, b.idIn.p, b.order.p, b.n, 0, 64, s.stream));
        k_karras<<<nblk(b.n), kB, 0, s.stream>>>(b.n, b.keyOut.p, b.segRange.p, b.child.p, b.parent.p);
        s.stats.kernelLaunches += 3;
        b.built = true;
        CK(cudaMemsetAsync(b.flags.p, 0, (size_t)b.n * sizeof(int32_t), s.stream));     // (refits leave the counters at zero)
    }
    // (512-thread blocks: a refit the device decides to skip costs one look at the flag per block)
    k_refit<<<nblk(b.n, 512), 512, 0, s.stream>>>(b.n, b.order.p, b.primLo.p, b.primHi.p, b.child.p, b.parent.p, b.box.p, b.flags.p, need);
    s.stats.kernelLaunches++;
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ queries
__device__ __forceinline__ bool box_contains(float4 lo, float4 hi, V3 p) {
    return p.x >= lo.x && p.x <= hi.x && p.y >= lo.y && p.y <= hi.y && p.z >= lo.z && p.z <= hi.z;
}
__device__ __forceinline__ float box_dist2(float4 lo, float4 hi, V3 p) {
    const float dx = p.x < lo.x ? lo.x - p.x : (p.x > hi.x ? p.x - hi.x : 0.f);
    const float dy = p.y < lo.y ? lo.y - p.y : (p.y > hi.y ? p.y - hi.y : 0.f);
    const float dz = p.z < lo.z ? lo.z - p.z : (p.z > hi.z ? p.z - hi.z : 0.f);
    return dx * dx + dy * dy + dz * dz;
}

struct BvhView { int n; const int2* child; const float4* box; const int32_t* order; };
static BvhView view_of(const Lbvh& b) { BvhView v; v.n = b.n; v.child = b.child.p; v.box = b.box.p; v.order = b.order.p; return v; }

static const int kStack = 64;

// DCD phase 1 (vertexCollisionDetection + tetIntersectionFunc, rule R1): the containing tet with the smallest global id.
// One WARP per surface vertex: the nodes whose box holds the point form a frontier kept in shared memory; every round the 32
// lanes test 32 frontier nodes at once (internal node: its children join the frontier; leaf: the tet itself is tested) and
// the smallest containing tet id is warp-reduced at the end. A point on the surface of a regular mesh sits in the boxes of
// ~50 tets and of their ancestors, ~250 node visits that one thread would walk one dependent load after the other.
static const int kFrontier = 384;          // frontier capacity per warp (overflow -> errFlags[3], like the traversal stacks)
// Candidate cache (goes with the lazy refit): when the trees are refitted (*need == 1) every surface vertex keeps the ids of
// the tets it could possibly enter before the next refit - the non-incident tets whose lazy box, inflated by the margin once
// more, holds the vertex: both the vertex and the tet stay within `margin` of where they are now for as long as the refit is
// skipped. Between refits a detection only tests those ~30 tets per vertex instead of walking the 10 M-leaf tree; the set is
// a superset of the tets a fresh traversal would test, and the tests themselves are the same, so results do not change.
// A vertex whose candidates do not fit (kCand) keeps walking the tree.
__global__ void __launch_bounds__(kB)
k_dcd_phase1(int nSlots, int intermediate, ColParamsDev cp, BvhView bv, const int32_t* __restrict__ surfSid, const int32_t* __restrict__ surfMesh,
             const MeshParamsDev* __restrict__ mp, const float4* __restrict__ pos, const int4* __restrict__ tetSid,
             const int32_t* __restrict__ tetMesh, uint8_t* penetrated, int32_t* cNum, int32_t* cHitTet, int32_t* cFace, int32_t* cFlags,
             int32_t* hitList, int32_t* nHits, int32_t* errFlags, const int32_t* __restrict__ need, float margin, int32_t* candCount,
             int32_t* cand, int32_t* overflowed) {
    __shared__ int32_t frontierAll[(kB / 32) * kFrontier];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t* frontier = frontierAll + warp * kFrontier;
    const int slot = blockIdx.x * (kB / 32) + warp;
    if (slot >= nSlots) return;
    const int mesh = surfMesh[slot];
    if (!mp[mesh].activeCollide) return;
    const int s = surfSid[slot];
    const V3 p = v3(pos[s]);
    const bool collect = *need != 0;             // the trees were refitted for this detection: rebuild this vertex's candidates
    bool queried = true;
    if (intermediate) queried = penetrated[slot] || !cp.allowCCD;
    else {
        if (lane == 0) penetrated[slot] = 0;
        queried = p.y < cp.offHeight;
    }
    if (!queried && !collect) return;
    int best = 0x7FFFFFFF;
    const int cached = collect ? -1 : candCount[slot];
    if (cached >= 0) {
        for (int i = lane; i < cached; i += 32) {
            const int t = cand[(size_t)slot * kCand + i];
            const int4 tv = __ldg(tetSid + t);
            if (t < best && point_in_tet(p, v3(pos[tv.x]), v3(pos[tv.y]), v3(pos[tv.z]), v3(pos[tv.w]))) best = t;
        }
    } else if (bv.n > 0) {
        // the frontier only ever holds nodes whose box (inflated by m2) contains the point: children are tested before they are queued
        const float m2 = collect ? margin : 0.f;
        const V3 pLo = v3(p.x - m2, p.y - m2, p.z - m2), pHi = v3(p.x + m2, p.y + m2, p.z + m2);
        auto holds = [&](float4 lo, float4 hi) { return pHi.x >= lo.x && pLo.x <= hi.x && pHi.y >= lo.y && pLo.y <= hi.y && pHi.z >= lo.z && pLo.z <= hi.z; };
        const int root = bv.n > 1 ? 0 : bv.n;               // single-leaf scene: the leaf itself
        int qn = holds(bv.box[2 * root], bv.box[2 * root + 1]) ? 1 : 0;
        int nCand = 0;
        if (lane == 0) frontier[0] = root;
        __syncwarp();
        while (qn > 0) {
            const int m = qn < 32 ? qn : 32;
            int node = -1;
            if (lane < m) node = frontier[qn - 1 - lane];
            qn -= m;
            __syncwarp();
            int c0 = -1, c1 = -1, keep = -1;
            if (node >= bv.n) {
                const int t = bv.order[node - bv.n];
                const int g = tetMesh[t];
                if (cp.handleSelfCollision || g != mesh) {
                    const int4 tv = __ldg(tetSid + t);
                    if (!(g == mesh && (tv.x == s || tv.y == s || tv.z == s || tv.w == s))) {
                        keep = t;
                        if (queried && t < best && point_in_tet(p, v3(pos[tv.x]), v3(pos[tv.y]), v3(pos[tv.z]), v3(pos[tv.w]))) best = t;
                    }
                }
            } else if (node >= 0) {
                const int2 ch = bv.child[node];
                if (holds(bv.box[2 * ch.x], bv.box[2 * ch.x + 1])) c0 = ch.x;
                if (holds(bv.box[2 * ch.y], bv.box[2 * ch.y + 1])) c1 = ch.y;
            }
            if (collect) {
                const unsigned km = __ballot_sync(0xffffffffu, keep >= 0);
                if (keep >= 0) { const int at = nCand + __popc(km & ((1u << lane) - 1u)); if (at < kCand) cand[(size_t)slot * kCand + at] = keep; }
                nCand += __popc(km);
            }
            const int mine = (c0 >= 0) + (c1 >= 0);
            int incl = mine;                                 // inclusive prefix sum of the pushes over the warp
            for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
            const int nPush = __shfl_sync(0xffffffffu, incl, 31);
            if (qn + nPush > kFrontier) { if (lane == 0) atomicAdd(errFlags + 3, 1); nCand = kCand + 1; break; }
            int at = qn + incl - mine;
            if (c0 >= 0) frontier[at++] = c0;
            if (c1 >= 0) frontier[at] = c1;
            qn += nPush;
            __syncwarp();
        }
        if (collect && lane == 0) { candCount[slot] = nCand <= kCand ? nCand : -1; if (nCand > kCand) atomicAdd(overflowed, 1); }
    }
    if (!queried) return;
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane) return;
    // CollisionDetectionResult::clear() + the new result
    if (best != 0x7FFFFFFF) {
        cNum[slot] = 1; cHitTet[slot] = best; cFace[slot] = -1; cFlags[slot] = (NotFound << 8);
        if (!intermediate) penetrated[slot] = 1;
        hitList[atomicAdd(nHits, 1)] = slot;
    } else {
        cNum[slot] = 0; cHitTet[slot] = -1; cFace[slot] = -1; cFlags[slot] = (NotFound << 8);
    }
}

struct TopoView {
    const int4 *tetSid, *tetNbr, *faceSid, *faceInfo, *faceNbr;
    const int32_t *surfNbrOff, *surfNbr, *slotOfSorted;
};

__device__ __forceinline__ V3 face_normal(const float4* __restrict__ pos, int4 fv) {
    const V3 a = v3(pos[fv.x]), b = v3(pos[fv.y]), c = v3(pos[fv.z]);
    const V3 n = cross(b - a, c - a);
    return n / norm(n);
}

__device__ bool edge_feasible(V3 p, const float4* __restrict__ pos, const TopoView& tp, int face, int4 fv, int edgeId, int e1, int e2, float eps) {
    const int nbr = comp4(__ldg(tp.faceNbr + face), edgeId);
    if (nbr == -1) return true;
    const V3 v1 = v3(pos[e1]), v2 = v3(pos[e2]);
    const V3 fN1 = -face_normal(pos, fv), fN2 = -face_normal(pos, __ldg(tp.faceSid + nbr));
    const V3 AP = p - v1, BP = p - v2, AB = v2 - v1, BA = -AB;
    const float relaxed = -(sqnorm(AB)) * eps - 0.f;
    if (dot(AP, AB) < relaxed) return false;
    if (dot(BP, BA) < relaxed) return false;
    if (dot(AP, cross(fN1, AB)) < relaxed) return false;
    if (dot(AP, cross(fN2, BA)) < relaxed) return false;
    return true;
}
__device__ bool vertex_feasible(V3 p, const float4* __restrict__ pos, const TopoView& tp, int vSid, float eps) {
    const V3 A = v3(pos[vSid]), AP = p - A;
    const int slot = tp.slotOfSorted[vSid];
    for (int q = tp.surfNbrOff[slot]; q < tp.surfNbrOff[slot + 1]; q++) {
        const V3 BA = A - v3(pos[tp.surfNbr[q]]);
        const float relaxed = -(dot(BA, BA)) * eps - 0.f;
        if (dot(AP, BA) < relaxed) return false;
    }
    return true;
}
__device__ bool feasible_region(V3 p, const float4* __restrict__ pos, const TopoView& tp, int face, int4 fv, int type, float eps) {
    switch (type) {
        case AtInterior: return true;
        case AtAB: return edge_feasible(p, pos, tp, face, fv, 0, fv.x, fv.y, eps);
        case AtBC: return edge_feasible(p, pos, tp, face, fv, 1, fv.y, fv.z, eps);
        case AtAC: return edge_feasible(p, pos, tp, face, fv, 2, fv.z, fv.x, eps);
        case AtA: return vertex_feasible(p, pos, tp, fv.x, eps);
        case AtB: return vertex_feasible(p, pos, tp, fv.y, eps);
        case AtC: return vertex_feasible(p, pos, tp, fv.z, eps);
        default: return false;
    }
}

__device__ __forceinline__ void project_2d(const float4* __restrict__ pos, P2 out[4], int4 tv, int inFace, V3 ax0, V3 ax1, V3 origin) {
    const int ids[4] = {comp4(tv, kTetFaces[inFace][0]), comp4(tv, kTetFaces[inFace][1]), comp4(tv, kTetFaces[inFace][2]), comp4(tv, inFace)};
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const V3 d = v3(pos[ids[i]]) - origin;
        out[i].x = dot(ax0, d); out[i].y = dot(ax1, d);
    }
}

static const int kTravStack = 192;

// tetrahedralTraverseTo (rule R4: entries are never dropped; an overflow is an error)
__device__ bool tet_traverse(const float4* __restrict__ pos, const TopoView& tp, V3 rayOrigin, V3 rayDir, float maxDis, int startTet, int startFace,
                             int targetTet, float eps, int32_t* errFlags) {
    if (startTet == targetTet) return true;
    V3 ax0, ax1;
    build_orthonormal_basis(rayDir, ax0, ax1);
    P2 pts[4];
    project_2d(pos, pts, __ldg(tp.tetSid + startTet), startFace, ax0, ax1, rayOrigin);
    bool possible[3];
    exit_face_selection(pts, possible, eps);
    int stack[kTravStack];   // (tet << 2) | face
    int sp = 0;
    int ring[32];
    int rBegin = 0, rSize = 0;
    ring[0] = startTet; rSize = 1;
    for (int i = 0; i < 3; i++) if (possible[i]) stack[sp++] = (startTet << 2) | kTetFaces[startFace][i];
    int nTrav = 0;
    while (sp) {
        const int e = stack[--sp];
        const int hitFace = e & 3, prevTet = e >> 2;
        const int4 ptv = __ldg(tp.tetSid + prevTet);
        const int f0 = comp4(ptv, kTetFaces[hitFace][0]), f1 = comp4(ptv, kTetFaces[hitFace][1]), f2 = comp4(ptv, kTetFaces[hitFace][2]);
        const int cur = comp4(__ldg(tp.tetNbr + prevTet), hitFace);
        if (cur == targetTet) return true;
        if (cur == -1) return false;
        bool seen = false;
        for (int i = 0; i < rSize; i++) if (ring[(rBegin + i) & 31] == cur) { seen = true; break; }
        if (seen) continue;
        if (rSize < 32) { ring[(rBegin + rSize) & 31] = cur; rSize++; }
        else { ring[rBegin] = cur; rBegin = (rBegin + 1) & 31; }
        const int4 ctv = __ldg(tp.tetSid + cur);
        const int newV = (ctv.x ^ ctv.y ^ ctv.z ^ ctv.w) ^ f0 ^ f1 ^ f2;
        int inFace = 3;
        if (newV == ctv.x) inFace = 0; else if (newV == ctv.y) inFace = 1; else if (newV == ctv.z) inFace = 2;
        project_2d(pos, pts, ctv, inFace, ax0, ax1, rayOrigin);
        exit_face_selection(pts, possible, eps);
        const int4 cnb = __ldg(tp.tetNbr + cur);
        for (int iF = 0; iF < 3; iF++) {
            if (!possible[iF]) continue;
            const int exitFace = kTetFaces[inFace][iF];
            if (comp4(cnb, exitFace) == targetTet) return true;
            if (sp >= kTravStack) { atomicAdd(errFlags + 3, 1); return false; }
            stack[sp++] = (cur << 2) | exitFace;
            if ((!(nTrav % 5)) && maxDis > 0.f) {
                float barys[3];
                origin_barycentric_2d(pts[kExit3[iF][0]], pts[kExit3[iF][1]], pts[kExit3[iF][2]], barys);
                const int ids[4] = {comp4(ctv, kTetFaces[inFace][0]), comp4(ctv, kTetFaces[inFace][1]), comp4(ctv, kTetFaces[inFace][2]), comp4(ctv, inFace)};
                const V3 ip = (barys[0] * v3(pos[ids[kExit3[iF][0]]]) + barys[1] * v3(pos[ids[kExit3[iF][1]]])) + barys[2] * v3(pos[ids[kExit3[iF][2]]]);
                if (sqnorm(ip - rayOrigin) > maxDis * maxDis) return false;
            }
        }
        ++nTrav;
    }
    return false;
}

// DCD phase 2 (closestPointQuery + closestPointQueryFunc, rule R2) for the vertices phase 1 found inside a tet.
// ONE WARP PER HIT. The query is long and data-dependent: without the reference's Embree-order-dependent cap on visited
// candidates (rule R2 drops it), a vertex whose nearby faces all fail the feasibility / traversal tests walks through
// hundreds of candidate faces, each with its own tetrahedral traversal (measured: 8 ms for a few thousand hits with one
// thread per hit). Here all lanes run the nearest-first BVH descent redundantly (uniform control flow, shared stack),
// collect up to 32 candidate leaves, then each lane evaluates ONE candidate; the best (smallest distance, ties -> smallest
// face id) is reduced over the warp and shrinks the search radius for the rest of the descent. The result is the minimum
// over all acceptable faces, independent of the visiting order, i.e. identical to the sequential query.
__global__ void __launch_bounds__(kB)
k_dcd_phase2(const int32_t* __restrict__ nHits, const int32_t* __restrict__ hitList, ColParamsDev cp, BvhView bv, const int2* __restrict__ segRange,
             TopoView tp, const int32_t* __restrict__ surfSid, const int32_t* __restrict__ surfMesh, const int32_t* __restrict__ tetMesh,
             const float4* __restrict__ pos, const int32_t* __restrict__ cHitTet, int32_t* cFace, int32_t* cFlags, int32_t* errFlags) {
    __shared__ int sStack[kB / 32][kStack];
    __shared__ int sCand[kB / 32][32];
    // rule R2' (the order-independent form of maxNumberOfBVHQuery): an acceptable face at distance d is a valid answer only
    // while fewer than `cap` faces are strictly nearer. So the `cap` smallest distances of the evaluated, not accepted faces
    // are kept (sRej, thinned to the cap smallest whenever it fills up); once there are cap of them their largest, `bound`,
    // limits the search like a found face does: nothing beyond it can be a valid answer, and a query whose surroundings hold
    // no acceptable face ends after the nearest few hundred faces instead of scanning the whole mesh.
    __shared__ float sRej[kB / 32][kRejCap];
    int nRej = 0;
    const int cap = (cp.maxBVHQuery > 0 && 2 * cp.maxBVHQuery <= kRejCap) ? cp.maxBVHQuery : 0;      // 0: no cap
    float bound = __int_as_float(0x7f800000);
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (i >= *nHits) return;                       // whole warps leave together
    const unsigned full = 0xffffffffu;
    const int slot = hitList[i];
    const int meshQ = surfMesh[slot];
    const int sQ = surfSid[slot];
    const V3 q = v3(pos[sQ]);
    const int embraceTet = cHitTet[slot];
    const int g = tetMesh[embraceTet];
    const float kInf = __int_as_float(0x7f800000);
    float radius = kInf;
    int bestFace = -1, bestType = NotFound;
    const int2 seg = segRange[g];
    V3 target = q;
    if (cp.shiftQueryToCenter) {
        const int4 tv = __ldg(tp.tetSid + embraceTet);
        V3 cen = (((v3(0.f, 0.f, 0.f) + v3(pos[tv.x])) + v3(pos[tv.y])) + v3(pos[tv.z])) + v3(pos[tv.w]);
        cen = cen * 0.25f;
        target = (1.f - cp.centerShiftLevel) * q + cp.centerShiftLevel * cen;
    }
    if (seg.y > seg.x) {
        int sp = 0;
        sStack[w][sp++] = (seg.y - seg.x > 1) ? seg.x : bv.n + seg.x;
        __syncwarp();
        for (;;) {
            int nCand = 0;
            while (sp > 0 && nCand < 32) {
                const int node = sStack[w][--sp];
                const float reach = fminf(radius, bound);
                const float r2 = reach == kInf ? 1e30f : reach * reach * 1.0001f + 1e-12f;
                if (box_dist2(bv.box[2 * node], bv.box[2 * node + 1], q) > r2) continue;
                if (node >= bv.n) { sCand[w][nCand++] = bv.order[node - bv.n]; continue; }
                const int2 ch = bv.child[node];
                if (sp + 2 > kStack) { if (lane == 0) atomicAdd(errFlags + 3, 1); continue; }
                // nearer child on top of the stack: the radius shrinks sooner
                const float dx = box_dist2(bv.box[2 * ch.x], bv.box[2 * ch.x + 1], q), dy = box_dist2(bv.box[2 * ch.y], bv.box[2 * ch.y + 1], q);
                __syncwarp();
                if (dx <= dy) { sStack[w][sp++] = ch.y; sStack[w][sp++] = ch.x; } else { sStack[w][sp++] = ch.x; sStack[w][sp++] = ch.y; }
                __syncwarp();
            }
            __syncwarp();
            if (nCand == 0) break;
            // ---- one candidate per lane
            float myD = kInf; int myFace = 0x7FFFFFFF, myType = NotFound;
            float myRej = -1.f;
            if (lane < nCand) {
                const int face = sCand[w][lane];
                const int4 fv = __ldg(tp.faceSid + face);
                const V3 a = v3(pos[fv.x]), b = v3(pos[fv.y]), c = v3(pos[fv.z]);
                int type; V3 bary;
                const V3 cpnt = closest_point_triangle(q, a, b, c, bary, type);
                const float d = norm(q - cpnt);
                bool ok = true;
                if (g == meshQ) {
                    int cpId = -1;
                    if (type == AtA) cpId = fv.x; else if (type == AtB) cpId = fv.y; else if (type == AtC) cpId = fv.z;
                    if (cpId != -1 && cpId == sQ) ok = false;
                }
                if (ok && !(d < radius || (d == radius && bestFace >= 0 && face < bestFace))) ok = false;
                const bool inReach = d <= bound;          // beyond the cap-th nearest rejected face nothing is a valid answer
                if (!inReach) ok = false;
                if (ok && cp.checkFeasibleRegion && !feasible_region(q, pos, tp, face, fv, type, cp.feasibleEps)) ok = false;
                if (ok && cp.checkTetTraverse && (g == meshQ || cp.traverseNonSelf)) {
                    V3 tracing;
                    if (type == AtInterior) tracing = cpnt;
                    else tracing = cpnt * (1.f - cp.centerShiftLevel) + (cp.centerShiftLevel / 3.0f) * ((a + b) + c);
                    V3 dir = target - tracing;
                    const float len = norm(dir);
                    dir = dir / len;
                    const float maxDis = cp.stopAfterPassing ? cp.maxSearchMult * len : -1.f;
                    const int4 fi = __ldg(tp.faceInfo + face);
                    if (!tet_traverse(pos, tp, tracing, dir, maxDis, fi.x, fi.y, embraceTet, cp.rayEps, errFlags)) ok = false;
                }
                if (ok) { myD = d; myFace = face; myType = type; }
                else if (inReach && d < radius) myRej = d;
            }
            if (cap) {      // log the rejected distances (warp-aggregated append) ...
                const unsigned rm = __ballot_sync(full, myRej >= 0.f);
                if (myRej >= 0.f) sRej[w][nRej + __popc(rm & ((1u << lane) - 1u))] = myRej;      // (room for 32 is kept below)
                nRej += __popc(rm);
                __syncwarp();
                // ... and keep only the cap smallest: the first time cap of them are known, and whenever the log has doubled
                if (nRej >= cap && (bound == kInf || nRej >= 2 * cap || nRej + 32 > kRejCap)) {
                    // value of the cap-th smallest entry: bisection on the bit patterns (distances are non-negative floats)
                    uint32_t lo = 0u, hi = 0x7f800000u;
                    while (lo < hi) {
                        const uint32_t mid = lo + ((hi - lo) >> 1);
                        int c = 0;
                        for (int i = lane; i < nRej; i += 32) c += __float_as_uint(sRej[w][i]) <= mid;
                        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(full, c, o);
                        if (c >= cap) hi = mid; else lo = mid + 1;
                    }
                    bound = fminf(bound, __uint_as_float(lo));
                    int kept = 0;
                    for (int base = 0; base < nRej; base += 32) {
                        const float e = base + lane < nRej ? sRej[w][base + lane] : kInf;
                        const bool keep = e <= bound;
                        const unsigned km = __ballot_sync(full, keep);
                        __syncwarp();
                        if (keep) sRej[w][kept + __popc(km & ((1u << lane) - 1u))] = e;
                        kept += __popc(km);
                        __syncwarp();
                    }
                    nRej = kept;
                    if (nRej + 32 > kRejCap) { if (lane == 0) atomicAdd(errFlags + 3, 1); nRej = kRejCap - 32; }      // > 1000 equal distances
                }
            }
            __syncwarp();
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float od = __shfl_xor_sync(full, myD, o);
                const int of = __shfl_xor_sync(full, myFace, o), ot = __shfl_xor_sync(full, myType, o);
                if (od < myD || (od == myD && of < myFace)) { myD = od; myFace = of; myType = ot; }
            }
            if (myFace != 0x7FFFFFFF) { radius = myD; bestFace = myFace; bestType = myType; }
        }
        if (cap && bestFace >= 0) {      // faces strictly nearer than the answer and not acceptable: cap of them void it
            int cnt = 0;
            for (int i = lane; i < nRej; i += 32) cnt += sRej[w][i] < radius;
            for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(full, cnt, o);
            if (cnt >= cap) { bestFace = -1; bestType = NotFound; }
        }
    }
    if (lane == 0) {
        if (bestFace >= 0) { cFace[slot] = bestFace; cFlags[slot] = 2 | (bestType << 8); }
        else { cFace[slot] = -1; cFlags[slot] = (NotFound << 8); }
    }
}

// CCD (vertexContinuousCollisionDetection + continuousTriPointIntersectionFunc, rule R3)
__global__ void __launch_bounds__(kB)
k_ccd(int nSlots, ColParamsDev cp, BvhView bv, const int4* __restrict__ faceSid, const int32_t* __restrict__ surfSid,
      const int32_t* __restrict__ surfMesh, const MeshParamsDev* __restrict__ mp, const float4* __restrict__ pos, const float4* __restrict__ prev,
      const uint8_t* __restrict__ penetrated, int32_t* cNum, int32_t* cHitTet, int32_t* cFace, int32_t* cFlags, float4* cF, int32_t* errFlags) {
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= nSlots) return;
    const int mesh = surfMesh[slot];
    if (!mp[mesh].activeCollide) return;
    const int s = surfSid[slot];
    const V3 p1 = v3(pos[s]);
    if (!(p1.y < cp.offHeight) || penetrated[slot]) return;
    const V3 p0 = v3(prev[s]);
    const V3 mid = 0.5f * (p1 + p0);
    const float len = norm(p1 - p0);
    const float radius = len * 0.5f;
    const float r2 = radius * radius * 1.0001f + 1e-12f;
    float bestDepth = 0.f;
    int bestFace = -1;
    if (bv.n > 0) {
        int stack[kStack];
        int sp = 0;
        stack[sp++] = bv.n > 1 ? 0 : bv.n;
        while (sp) {
            const int node = stack[--sp];
            if (box_dist2(bv.box[2 * node], bv.box[2 * node + 1], mid) > r2) continue;
            if (node >= bv.n) {
                const int face = bv.order[node - bv.n];
                const int4 fv = __ldg(faceSid + face);
                if (!mp[fv.w].activeCollide) continue;
                if (fv.w == mesh && (fv.x == s || fv.y == s || fv.z == s)) continue;
                const V3 x0[3] = {v3(prev[fv.x]), v3(prev[fv.y]), v3(prev[fv.z])};
                const V3 x1[3] = {v3(pos[fv.x]), v3(pos[fv.y]), v3(pos[fv.z])};
                float tt = -1.f; V3 bary;
                if (ccd::tri_point(tt, x0, x1, p0, p1, bary)) {
                    const float depth = tt * len;
                    if (bestFace < 0 || depth < bestDepth || (depth == bestDepth && face < bestFace)) { bestDepth = depth; bestFace = face; }
                }
            } else {
                const int2 ch = bv.child[node];
                if (sp + 2 > kStack) { atomicAdd(errFlags + 3, 1); continue; }
                stack[sp++] = ch.x; stack[sp++] = ch.y;
            }
        }
    }
    if (bestFace >= 0) {
        cNum[slot] = 1; cHitTet[slot] = -1; cFace[slot] = bestFace; cFlags[slot] = 1 | 2 | (AtInterior << 8);
        cF[slot] = make_float4(0.f, 0.f, 0.f, bestDepth);
    } else {
        cNum[slot] = 0; cHitTet[slot] = -1; cFace[slot] = -1; cFlags[slot] = 1 | (NotFound << 8);
    }
}

__global__ void k_clear_results(int nSlots, int32_t* cNum, int32_t* cHitTet, int32_t* cFace, int32_t* cFlags) {
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= nSlots) return;
    cNum[slot] = 0; cHitTet[slot] = -1; cFace[slot] = -1; cFlags[slot] = (NotFound << 8);
}

cudaError_t collision_dcd(Scene& s, bool intermediate) {
    CollisionState* c = s.cs;
    if (!c) return cudaSuccess;
    const int nS = (int)s.nSurfVerts;
    c->prepared = false;
    if (s.col.allowDCD) {
        bool rebuildTet = false, rebuildSurf = false, forced = false;
        if (!intermediate) {
            rebuildTet = (s.substep == 0) && !(s.frameId % s.phys.dcdTetMeshSceneBVHRebuildSteps) && s.frameId;
            rebuildSurf = (s.substep == 0) && !(s.frameId % s.phys.dcdSurfaceSceneBVHRebuildSteps) && s.frameId;
        }
        {
            const int N = (int)s.nVerts;
            const bool force = rebuildTet || rebuildSurf || !c->tetBvh.built || !c->surfBvh.built ||
                               activity_differs(s, c->tetBvh.activeAtBuild) || activity_differs(s, c->surfBvh.activeAtBuild);
            forced = force;
            CK(cudaMemsetAsync(c->needRefit.p, 0, 2 * sizeof(int32_t), s.stream));
            if (force) CK(cudaMemsetAsync(c->needRefit.p, 0x01, sizeof(int32_t), s.stream));
            else      // (the boxes in use were inflated by marginBuilt; a refit in this detection inflates by lazyMargin)
                k_disp_check<<<nblk(N), kB, 0, s.stream>>>(N, s.pos.p, c->refPos.p, c->marginBuilt * c->marginBuilt, c->needRefit.p);
            CK(update_bvh<0>(s, c->tetBvh, rebuildTet, false, c->lazyMargin, c->needRefit.p));
            CK(update_bvh<1>(s, c->surfBvh, rebuildSurf, true, c->lazyMargin, c->needRefit.p));
            k_copy_ref<<<capped(N), kB, 0, s.stream>>>(N, s.pos.p, c->refPos.p, c->needRefit.p);
            s.stats.kernelLaunches += 2;
        }
        int32_t* nHits = c->groupCount.p + c->G + 2;
        CK(cudaMemsetAsync(nHits, 0, sizeof(int32_t), s.stream));
        TopoView tp{c->tetSid.p, c->tetNbr.p, c->faceSid.p, c->faceInfo.p, c->faceNbr.p, c->surfNbrOff.p, c->surfNbr.p, c->slotOfSorted.p};
        k_dcd_phase1<<<nblk((int64_t)nS * 32), kB, 0, s.stream>>>(nS, intermediate ? 1 : 0, c->cp, view_of(c->tetBvh), c->surfSid.p, c->surfMesh.p, s.meshParams.p,
                                                  s.pos.p, c->tetSid.p, c->tetMesh.p, c->penetrated.p, c->cNum.p, c->cHitTet.p, c->cFace.p,
                                                  c->cFlags.p, c->hitList.p, nHits, s.errFlags.p, c->needRefit.p, c->lazyMargin, c->candCount.p, c->cand.p, c->needRefit.p + 1);
        // phase 2 runs over the compacted hit list; its length stays on the device (grid sized for the worst case, early exit)
        int32_t hits = 0, refit[2] = {0, 0};
        CK(cudaMemcpyAsync(&hits, nHits, sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream));
        CK(cudaMemcpyAsync(refit, c->needRefit.p, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream));
        CK(cudaStreamSynchronize(s.stream));
        if (refit[0]) {      // the trees were refitted with lazyMargin in this detection
            const int interval = c->detSinceRefit;
            c->detSinceRefit = 0;
            c->marginBuilt = c->lazyMargin;
            if (c->adaptive && !c->adaptFrozen && !forced) {
                if (refit[1] > nS / 50) { c->lazyMargin = std::max(c->marginBase, c->marginBuilt / 1.3f); c->marginCap = c->lazyMargin; }
                else if (c->marginBuilt < 0.999f * c->marginCap) { if (interval < 6) c->lazyMargin = std::min(c->marginCap, c->marginBuilt * 1.3f); }
                else if (interval < 3) { if (++c->shortAtCap >= 4) { c->lazyMargin = c->marginBase; c->adaptFrozen = true; } }
                else c->shortAtCap = 0;
            }
        }
        c->detSinceRefit++;
        if (hits > 0)
            k_dcd_phase2<<<nblk((int64_t)hits * 32), kB, 0, s.stream>>>(nHits, c->hitList.p, c->cp, view_of(c->surfBvh), c->surfBvh.segRange.p, tp,
                                                                      c->surfSid.p, c->surfMesh.p, c->tetMesh.p, s.pos.p, c->cHitTet.p, c->cFace.p,
                                                                      c->cFlags.p, s.errFlags.p);
        s.stats.kernelLaunches += 2;
    } else if (s.col.allowCCD && !intermediate) {
        k_clear_results<<<nblk(nS), kB, 0, s.stream>>>(nS, c->cNum.p, c->cHitTet.p, c->cFace.p, c->cFlags.p);
        s.stats.kernelLaunches++;
    }
    return cudaGetLastError();
}

cudaError_t collision_ccd(Scene& s) {
    CollisionState* c = s.cs;
    if (!c || !s.col.allowCCD) return cudaSuccess;
    const int nS = (int)s.nSurfVerts;
    c->prepared = false;
    const bool rebuild = (s.substep == 0) && !(s.frameId % s.phys.ccdBVHRebuildSteps) && s.frameId;
    CK(update_bvh<2>(s, c->ccdBvh, rebuild, false));
    k_ccd<<<nblk(nS), kB, 0, s.stream>>>(nS, c->cp, view_of(c->ccdBvh), c->faceSid.p, c->surfSid.p, c->surfMesh.p, s.meshParams.p, s.pos.p, s.prev.p,
                                       c->penetrated.p, c->cNum.p, c->cHitTet.p, c->cFace.p, c->cFlags.p, c->cF.p, s.errFlags.p);
    s.stats.kernelLaunches++;
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ prepare
__device__ __forceinline__ int group_of(const int32_t* __restrict__ groupStart, int G, int sid) {
    int lo = 0, hi = G;   // groupStart[lo] <= sid < groupStart[hi]
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (groupStart[mid] <= sid) lo = mid; else hi = mid; }
    return lo;
}

// pass 0 counts, pass 1 fills (relation slots are handed out from the back of each vertex's range)
template <int PASS>
__global__ void k_prepare(int nSlots, const int32_t* __restrict__ cNum, const int32_t* __restrict__ cFace, const int32_t* __restrict__ cFlags,
                          const int32_t* __restrict__ surfSid, const int4* __restrict__ faceSid, const int32_t* __restrict__ groupStart, int G,
                          int32_t* contactOfVertex, int32_t* relCount, const int32_t* __restrict__ relOff, int2* rel, uint8_t* colliding,
                          int32_t* groupCount, const int32_t* __restrict__ groupOff, int32_t* groupList) {
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= nSlots) return;
    if (!cNum[slot]) return;
    const int fl = cFlags[slot];
    if (PASS == 0) atomicAdd(groupCount + G + ((fl & 1) ? 1 : 0), 1);    // numActiveDCD / numActiveCCD
    if (!(fl & 2)) return;
    const int sq = surfSid[slot];
    const int4 fv = __ldg(faceSid + cFace[slot]);
    const int fvs[3] = {fv.x, fv.y, fv.z};
    int groups[4];
    int nG = 0;
    groups[nG++] = group_of(groupStart, G, sq);
    for (int k = 0; k < 3; k++) {
        const int g = group_of(groupStart, G, fvs[k]);
        bool has = false;
        for (int q = 0; q < nG; q++) has |= groups[q] == g;
        if (!has) groups[nG++] = g;
    }
    if (PASS == 0) {
        contactOfVertex[sq] = slot;
        colliding[sq] = 1;
        for (int k = 0; k < 3; k++) { colliding[fvs[k]] = 1; atomicAdd(relCount + fvs[k], 1); }
        for (int q = 0; q < nG; q++) atomicAdd(groupCount + groups[q], 1);
        atomicAdd(groupCount + G + 3, 1);
    } else {
        for (int k = 0; k < 3; k++) {
            const int at = atomicSub(relCount + fvs[k], 1) - 1;
            rel[relOff[fvs[k]] + at] = make_int2(slot, k);
        }
        for (int q = 0; q < nG; q++) groupList[groupOff[groups[q]] + atomicSub(groupCount + groups[q], 1) - 1] = slot;
    }
}

// relations of a vertex in ascending (contact slot, vertex order): the order the CPU path pushes them in.
// One thread per vertex owns its segment (an earlier version let "the contact holding the smallest entry" sort the segment:
// other contacts scanned it for the minimum WHILE it was being shifted, could elect themselves as well, and corrupted it).
__global__ void k_sort_relations(int nVerts, const int32_t* __restrict__ relOff, int2* rel) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nVerts) return;
    const int b = relOff[v], e = relOff[v + 1];
    if (e - b < 2) return;
    for (int i = b + 1; i < e; i++) {
        const int2 key = rel[i];
        int j = i - 1;
        while (j >= b && (rel[j].x > key.x || (rel[j].x == key.x && rel[j].y > key.y))) { rel[j + 1] = rel[j]; j--; }
        rel[j + 1] = key;
    }
}
__global__ void k_group_offsets(int G, const int32_t* __restrict__ groupCount, int32_t* groupOff) {
    if (blockIdx.x || threadIdx.x) return;
    int run = 0;
    for (int g = 0; g < G; g++) { groupOff[g] = run; run += groupCount[g]; }
    groupOff[G] = run;
}

cudaError_t collision_prepare(Scene& s) {
    CollisionState* c = s.cs;
    if (!c) return cudaSuccess;
    const int nS = (int)s.nSurfVerts;
    const int64_t N = s.nVerts;
    const int G = c->G;
    CK(cudaMemsetAsync(c->contactOfVertex.p, 0xff, N * sizeof(int32_t), s.stream));
    CK(cudaMemsetAsync(c->relCount.p, 0, (N + 1) * sizeof(int32_t), s.stream));
    CK(cudaMemsetAsync(c->colliding.p, 0, N, s.stream));
    CK(cudaMemsetAsync(c->groupCount.p, 0, (G + 2) * sizeof(int32_t), s.stream));
    CK(cudaMemsetAsync(c->groupCount.p + G + 3, 0, sizeof(int32_t), s.stream));
    k_prepare<0><<<nblk(nS), kB, 0, s.stream>>>(nS, c->cNum.p, c->cFace.p, c->cFlags.p, c->surfSid.p, c->faceSid.p, c->groupStart.p, G,
                                              c->contactOfVertex.p, c->relCount.p, c->relOff.p, c->rel.p, c->colliding.p, c->groupCount.p,
                                              c->groupOff.p, c->groupList.p);
    CK(cudaMemcpyAsync(c->h_groupCount.data(), c->groupCount.p, (G + 4) * sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream));
    CK(cudaStreamSynchronize(s.stream));
    s.stats.kernelLaunches++;
    s.stats.numContactsDCD = c->h_groupCount[G]; s.stats.numContactsCCD = c->h_groupCount[G + 1];
    c->nActive = c->h_groupCount[G + 3];
    int run = 0;
    for (int g = 0; g < G; g++) { c->h_groupOff[g] = run; run += c->h_groupCount[g]; }
    c->h_groupOff[G] = run;
    c->prepared = true;
    // (a captured colour loop always runs the contact variant with device-sized pre-passes: its tables must be valid,
    //  i.e. all-empty, when nothing is in contact)
    if (c->nActive == 0 && !collision_graphable(s)) return cudaGetLastError();
    CK(cub::DeviceScan::ExclusiveSum(c->scanTmp.p, c->scanTmpBytes, c->relCount.p, c->relOff.p, (int)(N + 1), s.stream));
    k_group_offsets<<<1, 32, 0, s.stream>>>(G, c->groupCount.p, c->groupOff.p);
    k_prepare<1><<<nblk(nS), kB, 0, s.stream>>>(nS, c->cNum.p, c->cFace.p, c->cFlags.p, c->surfSid.p, c->faceSid.p, c->groupStart.p, G,
                                              c->contactOfVertex.p, c->relCount.p, c->relOff.p, c->rel.p, c->colliding.p, c->groupCount.p,
                                              c->groupOff.p, c->groupList.p);
    k_sort_relations<<<nblk(N, 256), 256, 0, s.stream>>>((int)N, c->relOff.p, c->rel.p);
    s.stats.kernelLaunches += 4;
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ per-colour contact pre-pass
// updateCollisionInfo for the active contacts of one parallel group (VBDPhysics.cpp:1675-1680, 2205-2311)
__global__ void __launch_bounds__(kB)
k_update_infos(int n, const int32_t* __restrict__ list, const int32_t* __restrict__ groupOff, int group, ColParamsDev cp, const int32_t* __restrict__ surfSid, const int32_t* __restrict__ surfMesh,
               const int32_t* __restrict__ cFace, const int4* __restrict__ faceSid, const float4* __restrict__ pos, const float4* __restrict__ prev,
               const MeshParamsDev* __restrict__ mp, float4* cF, float4* cH0, float4* cH1, float* cB2) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (groupOff) {      // device-sized launch (captured colour loop): the group's run of the list is read here
        const int b = groupOff[group];
        n = groupOff[group + 1] - b;
        list += b;
    }
    if (i >= n) return;
    const int slot = list[i];
    const int sq = surfSid[slot];
    const int4 fv = __ldg(faceSid + cFace[slot]);
    const V3 a = v3(pos[fv.x]), b = v3(pos[fv.y]), c = v3(pos[fv.z]), p = v3(pos[sq]);
    V3 bary; int type;
    const V3 cpnt = closest_point_triangle(p, a, b, c, bary, type);
    V3 faceN = cross(b - a, c - a); faceN = faceN / norm(faceN);
    V3 n_ = faceN;
    if (type != AtInterior) {
        V3 d = cpnt - p;
        d = d / norm(d);
        if (dot(d, faceN) < 0) d = -d;
        n_ = d;
    }
    const V3 dx3 = p - v3(prev[sq]);
    const V3 dx0 = a - v3(prev[fv.x]), dx1 = b - v3(prev[fv.y]), dx2 = c - v3(prev[fv.z]);
    const V3 dx = dx3 - ((bary.x * dx0 + bary.y * dx1) + bary.z * dx2);
    V3 e0 = b - a; e0 = e0 / norm(e0);
    V3 e1 = cross(e0, n_); e1 = e1 / norm(e1);
    const float u0 = dot(e0, dx), u1 = dot(e1, dx);
    const MeshParamsDev* mh = mp + fv.w;
    const MeshParamsDev* mc = mp + surfMesh[slot];
    const float mu = (float)(((double)(mh->frictionDynamic + mc->frictionDynamic)) * 0.5);
    const float epsV = (float)(((double)(mh->frictionEpsV + mc->frictionEpsV)) * 0.5);
    const float epsU = epsV * cp.dt;
    const V3 diff = cpnt - p;
    const float depth = dot(diff, n_);
    const float k = cp.kCollision;
    const float lambda = depth * k;
    V3 F = v3(0.f, 0.f, 0.f);
    float H[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};   // 00 01 02 11 12 22
    if (depth > 0.f) {
        const float uNorm = sqrtf(u0 * u0 + u1 * u1);
        if (uNorm > 0) {
            const float f1 = (uNorm > epsU) ? 1 / uNorm : (-uNorm / epsU + 2) / epsU;
            const float sc = -mu * lambda * f1;
            F = v3(sc * (e0.x * u0 + e1.x * u1), sc * (e0.y * u0 + e1.y * u1), sc * (e0.z * u0 + e1.z * u1));
            const float kk = mu * lambda * f1;
            H[0] = kk * (e0.x * e0.x + e1.x * e1.x); H[1] = kk * (e0.x * e0.y + e1.x * e1.y); H[2] = kk * (e0.x * e0.z + e1.x * e1.z);
            H[3] = kk * (e0.y * e0.y + e1.y * e1.y); H[4] = kk * (e0.y * e0.z + e1.y * e1.z); H[5] = kk * (e0.z * e0.z + e1.z * e1.z);
        }
        F = F + lambda * n_;
        H[0] += (k * n_.x) * n_.x; H[1] += (k * n_.x) * n_.y; H[2] += (k * n_.x) * n_.z;
        H[3] += (k * n_.y) * n_.y; H[4] += (k * n_.y) * n_.z; H[5] += (k * n_.z) * n_.z;
    }
    cF[slot] = make_float4(F.x, F.y, F.z, depth);
    cH0[slot] = make_float4(H[0], H[1], H[2], H[3]);
    cH1[slot] = make_float4(H[4], H[5], bary.x, bary.y);
    cB2[slot] = bary.z;
}

// A colour loop captured into a CUDA graph cannot size the pre-pass from this substep's host-side counts: it launches
// over the capacity of the list instead and reads the group's run on the device. Worth it only while that capacity is a
// few hundred blocks (small scenes, which are the launch-bound ones).
static constexpr int64_t kGraphSurfCap = 1 << 18;
bool collision_graphable(const Scene& s) { return s.cs && s.nSurfVerts <= kGraphSurfCap; }

cudaError_t collision_update_infos(Scene& s, int group, bool deviceSized) {
    CollisionState* c = s.cs;
    if (!c || !c->prepared) return cudaSuccess;
    if (deviceSized) {
        k_update_infos<<<nblk(s.nSurfVerts), kB, 0, s.stream>>>(0, c->groupList.p, c->groupOff.p, group, c->cp, c->surfSid.p, c->surfMesh.p, c->cFace.p,
                                                              c->faceSid.p, s.pos.p, s.prev.p, s.meshParams.p, c->cF.p, c->cH0.p, c->cH1.p, c->cB2.p);
        s.stats.kernelLaunches++;
        return cudaSuccess;
    }
    const int n = c->h_groupCount[group];
    if (n <= 0) return cudaSuccess;
    k_update_infos<<<nblk(n), kB, 0, s.stream>>>(n, c->groupList.p + c->h_groupOff[group], nullptr, 0, c->cp, c->surfSid.p, c->surfMesh.p, c->cFace.p,
                                               c->faceSid.p, s.pos.p, s.prev.p, s.meshParams.p, c->cF.p, c->cH0.p, c->cH1.p, c->cB2.p);
    s.stats.kernelLaunches++;
    return cudaSuccess;
}

bool collision_contact_view(Scene& s, ContactView& cv, const uint8_t** collidingMask, bool force) {
    CollisionState* c = s.cs;
    if (!c || !c->prepared) { *collidingMask = nullptr; return false; }
    *collidingMask = c->colliding.p;
    if (c->nActive == 0 && !force) return false;
    cv.contactOfVertex = c->contactOfVertex.p; cv.relOff = c->relOff.p; cv.rel = c->rel.p;
    cv.cF = c->cF.p; cv.cH0 = c->cH0.p; cv.cH1 = c->cH1.p; cv.cB2 = c->cB2.p;
    return true;
}

// ------------------------------------------------------------------------------------------------ parity surfaces
int64_t collision_fetch_contacts(Scene& s, gaia_contact* out, int64_t cap) {
    CollisionState* c = s.cs;
    if (!c) return 0;
    const int64_t nS = s.nSurfVerts;
    std::vector<int32_t> num(nS), tet(nS), face(nS), flags(nS);
    if (cudaStreamSynchronize(s.stream) != cudaSuccess) return -1;
    if (cudaMemcpy(num.data(), c->cNum.p, nS * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    cudaMemcpy(tet.data(), c->cHitTet.p, nS * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(face.data(), c->cFace.p, nS * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(flags.data(), c->cFlags.p, nS * 4, cudaMemcpyDeviceToHost);
    auto mesh_of = [&](int64_t id, auto baseOf) {
        size_t lo = 0, hi = s.meshes.size();
        while (hi - lo > 1) { const size_t mid = (lo + hi) / 2; if (baseOf(s.meshes[mid]) <= id) lo = mid; else hi = mid; }
        return (int)lo;
    };
    int64_t cnt = 0;
    for (auto& m : s.meshes)
        for (int sv = 0; sv < m.nSurfV(); sv++) {
            const int64_t slot = m.surfVBase + sv;
            if (!num[slot]) continue;
            if (out && cnt < cap) {
                gaia_contact k;
                k.queryMesh = m.id; k.queryVertex = m.surfaceVIds[sv];
                k.fromCCD = flags[slot] & 1; k.shortestPathFound = (flags[slot] >> 1) & 1; k.closestPointType = flags[slot] >> 8;
                int hitMesh = -1;
                if (tet[slot] >= 0) { hitMesh = mesh_of(tet[slot], [](const HostMesh& h) { return h.tetBase; }); k.hitTet = (int32_t)(tet[slot] - s.meshes[hitMesh].tetBase); }
                else k.hitTet = -1;
                if (face[slot] >= 0) {
                    const int fm = mesh_of(face[slot], [](const HostMesh& h) { return h.surfFBase; });
                    if (hitMesh < 0) hitMesh = fm;
                    k.hitFace = (int32_t)(face[slot] - s.meshes[fm].surfFBase);
                } else k.hitFace = -1;
                k.hitMesh = hitMesh;
                out[cnt] = k;
            }
            cnt++;
        }
    return cnt;
}

// a15: the per-group active list (ActiveCollisionList::addToActiveCollisionList, M/VBD/ActiveCollisionList.cpp:43-110) as
// (mesh, vertex) pairs of the contacts' QUERY vertices, sorted ascending (the device fills the list with atomics; the
// canonical order of the parity surface is the sorted one). Returns the number of pairs, -1 on a CUDA error.
int64_t collision_fetch_active_list(Scene& s, int group, int32_t* out, int64_t cap) {
    CollisionState* c = s.cs;
    if (!c || !c->prepared || c->nActive == 0) return 0;
    if (group < 0 || group >= c->G) return -2;
    if (cudaStreamSynchronize(s.stream) != cudaSuccess) return -1;
    const int n = c->h_groupCount[group];
    if (n <= 0 || !out) return std::max(n, 0);
    std::vector<int32_t> slots(n);
    if (cudaMemcpy(slots.data(), c->groupList.p + c->h_groupOff[group], (size_t)n * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    std::vector<std::pair<int32_t, int32_t>> pairs(n);
    for (int i = 0; i < n; i++) {
        size_t lo = 0, hi = s.meshes.size();
        while (hi - lo > 1) { const size_t mid = (lo + hi) / 2; if (s.meshes[mid].surfVBase <= slots[i]) lo = mid; else hi = mid; }
        const HostMesh& m = s.meshes[lo];
        pairs[i] = {m.id, m.surfaceVIds[slots[i] - m.surfVBase]};
    }
    std::sort(pairs.begin(), pairs.end());
    for (int i = 0; i < n && i < cap; i++) { out[2 * i] = pairs[i].first; out[2 * i + 1] = pairs[i].second; }
    return n;
}

// a15: face-side relations of one mesh (prepareCollisionDataCPU, M/VBD/VBDPhysics.cpp:1793-1806): rows
// (vertex, queryMesh, queryVertex, face-vertex order), sorted ascending. Returns the number of rows.
int64_t collision_fetch_relations(Scene& s, int mesh, int32_t* out, int64_t cap) {
    CollisionState* c = s.cs;
    if (!c || !c->prepared || c->nActive == 0) return 0;
    if (cudaStreamSynchronize(s.stream) != cudaSuccess) return -1;
    const HostMesh& m = s.meshes[mesh];
    const int64_t N = s.nVerts;
    std::vector<int32_t> off(N + 1);
    if (cudaMemcpy(off.data(), c->relOff.p, (size_t)(N + 1) * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    std::vector<int2> rel((size_t)off[N]);
    if (off[N] && cudaMemcpy(rel.data(), c->rel.p, rel.size() * sizeof(int2), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    struct Row { int32_t v, qm, qv, k; };
    std::vector<Row> rows;
    for (int v = 0; v < m.nV; v++) {
        const int sid = m.sortedId[v];
        for (int q = off[sid]; q < off[sid + 1]; q++) {
            const int slot = rel[q].x;
            size_t lo = 0, hi = s.meshes.size();
            while (hi - lo > 1) { const size_t mid = (lo + hi) / 2; if (s.meshes[mid].surfVBase <= slot) lo = mid; else hi = mid; }
            const HostMesh& qmesh = s.meshes[lo];
            rows.push_back({v, qmesh.id, qmesh.surfaceVIds[slot - qmesh.surfVBase], rel[q].y});
        }
    }
    std::sort(rows.begin(), rows.end(), [](const Row& a, const Row& b) {
        if (a.v != b.v) return a.v < b.v; if (a.qm != b.qm) return a.qm < b.qm; if (a.qv != b.qv) return a.qv < b.qv; return a.k < b.k; });
    const int64_t n = (int64_t)rows.size();
    if (out) for (int64_t i = 0; i < n && i < cap; i++) { out[4 * i] = rows[i].v; out[4 * i + 1] = rows[i].qm; out[4 * i + 2] = rows[i].qv; out[4 * i + 3] = rows[i].k; }
    return n;
}

int64_t collision_leaf_order(Scene& s, int which, int32_t* out, int64_t cap) {
    CollisionState* c = s.cs;
    if (!c) return 0;
    Lbvh* b = which == 0 ? &c->tetBvh : (which == 1 ? &c->surfBvh : (which == 2 ? &c->ccdBvh : nullptr));
    if (!b) return -1;
    if (!b->built) return 0;
    if (out) {
        if (cap < b->n) return -1;
        cudaStreamSynchronize(s.stream);
        if (cudaMemcpy(out, b->order.p, (size_t)b->n * 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    }
    return b->n;
}

}  // namespace gb
