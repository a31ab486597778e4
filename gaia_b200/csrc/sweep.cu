// Per-colour VBD vertex solve over TILES, hand-written for sm_100a: persistent blocks fed by the bulk-copy engine.
//
// Reference semantics (CPU/TBB path) - the same list as solve.cu:
//   accumlateMaterialForceAndHessian2 (Neo-Hookean)    M/VBD/VBD_NeoHookean.cpp:379-581
//   accumlateInertiaForceAndHessian                    M/VBD/VBD_BaseMaterial.h:359-366
//   accumlateBoundaryForceAndHessian + friction        M/VBD/VBDPhysics.cpp:2821-2954, 2563-2592
//   accumlateCollisionForceAndHessian[PerCollision]    M/VBD/VBDPhysics.cpp:2956-3047
//   VBDStepWithCollision + solve3x3_psd_stable         M/VBD/VBDPhysics.cpp:2402-2483, 3P/CuMatrix/.../CuMatrix.h:235-267
// The kernel this replaces in the reference: VBDSolveParallelGroup_vertexSweep_2hierarchiesGPU, M/VBD/VBDPhysicsCompute.cu:803-908
// (one block per vertex, one thread per incident tet, 9x9 Hessians in shared memory).
//
// Design. A colour is a list of tiles (host_setup.cpp: build_tile_streams): <= 64 consecutive vertices
// whose per-tet records (8 B: det B', three 8-bit neighbour slots) and per-neighbour entries (8 B: sorted id, summed
// weight of the linear term) are contiguous in HBM. A block is persistent over the tiles t = first + blockIdx.x + k*gridDim.x:
//   * thread 0 is the producer: cp.async.bulk (UBLKCP) copies of the tile's records, neighbour entries, vertex headers and
//     own positions into a 2-deep shared-memory ring, completion counted by an mbarrier (complete_tx) - the streams never
//     pass through registers or the LSU/L1 path, and the copy of tile k+2 is in flight while tile k is computed;
//   * L lanes serve one vertex. The scattered part - gathering the ~14 distinct neighbour positions - is software
//     pipelined one round ahead in registers (issued before the current round's arithmetic, consumed after it), staged
//     into shared memory relative to the owner, and feeds the linear term  -mu * sum_nb W_nb (x_nb - x)  on the way;
//   * the volumetric term walks the records: three ld.shared.v4 per tet, n = (b-a)x(c-a), det F = (a.n) det B',
//     f -= lmbd (det F - a) n/6 ... (SURVEY.md §8(a) "math restatement"), 10 sums butterfly-reduced over the L lanes;
//   * the vertex-level tail (damping, inertia, world box + friction, contacts, 3x3 solve, in-place store, halo push)
//     runs transposed: thread t finishes vertex t of the tile, at full SIMT width.
// Arithmetic per vertex depends only on L (never on the tiling), and every contraction is written with _rn intrinsics,
// so partitions of a mesh and the single-GPU run produce the same bits.
#include "scene_impl.h"
#include <algorithm>
#include <cstring>

namespace gb {

#define CK(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return _e; } while (0)

static const int kSweepThreads = 256;
static const int kSweepBlocksPerSM = 3;

// ---- shared memory map (bytes). Rings: records 2 deep, neighbour entries / vertex headers / own positions 3 deep.
static const int kRecStage = kTileRecCap * 8, kNbwStage = kTileNbCap * 8, kVtxStage = kTileV * 16;
static const int kSmRec = 0;                                          // [2][kTileRecCap] uint2
static const int kSmNbw = kSmRec + 2 * kRecStage;                      // [3][kTileNbCap] int2
static const int kSmHdr = kSmNbw + 3 * kNbwStage;                      // [3][kTileV] int4
static const int kSmPos = kSmHdr + 3 * kVtxStage;                      // [3][kTileV] float4
static const int kSmRel = kSmPos + 3 * kVtxStage;                      // [kTileNbCap] float4: neighbour positions relative to the owner
static const int kTailGroup = kSweepThreads / kTileV;                  // tiles whose vertex-level tails run together, one vertex per thread
static const int kSmAcc = kSmRel + kTileNbCap * 16;                    // [kTailGroup][9][kTileV] float: incident-tet sums
static const int kSmInfo = kSmAcc + kTailGroup * 9 * kTileV * 4;       // [4] TileDesc
static const int kSmTail = kSmInfo + 4 * 32;                           // [kTailGroup] int2 {vBegin, nV} of the tiles awaiting their tail
static const int kSmBarRec = kSmTail + kTailGroup * 8;                         // [2] mbarrier: records of tile k landed
static const int kSmBarNb = kSmBarRec + 2 * 8;                         // [3] mbarrier: neighbour entries + headers + positions landed
static const int kSmTotal = kSmBarNb + 3 * 8;
static_assert(3 * (kSmTotal + 1024) <= 227 * 1024, "three blocks per SM");

__device__ __forceinline__ float3 f3(float4 v) { return make_float3(v.x, v.y, v.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return make_float3(a.x - b.x, a.y - b.y, a.z - b.z); }

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// global -> shared bulk copy (TMA engine, 1-D): size is a multiple of 16 bytes, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_inval(uint32_t bar) { asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(bar) : "memory"); }
// coherent (L1-cached, invalidated by the acquire fence of the grid barrier) 16-byte load: the multi-colour launch reads
// positions other blocks wrote earlier in the SAME launch, which ld.global.nc must not be used for
__device__ __forceinline__ float4 ldca128(const float4* p) {
    float4 r;
    asm volatile("ld.global.ca.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
    return r;
}
// ghost positions are written by ANOTHER GPU while this kernel runs: a strong load at system scope (served by the point of
// coherence, never by a line some cache picked up before the neighbour's store landed)
__device__ __forceinline__ float4 ldsys128(const float4* p) {
    float4 r;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 r;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
    return r;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
    uint2 r;
    asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(addr));
    return r;
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t r;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(r) : "r"(addr));
    return r;
}
__device__ __forceinline__ int4 lds128i(uint32_t addr) {
    int4 r;
    asm volatile("ld.shared.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
    return r;
}
__device__ __forceinline__ void sts128i(uint32_t addr, int4 v) {
    asm volatile("st.shared.v4.s32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void sts32f(uint32_t addr, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory"); }
__device__ __forceinline__ float lds32f(uint32_t addr) { return __uint_as_float(lds32(addr)); }
// predicated 16-byte store (no branch around it)
__device__ __forceinline__ void sts128_if(bool pred, uint32_t addr, float x, float y, float z, float w) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %5, 0;\n\t@p st.shared.v4.f32 [%0], {%1,%2,%3,%4};\n\t}"
                 ::"r"(addr), "f"(x), "f"(y), "f"(z), "f"(w), "r"((uint32_t)pred) : "memory");
}
__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

struct SweepArgs {
    int tileBegin, tileEnd;              // tiles of this colour
    float4* pos;
    const float4* prev;
    const float4* inertia;
    const int4* vtx2;
    const uint2* rec;
    const int2* nbw;
    const TileDesc* tiles;
    const MeshParamsDev* mp;
    StepParamsDev sp;
    int applyFriction;
    int needPrev;                        // some mesh has damping, or the world box has friction: the tail reads x_prev
    int allActive;                       // every mesh is active for the material solve (no per-vertex look-up needed)
    ContactView cv;
    int32_t* errFlags;
    HaloView hv;
    // MULTI: one launch sweeps nSteps colours in turn (colour of step j = (firstGroup + j) % nGroups), a grid-wide barrier
    // between them
    const int2* stepTiles;               // [nGroups] {tileBegin, tileEnd}
    int nSteps, firstGroup, nGroups;
    int frictionFromStep;                // steps >= this apply friction
    uint32_t* gridBar;                   // [0] arrivals, [1] exits (both 0 between launches)
};

// IPC-style mollified friction (computeVertexFriction, VBDPhysics.cpp:2563-2592) for an axis-aligned tangent basis
__device__ __forceinline__ void box_friction(float mu, float lambda, float ua, float ub, float epsU, float& fa, float& fb, float& haa) {
    const float uNorm = sqrtf(ua * ua + ub * ub);
    fa = 0.f; fb = 0.f; haa = 0.f;
    if (uNorm > 0.f) {
        const float f1 = (uNorm > epsU) ? 1.f / uNorm : (-uNorm / epsU + 2.f) / epsU;
        const float s = -mu * lambda * f1;
        fa = s * ua; fb = s * ub;
        haa = mu * lambda * f1;
    }
}

// The vertex-level part of VBDStepWithCollision once the incident-tet sums are known. (fx,fy,fz) = -f_elastic,
// s1 = A sum|m0|^2, g = sum A g g^T (already / 6). Returns true when a new position was written.
// LS: do not move the vertex; hand the descent direction back instead (xNew = {dx, +-step}: negative step = the 3x3 solve
// failed and dx is the force) - the line-search variant decides the step length itself.
template <bool HAS_CONTACTS, bool LS = false>
__device__ __forceinline__ bool vertex_tail(int s, float3 x, float4 ys, float3 xp, const MeshParamsDev* __restrict__ P,
                                            float fx, float fy, float fz, float s1, float g00, float g01, float g02, float g11, float g12, float g22,
                                            float4* pos, const StepParamsDev& sp, int applyFriction, const ContactView& cv, int32_t* errFlags, float4& xNew) {
    const float miu = P->miu, lmbd = P->lmbd, dampH = P->dampH, dampD = P->dampD;
    float F0 = -fx, F1 = -fy, F2 = -fz;
    const float hd = miu * s1;
    float h00 = lmbd * g00 + hd, h01 = lmbd * g01, h02 = lmbd * g02, h11 = lmbd * g11 + hd, h12 = lmbd * g12, h22 = lmbd * g22 + hd;
    if ((dampH != 0.f) | (dampD != 0.f)) {
        // dampingH = (H_vol*gamma_h + mu*|m|^2*A*gamma_d I)/dt ; f -= dampingH*(x - x_prev); H += dampingH
        const float idt = 1.f / sp.dt;
        const float3 d = x - xp;
        const float ch = lmbd * dampH * idt, cd = miu * s1 * dampD * idt;
        const float d00 = ch * g00 + cd, d01 = ch * g01, d02 = ch * g02, d11 = ch * g11 + cd, d12 = ch * g12, d22 = ch * g22 + cd;
        F0 -= d00 * d.x + d01 * d.y + d02 * d.z;
        F1 -= d01 * d.x + d11 * d.y + d12 * d.z;
        F2 -= d02 * d.x + d12 * d.y + d22 * d.z;
        h00 += d00; h01 += d01; h02 += d02; h11 += d11; h12 += d12; h22 += d22;
    }
    {   // inertia
        const float m = ys.w;
        const float w = m * sp.dtSqrReciprocal;
        F0 += (m * (ys.x - x.x)) * sp.dtSqrReciprocal;
        F1 += (m * (ys.y - x.y)) * sp.dtSqrReciprocal;
        F2 += (m * (ys.z - x.z)) * sp.dtSqrReciprocal;
        h00 += w; h11 += w; h22 += w;
    }
    if (HAS_CONTACTS && sp.useBowlGround) {
        // the inside of a sphere as the ground (accumlateBoundaryForceAndHessian, VBDPhysics.cpp:2824-2864), as written there -
        // including the sign of its Hessian. Compiled into the contact variant only (scenes with the bowl are routed to it with
        // empty contact tables when nothing touches), so the contact-free sweep is the same code as without this feature.
        const float3 rel = make_float3(x.x - sp.bowlCenter[0], x.y - sp.bowlCenter[1], x.z - sp.bowlCenter[2]);
        const float dist = sqrtf(rel.x * rel.x + rel.y * rel.y + rel.z * rel.z);
        if ((sp.bowlCap || !(rel.y > 0.f)) && !(dist < sp.bowlRadius)) {
            const float3 dir = make_float3(rel.x / dist, rel.y / dist, rel.z / dist);
            const float depth = dist - sp.bowlRadius;
            const float fk = -depth * sp.kBoundary;
            F0 += fk * dir.x; F1 += fk * dir.y; F2 += fk * dir.z;
            const float rd = sp.bowlRadius / dist, kB = sp.kBoundary;
            h00 += kB * ((-1.f + dir.x * dir.x) + rd * (1.f - dir.x * dir.x));
            h11 += kB * ((-1.f + dir.y * dir.y) + rd * (1.f - dir.y * dir.y));
            h22 += kB * ((-1.f + dir.z * dir.z) + rd * (1.f - dir.z * dir.z));
            h01 += kB * (dir.x * dir.y + rd * (0.f - dir.x * dir.y));
            h02 += kB * (dir.x * dir.z + rd * (0.f - dir.x * dir.z));
            h12 += kB * (dir.y * dir.z + rd * (0.f - dir.y * dir.z));
            if (applyFriction) {
                // T = [normalize(dir x e_z), normalize(dir x T0)], u = T^T (x - x_prev); computeVertexFriction (:2563-2592)
                float3 t0 = make_float3(dir.y, -dir.x, 0.f);
                const float n0 = sqrtf(t0.x * t0.x + t0.y * t0.y + t0.z * t0.z);
                t0 = make_float3(t0.x / n0, t0.y / n0, t0.z / n0);
                float3 t1 = make_float3(dir.y * t0.z - dir.z * t0.y, dir.z * t0.x - dir.x * t0.z, dir.x * t0.y - dir.y * t0.x);
                const float n1 = sqrtf(t1.x * t1.x + t1.y * t1.y + t1.z * t1.z);
                t1 = make_float3(t1.x / n1, t1.y / n1, t1.z / n1);
                const float3 dx = make_float3(x.x - xp.x, x.y - xp.y, x.z - xp.z);
                const float u0 = t0.x * dx.x + t0.y * dx.y + t0.z * dx.z, u1 = t1.x * dx.x + t1.y * dx.y + t1.z * dx.z;
                float fa, fb, haa;
                box_friction(sp.muBoundary, depth * sp.kBoundary, u0, u1, sp.epsVBoundary * sp.dt, fa, fb, haa);
                F0 += fa * t0.x + fb * t1.x; F1 += fa * t0.y + fb * t1.y; F2 += fa * t0.z + fb * t1.z;
                h00 += haa * (t0.x * t0.x + t1.x * t1.x); h11 += haa * (t0.y * t0.y + t1.y * t1.y); h22 += haa * (t0.z * t0.z + t1.z * t1.z);
                h01 += haa * (t0.x * t0.y + t1.x * t1.y); h02 += haa * (t0.x * t0.z + t1.x * t1.z); h12 += haa * (t0.y * t0.z + t1.y * t1.z);
            }
        }
    } else if (sp.usePlaneGround) {   // world-box penalty + friction, per violated axis (CPU semantics)
        const float xs[3] = {x.x, x.y, x.z};
        float Fb[3] = {0.f, 0.f, 0.f}, Hb[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int d = 0; d < 3; d++) {
            float depth = 0.f, sgn = 0.f;
            if (xs[d] < sp.boundsMin[d]) { depth = sp.boundsMin[d] - xs[d]; sgn = 1.f; }
            else if (xs[d] > sp.boundsMax[d]) { depth = xs[d] - sp.boundsMax[d]; sgn = -1.f; }
            if (sgn != 0.f) {
                Fb[d] += sgn * (depth * sp.kBoundary);
                Hb[d] += sp.kBoundary;
                if (applyFriction) {
                    const float dx[3] = {x.x - xp.x, x.y - xp.y, x.z - xp.z};
                    const int a = (d + 1) % 3, b2 = (d + 2) % 3;
                    float fa, fb, haa;
                    box_friction(sp.muBoundary, depth * sp.kBoundary, dx[a], dx[b2], sp.epsVBoundary * sp.dt, fa, fb, haa);
                    Fb[a] += fa; Fb[b2] += fb; Hb[a] += haa; Hb[b2] += haa;
                }
            }
        }
        F0 += Fb[0]; F1 += Fb[1]; F2 += Fb[2];
        h00 += Hb[0]; h11 += Hb[1]; h22 += Hb[2];
    }
    if (HAS_CONTACTS) {   // own contact with weight 1, contacts on faces this vertex belongs to with weight -bary
        const int own = __ldg(cv.contactOfVertex + s);
        const int rb = __ldg(cv.relOff + s), re = __ldg(cv.relOff + s + 1);
        for (int q = (own >= 0 ? -1 : 0); q < re - rb; q++) {
            int slot; float bw;
            if (q < 0) { slot = own; bw = 1.f; }
            else {
                const int2 r = __ldg(cv.rel + rb + q);
                slot = r.x;
                const float4 h1 = __ldg(cv.cH1 + slot);
                bw = -(r.y == 0 ? h1.z : (r.y == 1 ? h1.w : __ldg(cv.cB2 + slot)));
            }
            const float4 cf = __ldg(cv.cF + slot);
            if (cf.w > 0.f && bw != 0.f) {
                const float4 c0 = __ldg(cv.cH0 + slot), c1 = __ldg(cv.cH1 + slot);
                const float b2w = bw * bw;
                F0 += cf.x * bw; F1 += cf.y * bw; F2 += cf.z * bw;
                h00 += c0.x * b2w; h01 += c0.y * b2w; h02 += c0.z * b2w; h11 += c0.w * b2w; h12 += c1.x * b2w; h22 += c1.y * b2w;
            }
        }
    }
    // 3x3 solve (solve3x3_psd_stable) and in-place update
    if (F0 * F0 + F1 * F1 + F2 * F2 > 1e-10f) {
        const float i11 = h22 * h11 - h12 * h12;
        const float i12 = -(h22 * h01 - h12 * h02);
        const float i13 = h12 * h01 - h11 * h02;
        const float det = h00 * i11 + h01 * i12 + h02 * i13;
        float dx0, dx1, dx2, step;
        bool failed = false;
        if (fabsf(det) < 1e-5f * (fabsf(h00 * i11) + fabsf(h01 * i12) + fabsf(h02 * i13))) {
            dx0 = F0; dx1 = F1; dx2 = F2; step = sp.stepSizeGD; failed = true;
            atomicAdd(errFlags + 1, 1);
        } else {
            const float deti = 1.f / det;
            const float i22 = h22 * h00 - h02 * h02;
            const float i23 = -(h12 * h00 - h01 * h02);
            const float i33 = h11 * h00 - h01 * h01;
            dx0 = deti * (i11 * F0 + i12 * F1 + i13 * F2);
            dx1 = deti * (i12 * F0 + i22 * F1 + i23 * F2);
            dx2 = deti * (i13 * F0 + i23 * F1 + i33 * F2);
            step = sp.stepSize;
        }
        if (isnan(dx0) || isnan(dx1) || isnan(dx2)) { atomicAdd(errFlags + 0, 1); return false; }
        if (LS) { xNew = make_float4(dx0, dx1, dx2, failed ? -step : step); return true; }
        xNew = make_float4(x.x + step * dx0, x.y + step * dx1, x.z + step * dx2, 0.f);
        pos[s] = xNew;
        return true;
    }
    return false;
}

// Fused halo exchange, consumer side: wait until every neighbour partition has published `epoch` (its previous colour
// launch, pushes included, is complete). One lane polls; 5 s without progress = a neighbour died (errFlags[4]).
__device__ __forceinline__ void sweep_halo_wait(const HaloPeerTable* tab, uint32_t epoch, int nPeers, int32_t* errFlags) {
    for (int p = 0; p < nPeers; p++) {
        if (*(volatile int32_t*)(errFlags + 4)) return;
        const unsigned long long t0 = gtimer();
        for (;;) {
            uint32_t v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(tab->myFlag[p]) : "memory");
            if ((int32_t)(v - epoch) >= 0) break;
            if (gtimer() - t0 > 5000000000ull) { atomicAdd(errFlags + 4, 1); return; }
            __nanosleep(64);
        }
    }
}

// what a thread holds of the NEXT round while the current one is computed: its share of the vertex's neighbour gathers
template <int NQ>
struct Prefetch {
    float4 p[NQ];      // gathered neighbour positions (absolute)
    float w[NQ];       // their linear-term weights
    int hdrY;          // nbLocal | nNb << 16 of the vertex (0 when the lane has no vertex)
};

template <int L, bool HAS_CONTACTS, bool HALO, bool MULTI>
__global__ void __launch_bounds__(kSweepThreads, kSweepBlocksPerSM)
k_sweep_tiles(const SweepArgs a) {
    static_assert(!(MULTI && HAS_CONTACTS), "the contact pre-pass of a colour is a separate launch: one colour per launch");
    constexpr int VPR = kSweepThreads / L;          // vertices per round
    constexpr int NQ = 16 / L;                      // gathers held in registers per lane (16 neighbours per vertex)
    extern __shared__ __align__(128) unsigned char smem[];
    const uint32_t sb = smem_u32(smem);             // every shared-memory access below goes through 32-bit shared addresses
    const int tid = threadIdx.x, lane = tid & (L - 1), sub = tid / L;
    int tileBegin = a.tileBegin, tileEnd = a.tileEnd, applyFriction = a.applyFriction;
    if (MULTI) {
        const int2 t = __ldg(a.stepTiles + a.firstGroup);
        tileBegin = t.x; tileEnd = t.y; applyFriction = a.frictionFromStep <= 0;
    }
    int nMine = (tileEnd - tileBegin - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // tiles of this block
    if (nMine < 0) nMine = 0;
    int step = 0;                     // MULTI: colour step inside this launch
    const float4* posRO = a.pos;      // gathers only touch other colours: read-only for this colour

    if (HALO && blockIdx.x == 0 && tid < a.hv.nPeers) {
        // everything this stream launched before is complete (stream order): publish it to the neighbours
        const uint32_t done = *(volatile const uint32_t*)a.hv.ctl + (uint32_t)a.hv.epochOffset;
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.hv.tab->peerFlag[tid]), "r"(done) : "memory");
    }
    if (!MULTI && nMine <= 0) return;
    if (tid == 0) {
        for (int i = 0; i < 5; i++) mbar_init(sb + kSmBarRec + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // ---- producer (thread 0). Tile k: neighbour entries, headers and own positions -> slot k % 3 (one barrier: the data the
    // gathers need first); records -> slot k & 1 (their own barrier, needed half a tile later). info ring: slot k & 3.
    auto issue_nb = [&](int k) {
        const int4* dp = (const int4*)(a.tiles + tileBegin + (int)blockIdx.x + k * (int)gridDim.x);
        const int4 d0 = __ldg(dp), d1 = __ldg(dp + 1);       // {vBegin, nV, recBegin, recCount}, {nbBegin, nbCount, flags, -}
        const int s3 = k % 3;
        sts128i(sb + kSmInfo + (k & 3) * 32, d0); sts128i(sb + kSmInfo + (k & 3) * 32 + 16, d1);
        const uint32_t bar = sb + kSmBarNb + 8 * s3;
        const uint32_t bNb = (uint32_t)d1.y * 8u, bV = (uint32_t)d0.y * 16u;
        mbar_expect_tx(bar, bNb + 2 * bV);
        if (bNb) bulk_g2s(sb + kSmNbw + s3 * kNbwStage, a.nbw + d1.x, bNb, bar);
        bulk_g2s(sb + kSmHdr + s3 * kVtxStage, a.vtx2 + d0.x, bV, bar);
        bulk_g2s(sb + kSmPos + s3 * kVtxStage, a.pos + d0.x, bV, bar);
    };
    auto issue_rec = [&](int k) {
        const int4 d0 = __ldg((const int4*)(a.tiles + tileBegin + (int)blockIdx.x + k * (int)gridDim.x));
        const uint32_t bar = sb + kSmBarRec + 8 * (k & 1);
        const uint32_t bRec = (uint32_t)d0.w * 8u;
        mbar_expect_tx(bar, bRec);
        if (bRec) bulk_g2s(sb + kSmRec + (k & 1) * kRecStage, a.rec + d0.z, bRec, bar);
    };
    auto prologue = [&]() {
        if (nMine > 0) { issue_nb(0); issue_rec(0); }
        if (nMine > 1) { issue_nb(1); issue_rec(1); }
        if (nMine > 2) issue_nb(2);
    };
    if (tid == 0) prologue();
    // (returns false when the copies never arrived - a watchdog for a sick device: the block leaves instead of spinning)
    auto wait_bar = [&](uint32_t bar, uint32_t parity) -> bool {
        if (mbar_try_wait(bar, parity)) return true;
        const unsigned long long t0 = gtimer();
        while (!mbar_try_wait(bar, parity))
            if (gtimer() - t0 > 4000000000ull) { if (tid == 0) atomicAdd(a.errFlags + 5, 1); return false; }
        return true;
    };

    bool haloWaited = false;
    // ---- gathers of round r of the tile in ring slots (s3, s4): wait for the tile's neighbour entries (first round), then
    // request this lane's positions
    auto prefetch = [&](int s3, int s4, uint32_t parNb, int r, Prefetch<NQ>& pf) -> bool {
        if (r == 0 && !wait_bar(sb + kSmBarNb + 8 * s3, parNb)) return false;
        bool ghost = false;
        if (HALO) {
            // a tile that holds boundary vertices also holds vertices whose neighbours are ghosts: wait for the neighbour
            // partitions' previous colour before gathering, and gather past L1
            ghost = (lds32(sb + kSmInfo + s4 * 32 + 24) & 1u) != 0;
            if (ghost && !haloWaited) {
                if (tid == 0 && !(a.hv.debug & 1)) sweep_halo_wait(a.hv.tab, *(volatile const uint32_t*)a.hv.ctl + (uint32_t)(a.hv.epochOffset + step), a.hv.nPeers, a.errFlags);
                __syncthreads();
                haloWaited = true;
            }
        }
        const int v = r * VPR + sub;
        pf.hdrY = 0;
        if (v < (int)lds32(sb + kSmInfo + s4 * 32 + 4)) {
            const int hw = (int)lds32(sb + kSmHdr + s3 * kVtxStage + v * 16 + 12);
            if (hw >= 0 && (a.allActive || a.mp[hw & 0x00FFFFFF].activeSolve)) pf.hdrY = (int)lds32(sb + kSmHdr + s3 * kVtxStage + v * 16 + 4);
        }
        const int nNb = pf.hdrY >> 16;
        const uint32_t nbAddr = sb + kSmNbw + s3 * kNbwStage + ((pf.hdrY & 0xFFFF) + lane) * 8;
#pragma unroll
        for (int q = 0; q < NQ; q++) {
            pf.p[q] = make_float4(0.f, 0.f, 0.f, 0.f);
            pf.w[q] = 0.f;
            if (lane + q * L < nNb) {
                const uint2 e = lds64(nbAddr + q * (L * 8));
                pf.w[q] = __uint_as_float(e.y);
                pf.p[q] = (HALO && ghost) ? __ldcg(posRO + e.x) : (MULTI ? ldca128(posRO + e.x) : __ldg(posRO + e.x));
            }
        }
        return true;
    };

    Prefetch<NQ> pf;
#pragma unroll 1
    for (;;) {      // (one pass unless MULTI)
    if (nMine > 0 && !prefetch(0, 0, 0u, 0, pf)) return;

    int s3 = 0;                 // k % 3
    uint32_t parNb = 0;         // (k / 3) & 1
#pragma unroll 1
    for (int k = 0; k < nMine; k++) {
        const int st = k & 1, s4 = k & 3;
        const int s3n = (s3 == 2) ? 0 : s3 + 1;
        const uint32_t parNbN = (s3 == 2) ? parNb ^ 1u : parNb;
        const int vBegin = (int)lds32(sb + kSmInfo + s4 * 32), nV = (int)lds32(sb + kSmInfo + s4 * 32 + 4);
        const bool ghostTile = HALO && (lds32(sb + kSmInfo + s4 * 32 + 24) & 1u);
        const int rounds = (nV + VPR - 1) / VPR;
        const int tg = k % kTailGroup;                 // position of this tile in its tail group
        const uint32_t accAddr = sb + kSmAcc + tg * (9 * kTileV * 4);
        // the tail's per-vertex inputs are read from global memory: pulled into L1 now
        if (tid < nV) {
            asm volatile("prefetch.global.L1 [%0];" ::"l"(a.inertia + vBegin + tid));
            if (a.needPrev) asm volatile("prefetch.global.L1 [%0];" ::"l"(a.prev + vBegin + tid));
            asm volatile("prefetch.global.L1 [%0];" ::"l"(a.vtx2 + vBegin + tid));
            asm volatile("prefetch.global.L1 [%0];" ::"l"(a.pos + vBegin + tid));
        }
        if (tid == 0) asm volatile("st.shared.v2.s32 [%0], {%1,%2};" ::"r"(sb + kSmTail + tg * 8), "r"(vBegin), "r"(nV) : "memory");

#pragma unroll 1
        for (int r = 0; r < rounds; r++) {
            // ---- current round: stage the gathered neighbours relative to the owner, linear term on the way
            const int v = r * VPR + sub;
            const int nbLocal = pf.hdrY & 0xFFFF, nNb = pf.hdrY >> 16;
            const bool active = nNb > 0;       // (a solved vertex of a tet mesh always has neighbours)
            int4 h = make_int4(0, 0, 0, 0);
            float3 x = make_float3(0.f, 0.f, 0.f);
            if (active) {
                h = lds128i(sb + kSmHdr + s3 * kVtxStage + v * 16);
                x = f3(lds128(sb + kSmPos + s3 * kVtxStage + v * 16));
            }
            const uint32_t relAddr = sb + kSmRel + nbLocal * 16;
            float lx = 0.f, ly = 0.f, lz = 0.f;
#pragma unroll
            for (int q = 0; q < NQ; q++) {
                // (lanes without a neighbour hold p = 0, w = 0: the arithmetic runs unconditionally, only the store is predicated.
                //  The loaded .w is stored too: all four registers of the gather stay live until here, so the compiler cannot
                //  hand one of them to another value and stall on the in-flight load)
                const float rx = __fsub_rn(pf.p[q].x, x.x), ry = __fsub_rn(pf.p[q].y, x.y), rz = __fsub_rn(pf.p[q].z, x.z);
                sts128_if(lane + q * L < nNb, relAddr + (lane + q * L) * 16, rx, ry, rz, pf.p[q].w);
                lx = __fmaf_rn(pf.w[q], rx, lx); ly = __fmaf_rn(pf.w[q], ry, ly); lz = __fmaf_rn(pf.w[q], rz, lz);
            }
            if (nNb > NQ * L) {      // vertices with more than 16 distinct neighbours: the rest is gathered in place
                const uint32_t nbAddr = sb + kSmNbw + s3 * kNbwStage + nbLocal * 8;
                for (int i = lane + NQ * L; i < nNb; i += L) {
                    const uint2 e = lds64(nbAddr + i * 8);
                    const float4 p = ghostTile ? __ldcg(posRO + e.x) : (MULTI ? ldca128(posRO + e.x) : __ldg(posRO + e.x));
                    const float w = __uint_as_float(e.y);
                    const float rx = __fsub_rn(p.x, x.x), ry = __fsub_rn(p.y, x.y), rz = __fsub_rn(p.z, x.z);
                    sts128_if(true, relAddr + i * 16, rx, ry, rz, p.w);
                    lx = __fmaf_rn(w, rx, lx); ly = __fmaf_rn(w, ry, ly); lz = __fmaf_rn(w, rz, lz);
                }
            }
            // ---- next round's gathers go out now; they land while this round's records are processed
            if (r + 1 < rounds) prefetch(s3, s4, parNb, r + 1, pf);
            else if (k + 1 < nMine) { if (!prefetch(s3n, (k + 1) & 3, parNbN, 0, pf)) return; }
            if (r == 0 && !wait_bar(sb + kSmBarRec + 8 * st, (uint32_t)(k >> 1) & 1u)) return;
            __syncwarp();

            // ---- volumetric term over the incident tets
            float miu = 0.f, lmbd = 0.f, aConst = 0.f;
            if (active) { const MeshParamsDev* P = a.mp + (h.w & 0x00FFFFFF); miu = P->miu; lmbd = P->lmbd; aConst = P->a; }
            float fx = 0.f, fy = 0.f, fz = 0.f, g00 = 0.f, g01 = 0.f, g02 = 0.f, g11 = 0.f, g12 = 0.f, g22 = 0.f;
            const int nRec = active ? (h.x >> 16) : 0;
            uint32_t recAddr = sb + kSmRec + st * kRecStage + ((h.x & 0xFFFF) + lane) * 8;      // this lane's next record
            // The lane's records follow triangle strips over the vertex's link (host_setup.cpp: plan_strips): step t keeps two
            // of the three neighbour positions in registers and replaces register t % 3 by ONE shared-memory read; a restart
            // record (bit 30) reloads all three. Bit 31: (p0,p1,p2) is an odd permutation of the tet's (a,b,c): n flips.
            const int cnt = (nRec - lane + L - 1) / L;
            float3 p0 = make_float3(0.f, 0.f, 0.f), p1 = p0, p2 = p0;
            uint2 R = make_uint2(0u, 0u);
            if (cnt > 0) R = lds64(recAddr);
#define GAIA_STRIP_STEP(PN, PB, PC, T)                                                                                        \
            {                                                                                                                 \
                uint2 Rn = R;                                                                                                 \
                recAddr += L * 8;                                                                                             \
                if ((T) + 1 < cnt) Rn = lds64(recAddr);                   /* requested before this one is consumed */         \
                const uint32_t I = R.y;                                                                                       \
                PN = f3(lds128(relAddr + ((I & 0xFFu) << 4)));                                                                \
                if (I & 0x40000000u) {                                                                                        \
                    PB = f3(lds128(relAddr + ((I >> 4) & 0xFF0u)));                                                           \
                    PC = f3(lds128(relAddr + ((I >> 12) & 0xFF0u)));                                                          \
                }                                                                                                             \
                /* n = (p1-p0) x (p2-p0) = +-(b-a) x (c-a); det(Ds') = a.n = p0.n; det F = det(Ds') det(B'); 6 A g g^T = det(B') n n^T */ \
                const float3 u = p1 - p0, w = p2 - p0;                                                                        \
                float3 n;                                                                                                     \
                n.x = __fmaf_rn(u.y, w.z, -__fmul_rn(u.z, w.y));                                                              \
                n.y = __fmaf_rn(u.z, w.x, -__fmul_rn(u.x, w.z));                                                              \
                n.z = __fmaf_rn(u.x, w.y, -__fmul_rn(u.y, w.x));                                                              \
                const float detDs = __fmaf_rn(p0.z, n.z, __fmaf_rn(p0.y, n.y, __fmul_rn(p0.x, n.x)));                         \
                const uint32_t flip = I & 0x80000000u;                                                                        \
                const float dB = __uint_as_float(R.x);                                                                        \
                const float kk = __fmaf_rn(detDs, __uint_as_float(R.x ^ flip), -aConst);        /* det F - a */              \
                const float ks = __uint_as_float(__float_as_uint(kk) ^ flip);                   /* (det F - a) * sign of n */ \
                fx = __fmaf_rn(ks, n.x, fx); fy = __fmaf_rn(ks, n.y, fy); fz = __fmaf_rn(ks, n.z, fz);                        \
                const float nx = __fmul_rn(dB, n.x), ny = __fmul_rn(dB, n.y), nz = __fmul_rn(dB, n.z);                        \
                g00 = __fmaf_rn(nx, n.x, g00); g01 = __fmaf_rn(nx, n.y, g01); g02 = __fmaf_rn(nx, n.z, g02);                  \
                g11 = __fmaf_rn(ny, n.y, g11); g12 = __fmaf_rn(ny, n.z, g12); g22 = __fmaf_rn(nz, n.z, g22);                  \
                R = Rn;                                                                                                       \
            }
#pragma unroll 1
            for (int t = 0; t < cnt; t += 3) {
                GAIA_STRIP_STEP(p0, p1, p2, t)
                if (t + 1 < cnt) GAIA_STRIP_STEP(p1, p2, p0, t + 1)
                if (t + 2 < cnt) GAIA_STRIP_STEP(p2, p0, p1, t + 2)
            }
#undef GAIA_STRIP_STEP
            // -f = mu * sum_nb W (x_nb - x) + lmbd * sum_tets (det F - a) * (-n / 6)
            const float lm6 = lmbd * (-1.f / 6.f);
            fx = __fmaf_rn(lm6, fx, __fmul_rn(miu, lx)); fy = __fmaf_rn(lm6, fy, __fmul_rn(miu, ly)); fz = __fmaf_rn(lm6, fz, __fmul_rn(miu, lz));
#pragma unroll
            for (int o = L / 2; o > 0; o >>= 1) {
                fx += __shfl_xor_sync(0xffffffffu, fx, o); fy += __shfl_xor_sync(0xffffffffu, fy, o); fz += __shfl_xor_sync(0xffffffffu, fz, o);
                g00 += __shfl_xor_sync(0xffffffffu, g00, o); g01 += __shfl_xor_sync(0xffffffffu, g01, o); g02 += __shfl_xor_sync(0xffffffffu, g02, o);
                g11 += __shfl_xor_sync(0xffffffffu, g11, o); g12 += __shfl_xor_sync(0xffffffffu, g12, o); g22 += __shfl_xor_sync(0xffffffffu, g22, o);
            }
            if (lane == 0 && v < nV) {
                const uint32_t aa = accAddr + v * 4;
                sts32f(aa + 0 * kTileV * 4, fx); sts32f(aa + 1 * kTileV * 4, fy); sts32f(aa + 2 * kTileV * 4, fz);
                sts32f(aa + 3 * kTileV * 4, g00); sts32f(aa + 4 * kTileV * 4, g01); sts32f(aa + 5 * kTileV * 4, g02);
                sts32f(aa + 6 * kTileV * 4, g11); sts32f(aa + 7 * kTileV * 4, g12); sts32f(aa + 8 * kTileV * 4, g22);
            }
        }
        __syncthreads();      // sums complete; this tile's record / neighbour slots and the staging rows are free again
        if (tid == 0) {
            if (k + 2 < nMine) issue_rec(k + 2);
            if (k + 3 < nMine) issue_nb(k + 3);
        }

        // ---- transposed tails of the last kTailGroup tiles, one vertex per thread (full SIMT width)
        if (tg == kTailGroup - 1 || k == nMine - 1) {
            const int tt = tid / kTileV, vt = tid % kTileV;
            int tvx, tvy;
            asm volatile("ld.shared.v2.s32 {%0,%1}, [%2];" : "=r"(tvx), "=r"(tvy) : "r"(sb + kSmTail + tt * 8));
            const uint32_t ac = sb + kSmAcc + tt * (9 * kTileV * 4) + vt * 4;
            bool pushed = false;
            if (tt <= tg && vt < tvy) {
                const int s = tvx + vt;
                const int4 h = __ldg(a.vtx2 + s);
                const MeshParamsDev* P = a.mp + (h.w & 0x00FFFFFF);
                if (h.w >= 0 && P->activeSolve) {
                    const float3 x = f3(a.pos[s]);
                    const float4 ys = __ldg(a.inertia + s);
                    float4 xpv = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (a.needPrev) xpv = __ldg(a.prev + s);
                    float4 xNew;
                    const bool wrote = vertex_tail<HAS_CONTACTS>(s, x, ys, f3(xpv), P, lds32f(ac + 0 * kTileV * 4), lds32f(ac + 1 * kTileV * 4), lds32f(ac + 2 * kTileV * 4),
                                           __int_as_float(h.z), lds32f(ac + 3 * kTileV * 4) * (1.f / 6.f), lds32f(ac + 4 * kTileV * 4) * (1.f / 6.f),
                                           lds32f(ac + 5 * kTileV * 4) * (1.f / 6.f), lds32f(ac + 6 * kTileV * 4) * (1.f / 6.f), lds32f(ac + 7 * kTileV * 4) * (1.f / 6.f),
                                           lds32f(ac + 8 * kTileV * 4) * (1.f / 6.f), a.pos, a.sp, applyFriction, a.cv, a.errFlags, xNew);
                    if (HALO && wrote && (h.w & kHaloBoundaryBit) && !(a.hv.debug & 2)) {
                        const int si = __ldg(a.hv.sendIdx + s);
                        for (int q = 0; q < (si & 15); q++) {
                            const int2 en = __ldg(a.hv.sendEnt + (si >> 4) + q);
                            a.hv.tab->peerPos[en.x][en.y] = xNew;
                        }
                        pushed = true;
                    }
                }
            }
            // the next launch of the stream publishes the epoch; the stores must be performed system-wide before it does
            if (HALO && pushed) __threadfence_system();
            __syncthreads();      // the sums are consumed: the next group may overwrite them
        }
        s3 = s3n; parNb = parNbN;
    }
    if (!MULTI) break;
    if (++step == a.nSteps) break;
    // ---- end of a colour: every block arrives at the grid barrier, re-arms its pipeline for the next colour and starts that
    // colour's stream copies (records, neighbour entries, headers and OWN positions do not depend on the colour just swept)
    // before it waits for the other blocks; only the neighbour gathers need the barrier
    {
        const int2 t = __ldg(a.stepTiles + (a.firstGroup + step) % a.nGroups);
        tileBegin = t.x; tileEnd = t.y; applyFriction = step >= a.frictionFromStep;
        nMine = (tileEnd - tileBegin - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
        if (nMine < 0) nMine = 0;
    }
    __syncthreads();      // (all tails of this block are done, nobody waits on an mbarrier any more)
    __shared__ int dead;
    if (tid == 0) {
        dead = 0;
        __threadfence();
        asm volatile("fence.proxy.async;" ::: "memory");      // the positions written here are read by other blocks' bulk copies
        const uint32_t arrived = atomicAdd(a.gridBar, 1u) + 1u;
        const uint32_t target = (uint32_t)step * gridDim.x;
        if (HALO && arrived == target) {      // last block of the colour: publish it to the neighbour partitions
            const uint32_t done = *(volatile const uint32_t*)a.hv.ctl + (uint32_t)(a.hv.epochOffset + step);
            __threadfence_system();
            for (int p = 0; p < a.hv.nPeers; p++)
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.hv.tab->peerFlag[p]), "r"(done) : "memory");
        }
        for (int i = 0; i < 5; i++) { mbar_inval(sb + kSmBarRec + 8 * i); mbar_init(sb + kSmBarRec + 8 * i, 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        prologue();
        if (arrived != target) {
            const unsigned long long t0 = gtimer();
            uint32_t spins = 0;
            while ((int32_t)(ld_acquire_gpu(a.gridBar) - target) < 0) {
                if ((++spins & 1023u) == 0 && gtimer() - t0 > 4000000000ull) { atomicAdd(a.errFlags + 5, 1); dead = 1; break; }
            }
        }
        __threadfence();      // acquire: the other blocks' positions are visible, stale L1 lines are dropped
        asm volatile("fence.proxy.async;" ::: "memory");
    }
    __syncthreads();
    if (dead) return;
    haloWaited = false;
    }
    if (MULTI && tid == 0) {      // last block out re-arms the barrier for the next launch
        __syncwarp();
        if (atomicAdd(a.gridBar + 1, 1u) == gridDim.x - 1) { a.gridBar[0] = 0u; a.gridBar[1] = 0u; __threadfence(); }
    }
}

// ---------------------------------------------------------------------------------------------------------------- host
cudaError_t sweep_upload(Scene& s) {
    CK(s.vtx2.alloc(s.h_vtx2.size())); CK(s.rec.alloc(s.h_rec.size() + 2)); CK(s.nbw.alloc(s.h_nbw.size() + 2)); CK(s.tiles.alloc(s.h_tiles.size() + 1));
    CK(cudaMemcpyAsync(s.vtx2.p, s.h_vtx2.data(), s.h_vtx2.size() * sizeof(int4), cudaMemcpyHostToDevice, s.stream));
    if (!s.h_rec.empty()) CK(cudaMemcpyAsync(s.rec.p, s.h_rec.data(), s.h_rec.size() * sizeof(uint2), cudaMemcpyHostToDevice, s.stream));
    if (!s.h_nbw.empty()) CK(cudaMemcpyAsync(s.nbw.p, s.h_nbw.data(), s.h_nbw.size() * sizeof(int2), cudaMemcpyHostToDevice, s.stream));
    if (!s.h_tiles.empty()) CK(cudaMemcpyAsync(s.tiles.p, s.h_tiles.data(), s.h_tiles.size() * sizeof(TileDesc), cudaMemcpyHostToDevice, s.stream));
    {   // colour -> tile range, and the grid barrier of the multi-colour launch
        const int G = (int)s.groupTileStart.size() - 1;
        std::vector<int2> gt(std::max(G, 1), make_int2(0, 0));
        for (int g = 0; g < G; g++) gt[g] = make_int2(s.groupTileStart[g], s.groupTileStart[g + 1]);
        CK(s.groupTilesDev.alloc(gt.size()));
        CK(cudaMemcpyAsync(s.groupTilesDev.p, gt.data(), gt.size() * sizeof(int2), cudaMemcpyHostToDevice, s.stream));
        CK(s.gridBar.alloc(2));
        CK(cudaMemsetAsync(s.gridBar.p, 0, 2 * sizeof(uint32_t), s.stream));
    }
    CK(cudaStreamSynchronize(s.stream));
    std::vector<uint2>().swap(s.h_rec);
    std::vector<int2>().swap(s.h_nbw);
    std::vector<int4>().swap(s.h_vtx2);
    return cudaSuccess;
}

// HALO: a tile that holds boundary vertices (vertices this partition sends = vertices with ghost neighbours) waits for the
// neighbours' epoch and gathers past L1
__global__ void k_tile_halo_flags(int nTiles, TileDesc* tiles, const int32_t* __restrict__ boundaryPrefix) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nTiles) return;
    const int b = tiles[t].vBegin, e = b + tiles[t].nV;
    tiles[t].flags = (boundaryPrefix[e] != boundaryPrefix[b]) ? 1 : 0;
}
cudaError_t sweep_mark_halo_tiles(Scene& s, const int32_t* boundaryPrefixDev) {
    const int n = (int)s.h_tiles.size();
    if (n) k_tile_halo_flags<<<(n + 255) / 256, 256, 0, s.stream>>>(n, s.tiles.p, boundaryPrefixDev);
    return cudaStreamSynchronize(s.stream);
}

template <int L, bool HC, bool HALO, bool MULTI>
static cudaError_t launch_tiles(Scene& s, const SweepArgs& a, int nTiles) {
    if (nTiles <= 0) return cudaSuccess;
    const uint32_t variant = 1u << ((MULTI ? 4 : 0) + (HC ? 2 : 0) + (HALO ? 1 : 0));
    if (!(s.sweepAttrSet & variant)) {      // per scene = per device (a second scene on another GPU sets its own)
        CK(cudaFuncSetAttribute(k_sweep_tiles<L, HC, HALO, MULTI>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmTotal));
        CK(cudaFuncSetAttribute(k_sweep_tiles<L, HC, HALO, MULTI>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        s.sweepAttrSet |= variant;
    }
    const int grid = std::min(nTiles, s.numSMs * kSweepBlocksPerSM);
    if (MULTI) {
        // the blocks wait for one another between colours: a cooperative launch, so that all of them are resident
        void* params[] = {(void*)&a};
        return cudaLaunchCooperativeKernel((const void*)k_sweep_tiles<L, HC, HALO, MULTI>, dim3(grid), dim3(kSweepThreads), params, kSmTotal, s.stream);
    }
    k_sweep_tiles<L, HC, HALO, MULTI><<<grid, kSweepThreads, kSmTotal, s.stream>>>(a);
    return cudaGetLastError();
}

static void fill_common(Scene& s, SweepArgs& a) {
    a.pos = s.pos.p; a.prev = s.prev.p; a.inertia = s.inertia.p; a.vtx2 = s.vtx2.p; a.rec = s.rec.p; a.nbw = s.nbw.p; a.tiles = s.tiles.p;
    a.mp = s.meshParams.p; a.sp = s.stepParams; a.errFlags = s.errFlags.p;
    a.needPrev = 0;
    for (auto& m : s.meshes) if (m.obj.dampingHydrostatic != 0.f || m.obj.dampingDeviatoric != 0.f) a.needPrev = 1;
    if ((s.phys.usePlaneGround || s.phys.useBowlGround) && s.phys.boundaryFrictionDynamic != 0.f) a.needPrev = 1;
    a.allActive = 1;
    for (auto& m : s.meshes) if (!m.activeForMaterialSolve) a.allActive = 0;
}

// Can `nSteps` colours run as ONE launch? (contact-free colours only; all blocks must be resident at once)
bool sweep_multi_available(Scene& s) {
    if (s.multiStep < 0) {
        int perSM = 0;
        const bool ok = cudaFuncSetAttribute(k_sweep_tiles<kSweepLanes, false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmTotal) == cudaSuccess &&
                        cudaFuncSetAttribute(k_sweep_tiles<kSweepLanes, false, false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared) == cudaSuccess &&
                        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, k_sweep_tiles<kSweepLanes, false, false, true>, kSweepThreads, kSmTotal) == cudaSuccess;
        int coop = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, s.device);
        s.multiStep = (ok && coop && perSM >= kSweepBlocksPerSM) ? 1 : 0;
        cudaGetLastError();
    }
    return s.multiStep == 1;
}

// colours firstGroup, firstGroup+1, ... (cyclically) for nSteps steps in one launch; friction from step frictionFromStep on
cudaError_t sweep_launch_multi(Scene& s, int firstGroup, int nSteps, int frictionFromStep, const HaloView* hv) {
    SweepArgs a;
    std::memset(&a, 0, sizeof(a));
    fill_common(s, a);
    const int G = (int)s.groupTileStart.size() - 1;
    a.stepTiles = s.groupTilesDev.p; a.nSteps = nSteps; a.firstGroup = firstGroup; a.nGroups = G; a.frictionFromStep = frictionFromStep;
    a.gridBar = s.gridBar.p;
    int maxTiles = 0;
    for (int g = 0; g < G; g++) maxTiles = std::max(maxTiles, s.groupTileStart[g + 1] - s.groupTileStart[g]);
    if (hv) { a.hv = *hv; return launch_tiles<kSweepLanes, false, true, true>(s, a, maxTiles); }
    return launch_tiles<kSweepLanes, false, false, true>(s, a, maxTiles);
}

cudaError_t sweep_launch(Scene& s, int group, int applyFriction, bool hasContacts, const ContactView& cv, const HaloView* hv) {
    SweepArgs a;
    std::memset(&a, 0, sizeof(a));
    a.tileBegin = s.groupTileStart[group]; a.tileEnd = s.groupTileStart[group + 1];
    fill_common(s, a);
    a.applyFriction = applyFriction; a.cv = cv;
    const int nTiles = a.tileEnd - a.tileBegin;
    if (hv) a.hv = *hv;
    if (hv) return hasContacts ? launch_tiles<kSweepLanes, true, true, false>(s, a, nTiles) : launch_tiles<kSweepLanes, false, true, false>(s, a, nTiles);
    return hasContacts ? launch_tiles<kSweepLanes, true, false, false>(s, a, nTiles) : launch_tiles<kSweepLanes, false, false, false>(s, a, nTiles);
}


// ------------------------------------------------------------------------------------------------ line-search variant
// APPLY_LOCAL_LINE_SEARCH (M/VBD/VBDPhysics.cpp:25, 2467-2474; compiled OUT in the reference, named by north_star): after the
// 3x3 solve the step length is found by backtracking on the vertex merit energy - inertia + Neo-Hookean energy of the
// incident tets (evaluateVertexMeritEnergy, M/VBD/VBD_NeoHookean.cpp:90-124; backTracingLineSearchVBD, VBD_NeoHookean.h:115-153).
// The energy needs |F|^2 and det F of every incident tet at up to maxIters trial positions, which the tile records (det B',
// slots, summed weights) do not hold, so this variant is a kernel of its own over plain per-tet tables (Dm^-1, rest volume,
// vertex ids; uploaded when the option is switched on): one thread per vertex, incident tets in the reference's order
// (ascending tet id), F = Ds Dm^-1 as the reference forms it. It is a feature switch, not a fast path.
struct LsArgs {
    int vBegin, vEnd;
    float4* pos;
    const float4* prev;
    const float4* inertia;
    const int4* vtx2;
    const int32_t* off;       // [N + 1] over sorted ids
    const int2* ent;          // {global tet, order of the vertex in it}
    const int4* tetSid;       // sorted ids of the tet's vertices
    const float4* tetDm;      // 3 per tet: Dm^-1 column-major [0..8], rest volume
    const MeshParamsDev* mp;
    StepParamsDev sp;
    int applyFriction, needPrev;
    ContactView cv;
    int32_t* errFlags;
    float tau, tauGD;
    int maxIters;
    unsigned long long* shrinks;
};

// energy of the incident tets with the vertex at x; DERIV: also dE/dx, s1 = sum A |m|^2, G = sum A g g^T
template <bool DERIV>
__device__ __forceinline__ float ls_elastic(const LsArgs& a, int s, int b, int e, float miu, float lmbd, float aConst, float3 x,
                                            float3& dEdx, float& s1, float (&G)[6]) {
    float E = 0.f;
    for (int q = b; q < e; q++) {
        const int2 en = __ldg(a.ent + q);
        const int4 tv = __ldg(a.tetSid + en.x);
        const float4 d0 = __ldg(a.tetDm + 3 * (size_t)en.x), d1 = __ldg(a.tetDm + 3 * (size_t)en.x + 1), d2 = __ldg(a.tetDm + 3 * (size_t)en.x + 2);
        const float Dm[9] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w, d2.x};
        const float A = d2.y;
        const int ids[4] = {tv.x, tv.y, tv.z, tv.w};
        float3 p[4];
#pragma unroll
        for (int k = 0; k < 4; k++) p[k] = (ids[k] == s) ? x : f3(a.pos[ids[k]]);
        float Ds[9];
#pragma unroll
        for (int c = 0; c < 3; c++) { const float3 d = p[c + 1] - p[0]; Ds[3 * c] = d.x; Ds[3 * c + 1] = d.y; Ds[3 * c + 2] = d.z; }
        float F[9];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int r = 0; r < 3; r++) F[r + 3 * c] = Ds[r] * Dm[3 * c] + Ds[r + 3] * Dm[3 * c + 1] + Ds[r + 6] * Dm[3 * c + 2];
        float sum = 0.f;
#pragma unroll
        for (int k = 0; k < 9; k++) sum += F[k] * F[k];
        const float PhiD = 0.5f * (sum - 3.f);
        const float detF = F[0] * (F[4] * F[8] - F[7] * F[5]) - F[3] * (F[1] * F[8] - F[7] * F[2]) + F[6] * (F[1] * F[5] - F[4] * F[2]);
        const float k_ = detF - aConst;
        E += A * (miu * PhiD + lmbd * (0.5f * (k_ * k_)));
        if (DERIV) {
            const float dd[9] = {F[4] * F[8] - F[7] * F[5], F[6] * F[5] - F[3] * F[8], F[3] * F[7] - F[6] * F[4],
                                 F[7] * F[2] - F[1] * F[8], F[0] * F[8] - F[6] * F[2], F[6] * F[1] - F[0] * F[7],
                                 F[1] * F[5] - F[4] * F[2], F[3] * F[2] - F[0] * F[5], F[0] * F[4] - F[3] * F[1]};
            float m1, m2, m3;
            switch (en.y) {
                case 0: m1 = -Dm[0] - Dm[1] - Dm[2]; m2 = -Dm[3] - Dm[4] - Dm[5]; m3 = -Dm[6] - Dm[7] - Dm[8]; break;
                case 1: m1 = Dm[0]; m2 = Dm[3]; m3 = Dm[6]; break;
                case 2: m1 = Dm[1]; m2 = Dm[4]; m3 = Dm[7]; break;
                default: m1 = Dm[2]; m2 = Dm[5]; m3 = Dm[8]; break;
            }
            const float lk = lmbd * k_;
            float dEdF[9];
#pragma unroll
            for (int k = 0; k < 9; k++) dEdF[k] = A * (miu * F[k] + lk * dd[k]);
            dEdx.x += dEdF[0] * m1 + dEdF[3] * m2 + dEdF[6] * m3;
            dEdx.y += dEdF[1] * m1 + dEdF[4] * m2 + dEdF[7] * m3;
            dEdx.z += dEdF[2] * m1 + dEdF[5] * m2 + dEdF[8] * m3;
            // the vertex block of the tet Hessian: A (mu |m|^2 I + lmbd g g^T), g = cof(F) m (det F is affine in one vertex)
            const float gx = dd[0] * m1 + dd[3] * m2 + dd[6] * m3, gy = dd[1] * m1 + dd[4] * m2 + dd[7] * m3, gz = dd[2] * m1 + dd[5] * m2 + dd[8] * m3;
            s1 += A * (m1 * m1 + m2 * m2 + m3 * m3);
            G[0] += A * gx * gx; G[1] += A * gx * gy; G[2] += A * gx * gz; G[3] += A * gy * gy; G[4] += A * gy * gz; G[5] += A * gz * gz;
        }
    }
    return E;
}

template <bool HAS_CONTACTS>
__global__ void __launch_bounds__(128)
k_sweep_linesearch(const LsArgs a) {
    const int s = a.vBegin + blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= a.vEnd) return;
    const int4 h = __ldg(a.vtx2 + s);
    if (h.w < 0) return;
    const MeshParamsDev* P = a.mp + (h.w & 0x00FFFFFF);
    if (!P->activeSolve) return;
    const float3 x = f3(a.pos[s]);
    const float4 ys = __ldg(a.inertia + s);
    float4 xpv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.needPrev) xpv = __ldg(a.prev + s);
    const int b = __ldg(a.off + s), e = __ldg(a.off + s + 1);
    const float miu = P->miu, lmbd = P->lmbd, aConst = P->a;
    float3 dEdx = make_float3(0.f, 0.f, 0.f);
    float s1 = 0.f, G[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const float Eel0 = ls_elastic<true>(a, s, b, e, miu, lmbd, aConst, x, dEdx, s1, G);
    float4 dir;
    if (!vertex_tail<HAS_CONTACTS, true>(s, x, ys, f3(xpv), P, dEdx.x, dEdx.y, dEdx.z, s1, G[0], G[1], G[2], G[3], G[4], G[5], a.pos, a.sp, a.applyFriction,
                                        a.cv, a.errFlags, dir)) return;
    // merit energy = inertia (0.5 * m * |x - y|^2 / dt^2, the product formed in double as the reference's literal makes it) + elastic
    auto inertiaE = [&](float3 q) {
        const float dx = q.x - ys.x, dy = q.y - ys.y, dz = q.z - ys.z;
        const float sq = dx * dx + dy * dy + dz * dz;
        return (float)(0.5 * (double)ys.w * (double)sq * (double)a.sp.dtSqrReciprocal);
    };
    const float E0 = inertiaE(x) + Eel0;
    float alpha = fabsf(dir.w);
    const float shrink = dir.w < 0.f ? a.tauGD : a.tau;
    float3 xt = x;
    float3 dummy3 = make_float3(0.f, 0.f, 0.f); float dummy1 = 0.f, dummy6[6];
    int rejected = 0;
    for (int it = 0; it < a.maxIters; it++) {
        xt = make_float3(alpha * dir.x + x.x, alpha * dir.y + x.y, alpha * dir.z + x.z);
        const float en = inertiaE(xt) + ls_elastic<false>(a, s, b, e, miu, lmbd, aConst, xt, dummy3, dummy1, dummy6);
        if (en < E0) break;          // first Wolfe condition with c = 0 (the call site passes 0)
        alpha = alpha * shrink;
        rejected++;
    }
    if (a.maxIters > 0) a.pos[s] = make_float4(xt.x, xt.y, xt.z, 0.f);      // the LAST trial position stays, accepted or not, as in the reference
    if (rejected) atomicAdd(a.shrinks, (unsigned long long)rejected);
}

// Eigen's 3x3 inverse (cofactors * 1/det) in float, as TetMeshFEM computes DmInv (M/TetMesh/TetMeshFEM.cpp:390-413)
static void ls_inv3(const float m[9] /*column-major*/, float r[9]) {
    auto M = [&](int i, int j) { return m[i + 3 * j]; };
    auto cof = [&](int i, int j) {
        const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
        return M(i1, j1) * M(i2, j2) - M(i1, j2) * M(i2, j1);
    };
    const float c00 = cof(0, 0), c10 = cof(1, 0), c20 = cof(2, 0);
    const float det = c00 * M(0, 0) + c10 * M(1, 0) + c20 * M(2, 0);
    const float invdet = 1.f / det;
    r[0 + 3 * 0] = c00 * invdet; r[0 + 3 * 1] = c10 * invdet; r[0 + 3 * 2] = c20 * invdet;
    r[1 + 3 * 0] = cof(0, 1) * invdet; r[1 + 3 * 1] = cof(1, 1) * invdet; r[1 + 3 * 2] = cof(2, 1) * invdet;
    r[2 + 3 * 0] = cof(0, 2) * invdet; r[2 + 3 * 1] = cof(1, 2) * invdet; r[2 + 3 * 2] = cof(2, 2) * invdet;
}

cudaError_t linesearch_upload(Scene& s) {
    if (s.lsOff.p) return cudaSuccess;
    const int64_t N = s.nVerts;
    int64_t nT = 0;
    for (auto& m : s.meshes) nT += m.nT;
    std::vector<int32_t> off(N + 1, 0);
    std::vector<int4> tetSid(std::max<int64_t>(nT, 1));
    std::vector<float4> dm(3 * std::max<int64_t>(nT, 1));
    for (auto& m : s.meshes)
        for (int t = 0; t < m.nT; t++) {
            const int32_t* tv = &m.tets[4 * (size_t)t];
            tetSid[m.tetBase + t] = make_int4(m.sortedId[tv[0]], m.sortedId[tv[1]], m.sortedId[tv[2]], m.sortedId[tv[3]]);
            float Dm[9], Di[9];
            const float* x0 = &m.pos0[3 * (size_t)tv[0]];
            for (int c = 0; c < 3; c++) for (int d = 0; d < 3; d++) Dm[d + 3 * c] = m.pos0[3 * (size_t)tv[c + 1] + d] - x0[d];
            ls_inv3(Dm, Di);
            float4* o = &dm[3 * (size_t)(m.tetBase + t)];
            o[0] = make_float4(Di[0], Di[1], Di[2], Di[3]); o[1] = make_float4(Di[4], Di[5], Di[6], Di[7]);
            o[2] = make_float4(Di[8], m.restVol[t], 0.f, 0.f);
            for (int k = 0; k < 4; k++) off[m.sortedId[tv[k]] + 1]++;
        }
    for (int64_t v = 0; v < N; v++) off[v + 1] += off[v];
    std::vector<int2> ent(std::max<int64_t>(off[N], 1));
    {
        std::vector<int32_t> at(off.begin(), off.end() - 1);
        for (auto& m : s.meshes)                    // tets ascending inside a mesh: the reference's incident-tet order
            for (int t = 0; t < m.nT; t++)
                for (int k = 0; k < 4; k++) ent[at[m.sortedId[m.tets[4 * (size_t)t + k]]]++] = make_int2((int)(m.tetBase + t), k);
    }
    CK(s.lsOff.alloc(off.size())); CK(s.lsEnt.alloc(ent.size())); CK(s.lsTetSid.alloc(tetSid.size())); CK(s.lsTetDm.alloc(dm.size())); CK(s.lsShrinks.alloc(1));
    CK(cudaMemcpy(s.lsOff.p, off.data(), off.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(s.lsEnt.p, ent.data(), ent.size() * sizeof(int2), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(s.lsTetSid.p, tetSid.data(), tetSid.size() * sizeof(int4), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(s.lsTetDm.p, dm.data(), dm.size() * sizeof(float4), cudaMemcpyHostToDevice));
    CK(cudaMemset(s.lsShrinks.p, 0, sizeof(unsigned long long)));
    return cudaSuccess;
}

// ------------------------------------------------------------------------------------------------ initializationType 5
// VBDBaseTetMesh::forwardStepSimplecticEuler (M/VBD/VBD_BaseMaterial.cpp:134-147, the initial guess of initializationType 5):
// the elastic force of EVERY vertex at the positions the substep starts from (VBDTetMeshNeoHookean::evaluateInternalForce /
// accumlateMaterialForce, VBD_NeoHookean.cpp:29-37, 781-870: no damping), then velocity = velocitiesPrev + dt f / mass and
// position += velocity dt - fixed vertices included, as in the reference. Jacobi over the mesh: the new positions go to a
// scratch array and are copied back by a second launch. Runs once per substep over the per-tet tables of the line search;
// scenes without a type-5 mesh launch nothing.
__global__ void __launch_bounds__(128)
k_symplectic_guess(const LsArgs a, int n, float4* __restrict__ vel, const float* __restrict__ mass, float4* __restrict__ out) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const MeshParamsDev* P = a.mp + (__ldg(a.vtx2 + s).w & 0x00FFFFFF);
    if (!P->activeSolve || P->initType != 5) return;
    const float3 x = f3(a.pos[s]);
    float3 dEdx = make_float3(0.f, 0.f, 0.f);
    float s1 = 0.f, G[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    ls_elastic<true>(a, s, __ldg(a.off + s), __ldg(a.off + s + 1), P->miu, P->lmbd, P->a, x, dEdx, s1, G);
    const float dt = a.sp.dt, invMass = 1.f / __ldg(mass + s);
    const float4 vp = vel[s];
    const float3 v = make_float3(__fadd_rn(vp.x, __fmul_rn(__fmul_rn(dt, -dEdx.x), invMass)), __fadd_rn(vp.y, __fmul_rn(__fmul_rn(dt, -dEdx.y), invMass)),
                                 __fadd_rn(vp.z, __fmul_rn(__fmul_rn(dt, -dEdx.z), invMass)));
    vel[s] = make_float4(v.x, v.y, v.z, 0.f);      // (the reference stores it too; updateVelocities overwrites it)
    out[s] = make_float4(__fadd_rn(x.x, __fmul_rn(v.x, dt)), __fadd_rn(x.y, __fmul_rn(v.y, dt)), __fadd_rn(x.z, __fmul_rn(v.z, dt)), 0.f);
}

__global__ void k_symplectic_apply(int n, const int4* __restrict__ vtx2, const MeshParamsDev* __restrict__ mp, const float4* __restrict__ in, float4* __restrict__ pos) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const MeshParamsDev* P = mp + (__ldg(vtx2 + s).w & 0x00FFFFFF);
    if (!P->activeSolve || P->initType != 5) return;
    pos[s] = in[s];
}

cudaError_t sweep_symplectic_guess(Scene& s) {
    if (!s.symScratch.p) return cudaSuccess;
    const int N = (int)s.nVerts;
    LsArgs a;
    std::memset(&a, 0, sizeof(a));
    a.pos = s.pos.p; a.vtx2 = s.vtx2.p; a.off = s.lsOff.p; a.ent = s.lsEnt.p; a.tetSid = s.lsTetSid.p; a.tetDm = s.lsTetDm.p; a.mp = s.meshParams.p; a.sp = s.stepParams;
    k_symplectic_guess<<<(N + 127) / 128, 128, 0, s.stream>>>(a, N, s.vel.p, s.mass.p, s.symScratch.p);
    k_symplectic_apply<<<(N + 255) / 256, 256, 0, s.stream>>>(N, s.vtx2.p, s.meshParams.p, s.symScratch.p, s.pos.p);
    s.stats.kernelLaunches += 2;
    return cudaGetLastError();
}

cudaError_t sweep_launch_linesearch(Scene& s, int group, int applyFriction, bool hasContacts, const ContactView& cv) {
    const int vb = s.groupStart[group], ve = s.groupStart[group + 1];
    if (ve <= vb) return cudaSuccess;
    SweepArgs c;
    std::memset(&c, 0, sizeof(c));
    fill_common(s, c);
    LsArgs a;
    std::memset(&a, 0, sizeof(a));
    a.vBegin = vb; a.vEnd = ve; a.pos = s.pos.p; a.prev = s.prev.p; a.inertia = s.inertia.p; a.vtx2 = s.vtx2.p;
    a.off = s.lsOff.p; a.ent = s.lsEnt.p; a.tetSid = s.lsTetSid.p; a.tetDm = s.lsTetDm.p; a.mp = s.meshParams.p; a.sp = s.stepParams;
    a.applyFriction = applyFriction; a.needPrev = c.needPrev; a.cv = cv; a.errFlags = s.errFlags.p;
    a.tau = s.lineSearch.tau; a.tauGD = s.lineSearch.tauGD; a.maxIters = s.lineSearch.maxIters; a.shrinks = s.lsShrinks.p;
    const int B = 128, G = (ve - vb + B - 1) / B;
    if (hasContacts) k_sweep_linesearch<true><<<G, B, 0, s.stream>>>(a);
    else k_sweep_linesearch<false><<<G, B, 0, s.stream>>>(a);
    return cudaGetLastError();
}

}  // namespace gb
