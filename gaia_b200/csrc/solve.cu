// Per-colour VBD vertex solve and the elementwise passes around it, hand-written for sm_100a.
//
// Reference semantics (CPU/TBB path):
//   applyInitialStep / computeInertia / forwardStep    M/VBD/VBD_BaseMaterial.h:201-259, VBD_BaseMaterial.cpp:7-132
//   accumlateMaterialForceAndHessian2 (Neo-Hookean)    M/VBD/VBD_NeoHookean.cpp:379-581
//   accumlateInertiaForceAndHessian                    M/VBD/VBD_BaseMaterial.h:359-366
//   accumlateBoundaryForceAndHessian + friction        M/VBD/VBDPhysics.cpp:2821-2954, 2563-2592
//   accumlateCollisionForceAndHessian[PerCollision]    M/VBD/VBDPhysics.cpp:2956-3047
//   VBDStepWithCollision + solve3x3_psd_stable         M/VBD/VBDPhysics.cpp:2402-2483, 3P/CuMatrix/.../CuMatrix.h:235-267
//   colour sweep loop                                  M/VBD/VBDPhysics.cpp:1659-1702
//   updateVelocities / dampVelocities                  M/VBD/VBDPhysics.cpp:2486-2561
//   Chebyshev accelerator (GPU-path feature)           M/VBD/VBDPhysics.cpp:3558-3574, VBDPhysicsCompute.cu:910-944
//
// Kernel design: the work is a sparse gather bounded by HBM, not a contraction — no tensor cores.
//   * vertices are colour-sorted, so a colour is one contiguous range [vBegin,vEnd) of every SoA array;
//   * each vertex is owned by a sub-warp of L lanes; lane l streams the incidence records l, l+L, ... of the vertex
//     (two coalesced 16-byte loads per record), gathers the three neighbour positions (16-byte loads that hit L2:
//     the position array of a 10M-tet mesh is 28 MB against 126 MB of L2) and accumulates the closed-form
//     Neo-Hookean force / Hessian of SURVEY.md §8(a) "math restatement" (≈100 flop instead of the 9x9 contraction);
//   * a butterfly shuffle reduces 10 partial sums; lane 0 adds inertia, boundary and contact terms, solves the 3x3
//     block and writes the new position in place (safe: a colour only reads other colours' positions).
#include "scene_impl.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>

namespace gb {

#define CK(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return _e; } while (0)

static const int kBlock = 256;

__device__ __forceinline__ float3 f3(float4 v) { return make_float3(v.x, v.y, v.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return make_float3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return make_float3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator*(float s, float3 a) { return make_float3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ float dot3(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float3 cross3(float3 a, float3 b) {
    return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

// streaming (read-once) loads: keep them out of L1 so the gathered positions stay cached
__device__ __forceinline__ int4 ld_stream(const int4* p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float4 ld_stream(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

// IPC-style mollified friction (computeVertexFriction, VBDPhysics.cpp:2563-2592) for an axis-aligned tangent basis
// T = [e_a, e_b]: adds force and the diagonal Hessian entries (a,a) and (b,b).
__device__ __forceinline__ void boundary_friction(float mu, float lambda, float ua, float ub, float epsU, float& fa, float& fb, float& haa) {
    const float uNorm = sqrtf(ua * ua + ub * ub);
    fa = 0.f; fb = 0.f; haa = 0.f;
    if (uNorm > 0.f) {
        const float f1 = (uNorm > epsU) ? 1.f / uNorm : (-uNorm / epsU + 2.f) / epsU;
        const float s = -mu * lambda * f1;
        fa = s * ua; fb = s * ub;
        haa = mu * lambda * f1;
    }
}

__device__ __forceinline__ uint32_t ld_stream(const uint32_t* p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}

static const int kNbMax = 32;              // neighbour positions staged in shared memory per vertex (more spill to global gathers)
static const int kNbStride = kNbMax + 1;   // +1 float4 of padding de-correlates the rows' banks

__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 r;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
    return r;
}

// The vertex-level part of VBDStepWithCollision after the incident-tet sums are known: damping, inertia, world box +
// friction, contacts, 3x3 solve, in-place update. Runs with one vertex per thread (all lanes busy).
template <bool HAS_CONTACTS>
__device__ __forceinline__ bool solve_tail(int s, float3 x, float4 ys, float miu, float lmbd, float dampH, float dampD,
                                           float fx, float fy, float fz, float s1, float g00, float g01, float g02, float g11, float g12, float g22,
                                           float4* pos, const float4* __restrict__ prev, const StepParamsDev& sp, int applyFriction,
                                           const ContactView& cv, int32_t* errFlags, float4& xNew) {
    // elastic force and Hessian: f = -(mu*fmu + lmbd*flm), H = mu*s1*I + lmbd*G
    float F0 = -fx, F1 = -fy, F2 = -fz;
    float hd = miu * s1;
    float h00 = lmbd * g00 + hd, h01 = lmbd * g01, h02 = lmbd * g02, h11 = lmbd * g11 + hd, h12 = lmbd * g12, h22 = lmbd * g22 + hd;
    const bool needPrev = (dampH != 0.f) | (dampD != 0.f);
    float3 xp = make_float3(0.f, 0.f, 0.f);
    bool havePrev = false;
    if (needPrev) {
        xp = f3(__ldg(prev + s)); havePrev = true;
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
    // inertia
    {
        const float m = ys.w;
        const float w = m * sp.dtSqrReciprocal;
        F0 += (m * (ys.x - x.x)) * sp.dtSqrReciprocal;
        F1 += (m * (ys.y - x.y)) * sp.dtSqrReciprocal;
        F2 += (m * (ys.z - x.z)) * sp.dtSqrReciprocal;
        h00 += w; h11 += w; h22 += w;
    }
    // world-box penalty + friction, per violated axis (CPU semantics)
    if (sp.usePlaneGround) {
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
                    if (!havePrev) { xp = f3(__ldg(prev + s)); havePrev = true; }
                    const float dx[3] = {x.x - xp.x, x.y - xp.y, x.z - xp.z};
                    const int a = (d + 1) % 3, b2 = (d + 2) % 3;
                    float fa, fb, haa;
                    boundary_friction(sp.muBoundary, depth * sp.kBoundary, dx[a], dx[b2], sp.epsVBoundary * sp.dt, fa, fb, haa);
                    Fb[a] += fa; Fb[b2] += fb; Hb[a] += haa; Hb[b2] += haa;
                }
            }
        }
        F0 += Fb[0]; F1 += Fb[1]; F2 += Fb[2];
        h00 += Hb[0]; h11 += Hb[1]; h22 += Hb[2];
    }
    // vertex-face contacts: own contact with weight 1, contacts on faces this vertex belongs to with weight -bary
    if (HAS_CONTACTS) {
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
        if (fabsf(det) < 1e-5f * (fabsf(h00 * i11) + fabsf(h01 * i12) + fabsf(h02 * i13))) {
            dx0 = F0; dx1 = F1; dx2 = F2; step = sp.stepSizeGD;
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
        xNew = make_float4(x.x + step * dx0, x.y + step * dx1, x.z + step * dx2, 0.f);
        pos[s] = xNew;
        return true;
    }
    return false;
}

// Neighbour position gather. HALO launches read past L1 (ld.global.cg): ghost slots are written by the neighbours over
// NVLink while the launch runs, and an L1 line fetched before the epoch was seen could hold the old value.
#ifndef GAIA_GATHER_CG
#define GAIA_GATHER_CG 0
#endif
template <bool HALO>
__device__ __forceinline__ float4 ld_nb(const float4* p) { return (HALO || GAIA_GATHER_CG) ? __ldcg(p) : __ldg(p); }

// Fused halo exchange, consumer side: wait until every neighbour partition has published `epoch` (its previous colour
// launch, pushes included, is complete). One lane polls; 5 s without progress = a neighbour died (errFlags[4]).
__device__ __forceinline__ void halo_wait(const HaloPeerTable* tab, uint32_t epoch, int nPeers, int32_t* errFlags) {
    for (int p = 0; p < nPeers; p++) {
        if (*(volatile int32_t*)(errFlags + 4)) return;
        unsigned long long t0;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
        for (;;) {
            uint32_t v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(tab->myFlag[p]) : "memory");
            if ((int32_t)(v - epoch) >= 0) break;
            unsigned long long t1;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
            if (t1 - t0 > 5000000000ull) { atomicAdd(errFlags + 4, 1); return; }
            __nanosleep(64);
        }
    }
}

// One block owns kBlock consecutive vertices of the colour and walks them in ROUNDS = L rounds of kBlock/L vertices:
// in a round each vertex is served by a sub-warp of L lanes (incident-tet sums), the sums are parked in shared memory,
// and after the last round every thread finishes ONE vertex (solve_tail) — so the scalar tail runs at full SIMT width.
//
// HALO (multi-GPU partitions): the exchange of this colour's boundary positions is part of the launch.
// Epochs count colour launches: launch number o of a batch (hv.epochOffset = o) first publishes "everything before me
// is complete" (base + o, base = ctl[0]) to the neighbours - stream order makes the previous launch's peer stores
// visible first - and a block that holds boundary vertices (bit 30 of the vertex header; found through a prefix count
// over the block's vertex range) waits until every neighbour has published the same epoch, gathers its neighbour
// positions past L1 (ghost slots are written by peers over NVLink while the launch runs) and after its solve stores the
// new positions straight into the neighbours' ghost slots. Blocks without boundary vertices never wait, so the NVLink
// latency hides behind them; nothing is added to their critical path (no barrier, no atomic).
template <int L, bool HAS_CONTACTS, int PF, int ROUNDS, bool GHOST>
__device__ __forceinline__ void
accumulate_block(int vBegin, int vEnd, float4* pos, const float4* __restrict__ prev, const float4* __restrict__ inertia,
               const int4* __restrict__ vtx, const float4* __restrict__ incW, const uint32_t* __restrict__ incIdx,
               const int32_t* __restrict__ nbList, const MeshParamsDev* __restrict__ mp,
               const StepParamsDev& sp, int applyFriction, const ContactView& cv, int32_t* errFlags, const HaloView& hv) {
    constexpr int VPR = kBlock / L;        // vertices per round
    constexpr int VPB = VPR * ROUNDS;      // vertices per block
    extern __shared__ float4 smem4[];
    float4* smNb = smem4;                                   // [VPR][kNbStride] neighbour positions relative to the owner
    float* acc = (float*)(smem4 + VPR * kNbStride);         // [10][VPB] incident-tet sums
    const int lane = threadIdx.x & (L - 1);
    const int sub = threadIdx.x / L;
    const uint32_t rowAddr = (uint32_t)__cvta_generic_to_shared(smNb + sub * kNbStride);
    float4* myNb = smNb + sub * kNbStride;
    const int blockBase = vBegin + blockIdx.x * VPB;
    const float4* posRO = pos;   // neighbour gathers only touch other colours: read-only for this launch
    bool waited = false;         // GHOST: the block has seen the neighbours' epoch

#pragma unroll 1
    for (int r = 0; r < ROUNDS; r++) {
        const int slot = r * VPR + sub;
        const int s = blockBase + slot;
        const bool inRange = s < vEnd;
        // ---- level 0 of the dependency chain: everything that only needs s
        int4 va = make_int4(0, 0, 0, (int)0x80000000);      // {incOff, nbOff, nInc | nNb << 16, meshId | fixed << 31}
        float4 xs = make_float4(0.f, 0.f, 0.f, 0.f);
        if (inRange) { va = __ldg(vtx + s); xs = pos[s]; }
        bool active = inRange && va.w >= 0;
        float miu = 0.f, lmbd = 0.f, aConst = 0.f;
        if (active) {
            const MeshParamsDev* P = mp + (va.w & 0x00FFFFFF);
            active = P->activeSolve != 0;
            miu = P->miu; lmbd = P->lmbd; aConst = P->a;
        }
        float fx = 0.f, fy = 0.f, fz = 0.f, s1 = 0.f, g00 = 0.f, g01 = 0.f, g02 = 0.f, g11 = 0.f, g12 = 0.f, g22 = 0.f;
        const float3 x = f3(xs);
        const int e = va.x + (active ? (va.z & 0xFFFF) : 0), nb0 = va.y;
        const int nnbAll = active ? (va.z >> 16) : 0;

        // ---- level 1: first incidence records and the neighbour ids (both only depend on va)
        int j = va.x + lane;
        float4 W = make_float4(0.f, 0.f, 0.f, 0.f), W1 = W, W2 = W;
        uint32_t I = 0, I1 = 0, I2 = 0;
        if (j < e) { W = ld_stream(incW + j); I = ld_stream(incIdx + j); }
        if (PF >= 1 && j + L < e) { W1 = ld_stream(incW + j + L); I1 = ld_stream(incIdx + j + L); }
        if (PF >= 2 && j + 2 * L < e) { W2 = ld_stream(incW + j + 2 * L); I2 = ld_stream(incIdx + j + 2 * L); }
        {
            const int nnb = min(nnbAll, kNbMax);
            int idx[kNbMax / L];
#pragma unroll
            for (int q = 0; q < kNbMax / L; q++) { const int i = lane + q * L; idx[q] = (i < nnb) ? (int)ld_stream((const uint32_t*)nbList + nb0 + i) : -1; }
            // ---- level 2: stage the distinct neighbour positions once (each is used by ~5 incident tets); L2 hits
            if (GHOST) {
                if (!waited) {
                    if (threadIdx.x == 0 && !(hv.debug & 1)) halo_wait(hv.tab, *(volatile const uint32_t*)hv.ctl + (uint32_t)hv.epochOffset, hv.nPeers, errFlags);
                    __syncthreads();
                    waited = true;
                }
#pragma unroll
                for (int q = 0; q < kNbMax / L; q++)
                    if (idx[q] >= 0) {
                        const float4 p = __ldcg(posRO + idx[q]);
                        myNb[lane + q * L] = make_float4(p.x - x.x, p.y - x.y, p.z - x.z, 0.f);
                    }
            } else {
#pragma unroll
                for (int q = 0; q < kNbMax / L; q++)
                    if (idx[q] >= 0) {
                        const float4 p = __ldg(posRO + idx[q]);
                        myNb[lane + q * L] = make_float4(p.x - x.x, p.y - x.y, p.z - x.z, 0.f);
                    }
            }
        }
        __syncwarp();
        const float lm6 = lmbd * (-1.f / 6.f);
        for (; j < e; j += L) {
            // software pipeline: the record PF+1 iterations ahead is requested before this one is consumed
            float4 Wn = W; uint32_t In = 0;
            const int jn = j + (PF + 1) * L;
            if (jn < e) { Wn = ld_stream(incW + jn); In = ld_stream(incIdx + jn); }
            float3 pa, pb, pc;
            if (nnbAll <= kNbMax) {
                pa = f3(lds128(rowAddr + ((I & 0xFFu) << 4)));
                pb = f3(lds128(rowAddr + ((I >> 4) & 0xFF0u)));
                pc = f3(lds128(rowAddr + ((I >> 12) & 0xFF0u)));
            } else {
                pa = f3(ld_nb<GHOST>(posRO + __ldg(nbList + nb0 + (I & 255)))) - x;
                pb = f3(ld_nb<GHOST>(posRO + __ldg(nbList + nb0 + ((I >> 8) & 255)))) - x;
                pc = f3(ld_nb<GHOST>(posRO + __ldg(nbList + nb0 + ((I >> 16) & 255)))) - x;
            }
            // n = cof(Ds')(1,1,1) = b x c + c x a + a x b = (b-a) x (c-a);  det(Ds') = a . n;  A*g = -n/6
            // (explicit _rn intrinsics: the contraction pattern is fixed, so every instantiation of this body - any L, PF,
            // ROUNDS, with or without the ghost path - produces the same bits)
            const float3 u = pb - pa, v = pc - pa;
            float3 n;
            n.x = __fmaf_rn(u.y, v.z, -__fmul_rn(u.z, v.y));
            n.y = __fmaf_rn(u.z, v.x, -__fmul_rn(u.x, v.z));
            n.z = __fmaf_rn(u.x, v.y, -__fmul_rn(u.y, v.x));
            const float detDs = __fmaf_rn(pa.z, n.z, __fmaf_rn(pa.y, n.y, __fmul_rn(pa.x, n.x)));
            const float k = __fmaf_rn(detDs, W.w, -aConst);            // det F - a,  det F = det(Ds') det(B')
            const float c6 = __fmul_rn(lm6, k);
            // A*(mu F m0 + lmbd k g)
            fx = __fmaf_rn(c6, n.x, __fmaf_rn(miu, __fmaf_rn(W.z, pc.x, __fmaf_rn(W.y, pb.x, __fmul_rn(W.x, pa.x))), fx));
            fy = __fmaf_rn(c6, n.y, __fmaf_rn(miu, __fmaf_rn(W.z, pc.y, __fmaf_rn(W.y, pb.y, __fmul_rn(W.x, pa.y))), fy));
            fz = __fmaf_rn(c6, n.z, __fmaf_rn(miu, __fmaf_rn(W.z, pc.z, __fmaf_rn(W.y, pb.z, __fmul_rn(W.x, pa.z))), fz));
            s1 = __fsub_rn(s1, __fadd_rn(__fadd_rn(W.x, W.y), W.z));   // A*|m0|^2
            const float nx = __fmul_rn(W.w, n.x), ny = __fmul_rn(W.w, n.y), nz = __fmul_rn(W.w, n.z);   // 6 A g g^T = det(B') n n^T
            g00 = __fmaf_rn(nx, n.x, g00); g01 = __fmaf_rn(nx, n.y, g01); g02 = __fmaf_rn(nx, n.z, g02);
            g11 = __fmaf_rn(ny, n.y, g11); g12 = __fmaf_rn(ny, n.z, g12); g22 = __fmaf_rn(nz, n.z, g22);
            if (PF == 0) { W = Wn; I = In; }
            else if (PF == 1) { W = W1; I = I1; W1 = Wn; I1 = In; }
            else { W = W1; I = I1; W1 = W2; I1 = I2; W2 = Wn; I2 = In; }
        }
#pragma unroll
        for (int o = L / 2; o > 0; o >>= 1) {
            fx += __shfl_xor_sync(0xffffffffu, fx, o); fy += __shfl_xor_sync(0xffffffffu, fy, o); fz += __shfl_xor_sync(0xffffffffu, fz, o);
            s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            g00 += __shfl_xor_sync(0xffffffffu, g00, o); g01 += __shfl_xor_sync(0xffffffffu, g01, o); g02 += __shfl_xor_sync(0xffffffffu, g02, o);
            g11 += __shfl_xor_sync(0xffffffffu, g11, o); g12 += __shfl_xor_sync(0xffffffffu, g12, o); g22 += __shfl_xor_sync(0xffffffffu, g22, o);
        }
        if (lane == 0) {
            acc[0 * VPB + slot] = fx; acc[1 * VPB + slot] = fy; acc[2 * VPB + slot] = fz; acc[3 * VPB + slot] = s1;
            acc[4 * VPB + slot] = g00; acc[5 * VPB + slot] = g01; acc[6 * VPB + slot] = g02;
            acc[7 * VPB + slot] = g11; acc[8 * VPB + slot] = g12; acc[9 * VPB + slot] = g22;
        }
        __syncwarp();      // the row of smNb is rewritten by the next round
    }
}

template <int L, bool HAS_CONTACTS, int PF, int ROUNDS, bool HALO>
__global__ void __launch_bounds__(kBlock, 4)
k_solve_colour(int vBegin, int vEnd, float4* pos, const float4* __restrict__ prev, const float4* __restrict__ inertia,
               const int4* __restrict__ vtx, const float4* __restrict__ incW, const uint32_t* __restrict__ incIdx,
               const int32_t* __restrict__ nbList, const MeshParamsDev* __restrict__ mp,
               StepParamsDev sp, int applyFriction, ContactView cv, int32_t* errFlags, HaloView hv) {
    constexpr int VPR = kBlock / L;
    constexpr int VPB = VPR * ROUNDS;
    extern __shared__ float4 smem4[];
    float* acc = (float*)(smem4 + VPR * kNbStride);         // [10][VPB] incident-tet sums
    const int blockBase = vBegin + blockIdx.x * VPB;
    bool ghostBlock = false;
    if (HALO) {
        if (blockIdx.x == 0 && threadIdx.x < hv.nPeers) {
            const uint32_t done = *(volatile const uint32_t*)hv.ctl + (uint32_t)hv.epochOffset;
            __threadfence_system();
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(hv.tab->peerFlag[threadIdx.x]), "r"(done) : "memory");
        }
        // a block that holds boundary vertices takes the ghost-safe copy of the accumulation (wait + gathers past L1);
        // all other blocks run exactly the single-GPU code. Both copies use the same pinned arithmetic.
        ghostBlock = __ldg(hv.boundaryPrefix + min(blockBase + VPB, vEnd)) != __ldg(hv.boundaryPrefix + min(blockBase, vEnd));
    }
    if (HALO && ghostBlock) accumulate_block<L, HAS_CONTACTS, PF, ROUNDS, true>(vBegin, vEnd, pos, prev, inertia, vtx, incW, incIdx, nbList, mp, sp, applyFriction, cv, errFlags, hv);
    else accumulate_block<L, HAS_CONTACTS, PF, ROUNDS, false>(vBegin, vEnd, pos, prev, inertia, vtx, incW, incIdx, nbList, mp, sp, applyFriction, cv, errFlags, hv);
    bool pushed = false;         // HALO: this thread stored into a neighbour's memory
    __syncthreads();
    // ---- transposed tail: thread t finishes vertex blockBase + t
    for (int slot = threadIdx.x; slot < VPB; slot += kBlock) {
        const int s = blockBase + slot;
        if (s >= vEnd) continue;
        const int4 va = __ldg(vtx + s);
        if (va.w < 0) continue;
        const MeshParamsDev* P = mp + (va.w & 0x00FFFFFF);
        if (!P->activeSolve) continue;
        const float4 xs = pos[s];
        const float4 ys = __ldg(inertia + s);
        float4 xNew;
        const bool wrote = solve_tail<HAS_CONTACTS>(s, f3(xs), ys, P->miu, P->lmbd, P->dampH, P->dampD,
                                 acc[0 * VPB + slot], acc[1 * VPB + slot], acc[2 * VPB + slot], acc[3 * VPB + slot],
                                 acc[4 * VPB + slot] * (1.f / 6.f), acc[5 * VPB + slot] * (1.f / 6.f), acc[6 * VPB + slot] * (1.f / 6.f),
                                 acc[7 * VPB + slot] * (1.f / 6.f), acc[8 * VPB + slot] * (1.f / 6.f), acc[9 * VPB + slot] * (1.f / 6.f),
                                 pos, prev, sp, applyFriction, cv, errFlags, xNew);
        if (HALO && wrote && (va.w & kHaloBoundaryBit) && !(hv.debug & 2)) {
            const int si = __ldg(hv.sendIdx + s);
            for (int q = 0; q < (si & 15); q++) {
                const int2 en = __ldg(hv.sendEnt + (si >> 4) + q);
                hv.tab->peerPos[en.x][en.y] = xNew;
            }
            pushed = true;
        }
    }
    // the next launch of the stream publishes the epoch; the stores must be performed system-wide before it does
    if (HALO && pushed) __threadfence_system();
}


// End of a batch of fused colour launches: publish the last epoch and advance the base.
__global__ void k_halo_advance(HaloView hv, int n) {
    const uint32_t done = *(volatile const uint32_t*)hv.ctl + (uint32_t)n;
    if (threadIdx.x < hv.nPeers) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(hv.tab->peerFlag[threadIdx.x]), "r"(done) : "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) hv.ctl[0] = done;
}

// applyInitialStep for every vertex of every active mesh (one launch)
__global__ void __launch_bounds__(kBlock)
k_initial_step(int n, float4* __restrict__ pos, float4* __restrict__ prev, float4* __restrict__ inertia, const float4* __restrict__ vel,
               float4* __restrict__ velPrev, float4* __restrict__ accel, const int32_t* __restrict__ info, const float* __restrict__ mass,
               const MeshParamsDev* __restrict__ mp, const int32_t* __restrict__ meshSubsteps, StepParamsDev sp) {
    const int s = blockIdx.x * kBlock + threadIdx.x;
    if (s >= n) return;
    const int inf = __ldg(info + s);
    const int meshId = inf & 0x00FFFFFF;
    const MeshParamsDev P = mp[meshId];
    if (!P.activeSolve) return;
    const int nDone = __ldg(meshSubsteps + meshId);      // substeps this mesh has already taken
    const float3 x = f3(pos[s]);
    float3 v = f3(__ldg(vel + s));
    const float dt = sp.dt;
    prev[s] = make_float4(x.x, x.y, x.z, 0.f);
    const int type = P.initType;
    float3 vp = make_float3(0.f, 0.f, 0.f), acc = make_float3(0.f, 0.f, 0.f);
    if (type == 2 || type == 4 || type == 6) {
        // velocitiesPrev() = velocities() BEFORE gravity is added: it is the velocity the last substep ended with
        if (type == 6 && nDone >= 1) {
            const float3 old = f3(velPrev[s]);
            acc = make_float3((v.x - old.x) / dt, (v.y - old.y) / dt, (v.z - old.z) / dt);
            accel[s] = make_float4(acc.x, acc.y, acc.z, 0.f);
        }
        velPrev[s] = make_float4(v.x, v.y, v.z, 0.f);
        vp = v;
    }
    // explicit _rn intrinsics: no FMA contraction, so y is bit-identical to the reference's separate multiply and add
    if (sp.associateGravityWithInertia) {
        v.x = __fadd_rn(v.x, __fmul_rn(dt, sp.gravity[0])); v.y = __fadd_rn(v.y, __fmul_rn(dt, sp.gravity[1])); v.z = __fadd_rn(v.z, __fmul_rn(dt, sp.gravity[2]));
    }
    if (inf < 0) v = make_float3(0.f, 0.f, 0.f);
    const float3 y = make_float3(__fadd_rn(x.x, __fmul_rn(dt, v.x)), __fadd_rn(x.y, __fmul_rn(dt, v.y)), __fadd_rn(x.z, __fmul_rn(dt, v.z)));
    inertia[s] = make_float4(y.x, y.y, y.z, __ldg(mass + s));   // .w = lumped mass, read by the colour solve
    float3 xn = x;
    switch (type) {
        case 0: xn = y; break;
        case 2: xn = make_float3(__fadd_rn(x.x, __fmul_rn(dt, vp.x)), __fadd_rn(x.y, __fmul_rn(dt, vp.y)), __fadd_rn(x.z, __fmul_rn(dt, vp.z))); break;
        case 3: xn = make_float3(__fadd_rn(__fmul_rn(0.5f, x.x), __fmul_rn(0.5f, y.x)), __fadd_rn(__fmul_rn(0.5f, x.y), __fmul_rn(0.5f, y.y)),
                                 __fadd_rn(__fmul_rn(0.5f, x.z), __fmul_rn(0.5f, y.z))); break;
        case 4: {
            const float c = __fmul_rn(__fmul_rn(dt, dt), P.initRatio_g);
            xn = make_float3(__fadd_rn(__fadd_rn(x.x, __fmul_rn(dt, vp.x)), __fmul_rn(c, sp.gravity[0])),
                             __fadd_rn(__fadd_rn(x.y, __fmul_rn(dt, vp.y)), __fmul_rn(c, sp.gravity[1])),
                             __fadd_rn(__fadd_rn(x.z, __fmul_rn(dt, vp.z)), __fmul_rn(c, sp.gravity[2])));
        } break;
        case 6:
            if (nDone >= 1) {
                const float gn = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(sp.gravity[0], sp.gravity[0]), __fmul_rn(sp.gravity[1], sp.gravity[1])),
                                                      __fmul_rn(sp.gravity[2], sp.gravity[2])));
                const float3 gd = make_float3(sp.gravity[0] / gn, sp.gravity[1] / gn, sp.gravity[2] / gn);
                float ac = __fadd_rn(__fadd_rn(__fmul_rn(acc.x, gd.x), __fmul_rn(acc.y, gd.y)), __fmul_rn(acc.z, gd.z));
                ac = ac < gn ? ac : gn;
                ac = ac > 1e-5f ? ac : 0.f;
                const float c = __fmul_rn(dt, dt);
                xn = make_float3(__fadd_rn(__fadd_rn(x.x, __fmul_rn(dt, vp.x)), __fmul_rn(__fmul_rn(c, gd.x), ac)),
                                 __fadd_rn(__fadd_rn(x.y, __fmul_rn(dt, vp.y)), __fmul_rn(__fmul_rn(c, gd.y), ac)),
                                 __fadd_rn(__fadd_rn(x.z, __fmul_rn(dt, vp.z)), __fmul_rn(__fmul_rn(c, gd.z), ac)));
            }
            break;
        default: break;
    }
    if (type != 1) pos[s] = make_float4(xn.x, xn.y, xn.z, 0.f);
}

__global__ void k_count_substeps(int nMeshes, const MeshParamsDev* __restrict__ mp, int32_t* meshSubsteps) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m < nMeshes && mp[m].activeSolve) meshSubsteps[m] += 1;
}

// updateVelocities + dampVelocities
__global__ void __launch_bounds__(kBlock)
k_update_velocity(int n, const float4* __restrict__ pos, const float4* __restrict__ prev, float4* __restrict__ vel,
                  const int32_t* __restrict__ info, const MeshParamsDev* __restrict__ mp, StepParamsDev sp) {
    const int s = blockIdx.x * kBlock + threadIdx.x;
    if (s >= n) return;
    const int inf = __ldg(info + s);
    const MeshParamsDev* P = mp + (inf & 0x00FFFFFF);
    if (!P->activeSolve) return;
    const float3 x = f3(__ldg(pos + s)), xp = f3(__ldg(prev + s));
    float3 v = make_float3((x.x - xp.x) / sp.dt, (x.y - xp.y) / sp.dt, (x.z - xp.z) / sp.dt);
    const float vMag = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(v.x, v.x), __fmul_rn(v.y, v.y)), __fmul_rn(v.z, v.z)));
    const float maxV = P->maxVel;
    if (P->hasNoGravZone && maxV > 0.f && x.y > P->noGravThreshold) {
        if (vMag > maxV) { const float r = maxV / vMag; v = r * v; }
    } else if (vMag > 1e-6f) {
        float vNew = __fsub_rn(__fmul_rn(vMag, P->expVelDamp), P->constVelDamp);
        vNew = vNew > 1e-6f ? vNew : 0.f;
        const float r = vNew / vMag;
        v = r * v;
    } else {
        v = make_float3(0.f, 0.f, 0.f);
    }
    if (inf < 0) v = make_float3(0.f, 0.f, 0.f);
    vel[s] = make_float4(v.x, v.y, v.z, 0.f);
}

// Chebyshev: x <- x_{k-2} + omega (x - x_{k-2}) for non-colliding vertices; shift the history
__global__ void __launch_bounds__(kBlock)
k_accelerate(int n, float4* __restrict__ pos, float4* __restrict__ pp, float4* __restrict__ pIter, const int32_t* __restrict__ info,
             const MeshParamsDev* __restrict__ mp, const uint8_t* __restrict__ colliding, float omega, int init) {
    const int s = blockIdx.x * kBlock + threadIdx.x;
    if (s >= n) return;
    const int inf = __ldg(info + s);
    if (inf < 0 || !mp[inf & 0x00FFFFFF].activeSolve) return;
    float4 x = pos[s];
    if (init) { pp[s] = x; pIter[s] = x; return; }
    const float4 a = pp[s];
    if (!(colliding && colliding[s])) {
        x = make_float4(a.x + omega * (x.x - a.x), a.y + omega * (x.y - a.y), a.z + omega * (x.z - a.z), 0.f);
        pos[s] = x;
    }
    pp[s] = pIter[s];
    pIter[s] = x;
}

// DeformerRotator / DeformerTranslater (VBDDeformer.h:45-131, 133-177)
struct Rot9 { float r[9]; };   // row-major
__global__ void k_apply_rotator(int n, const int32_t* __restrict__ sid, const float4* __restrict__ root, const float4* __restrict__ rootP,
                                Rot9 R, float4* __restrict__ pos) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 o = root[i], q = rootP[i];
    const float x = __fadd_rn(__fadd_rn(__fmul_rn(R.r[0], q.x), __fmul_rn(R.r[1], q.y)), __fmul_rn(R.r[2], q.z));
    const float y = __fadd_rn(__fadd_rn(__fmul_rn(R.r[3], q.x), __fmul_rn(R.r[4], q.y)), __fmul_rn(R.r[5], q.z));
    const float z = __fadd_rn(__fadd_rn(__fmul_rn(R.r[6], q.x), __fmul_rn(R.r[7], q.y)), __fmul_rn(R.r[8], q.z));
    pos[sid[i]] = make_float4(__fadd_rn(o.x, x), __fadd_rn(o.y, y), __fadd_rn(o.z, z), 0.f);
}
__global__ void k_apply_translator(int n, const int32_t* __restrict__ sid, const float4* __restrict__ p0, float3 d, float4* __restrict__ pos) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 o = p0[i];
    pos[sid[i]] = make_float4(__fadd_rn(d.x, o.x), __fadd_rn(d.y, o.y), __fadd_rn(d.z, o.z), 0.f);
}

// sorted <-> original order transfers (xyz packed on the original side)
__global__ void k_gather_xyz(int64_t count, int64_t origBegin, const float4* __restrict__ src, const int32_t* __restrict__ o2s, float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const float4 v = src[o2s[origBegin + i]];
    out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z;
}
__global__ void k_scatter_xyz(int64_t count, int64_t origBegin, float4* __restrict__ dst, const int32_t* __restrict__ o2s, const float* __restrict__ in) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    dst[o2s[origBegin + i]] = make_float4(in[3 * i], in[3 * i + 1], in[3 * i + 2], 0.f);
}

static inline int nblk(int64_t n, int per = kBlock) { return (int)((n + per - 1) / per); }

Scene::~Scene() {
    halo_destroy(*this);
    collision_destroy(*this);
    for (Deformer* d : deformers) delete d;
    for (auto& r : hostRegs) cudaHostUnregister(r.first);
    if (pinnedIO) cudaFreeHost(pinnedIO);
    if (iterGraphExec) cudaGraphExecDestroy(iterGraphExec);
    for (cudaEvent_t e : profEv) if (e) cudaEventDestroy(e);
    if (evA) cudaEventDestroy(evA);
    if (evP0) cudaEventDestroy(evP0);
    if (evP1) cudaEventDestroy(evP1);
    if (evB) cudaEventDestroy(evB);
    if (stream) cudaStreamDestroy(stream);
}

static MeshParamsDev make_mesh_params(const HostMesh& m) {
    MeshParamsDev p;
    std::memset(&p, 0, sizeof(p));
    p.miu = m.obj.miu; p.lmbd = m.obj.lmbd; p.a = 1 + m.obj.miu / m.obj.lmbd;
    p.dampH = m.obj.dampingHydrostatic; p.dampD = m.obj.dampingDeviatoric;
    p.frictionDynamic = m.obj.frictionDynamic; p.frictionEpsV = m.obj.frictionEpsV;
    p.expVelDamp = m.obj.exponentialVelDamping; p.constVelDamp = m.obj.constantVelDamping;
    p.maxVel = m.obj.maxVelocityMagnitude; p.noGravThreshold = m.obj.noGravZoneThreshold;
    p.hasNoGravZone = m.obj.hasNoGravZone; p.initType = m.obj.initializationType; p.initRatio_g = m.obj.initRatio_g;
    p.activeSolve = m.activeForMaterialSolve; p.activeCollide = m.activeForCollision;
    p.vertBase = (int32_t)m.vertBase; p.tetBase = (int32_t)m.tetBase; p.surfVBase = (int32_t)m.surfVBase; p.surfFBase = (int32_t)m.surfFBase;
    return p;
}

void sync_mesh_params(Scene& s) {
    if (!s.meshParamsDirty) return;
    s.h_meshParams.resize(s.meshes.size());
    for (size_t i = 0; i < s.meshes.size(); i++) s.h_meshParams[i] = make_mesh_params(s.meshes[i]);
    cudaMemcpyAsync(s.meshParams.p, s.h_meshParams.data(), s.h_meshParams.size() * sizeof(MeshParamsDev), cudaMemcpyHostToDevice, s.stream);
    cudaStreamSynchronize(s.stream);
    s.meshParamsDirty = false;
    if (s.iterGraphExec) { cudaGraphExecDestroy(s.iterGraphExec); s.iterGraphExec = nullptr; }   // per-iteration accounting changed
}

cudaError_t device_upload(Scene& s) {
    const int64_t N = s.nVerts;
    CK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&s.evA)); CK(cudaEventCreate(&s.evB)); CK(cudaEventCreate(&s.evP0)); CK(cudaEventCreate(&s.evP1));
    CK(s.pos.alloc(N + 1)); CK(s.prev.alloc(N + 1)); CK(s.inertia.alloc(N + 1)); CK(s.vel.alloc(N + 1));
    CK(s.mass.alloc(N)); CK(s.info.alloc(N)); CK(s.incOff.alloc(N + 1));
    CK(s.incW.alloc(s.nInc)); CK(s.incIdx.alloc(s.nInc)); CK(s.nbOff.alloc(N + 1)); CK(s.nbList.alloc(s.nNb)); CK(s.vtx.alloc(N));
    CK(s.sorted2orig.alloc(N)); CK(s.orig2sorted.alloc(N));
    CK(s.meshParams.alloc(s.meshes.size())); CK(s.errFlags.alloc(8));
    CK(s.ioStage.alloc((size_t)(3 * N + 3) / 4 + 1));
    bool needVelPrev = false, needAccel = false;
    for (auto& m : s.meshes) {
        const int t = m.obj.initializationType;
        if (t == 2 || t == 4 || t == 6) needVelPrev = true;
        if (t == 6) needAccel = true;
    }
    if (needVelPrev) CK(s.velPrev.alloc(N + 1));
    if (needAccel) CK(s.accel.alloc(N + 1));
    if (s.phys.useAccelerator) { CK(s.accPrevPrev.alloc(N + 1)); CK(s.accPrevIter.alloc(N + 1)); }
    CK(cudaMemsetAsync(s.errFlags.p, 0, 8 * sizeof(int32_t), s.stream));
    CK(s.meshSubsteps.alloc(s.meshes.size()));
    CK(cudaMemsetAsync(s.meshSubsteps.p, 0, s.meshes.size() * sizeof(int32_t), s.stream));
    // host images in sorted order
    std::vector<float4> hpos(N + 1), hvel(N + 1);
    std::vector<float> hmass(N);
    std::vector<int32_t> hinfo(N), o2s(N);
    for (auto& m : s.meshes)
        for (int v = 0; v < m.nV; v++) {
            const int32_t sid = m.sortedId[v];
            hpos[sid] = make_float4(m.pos0[3 * (size_t)v], m.pos0[3 * (size_t)v + 1], m.pos0[3 * (size_t)v + 2], 0.f);
            hvel[sid] = make_float4(m.obj.initialVelocity[0], m.obj.initialVelocity[1], m.obj.initialVelocity[2], 0.f);
            hmass[sid] = m.mass[v];
            hinfo[sid] = (m.id & 0x00FFFFFF) | ((m.fixedMask[v] || m.ghostMask[v]) ? (int32_t)0x80000000 : 0);
            o2s[m.vertBase + v] = sid;
        }
    hpos[N] = make_float4(0, 0, 0, 0); hvel[N] = hpos[N];
    CK(cudaMemcpyAsync(s.pos.p, hpos.data(), (N + 1) * sizeof(float4), cudaMemcpyHostToDevice, s.stream));
    CK(cudaMemsetAsync(s.prev.p, 0, (N + 1) * sizeof(float4), s.stream));
    CK(cudaMemsetAsync(s.inertia.p, 0, (N + 1) * sizeof(float4), s.stream));
    CK(cudaMemcpyAsync(s.vel.p, hvel.data(), (N + 1) * sizeof(float4), cudaMemcpyHostToDevice, s.stream));
    if (needVelPrev) CK(cudaMemcpyAsync(s.velPrev.p, hvel.data(), (N + 1) * sizeof(float4), cudaMemcpyHostToDevice, s.stream));
    if (needAccel) CK(cudaMemsetAsync(s.accel.p, 0, (N + 1) * sizeof(float4), s.stream));
    CK(cudaMemcpyAsync(s.mass.p, hmass.data(), N * sizeof(float), cudaMemcpyHostToDevice, s.stream));
    CK(cudaMemcpyAsync(s.info.p, hinfo.data(), N * sizeof(int32_t), cudaMemcpyHostToDevice, s.stream));
    CK(cudaMemcpyAsync(s.incOff.p, s.h_incOff.data(), (N + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, s.stream));
    CK(cudaMemcpyAsync(s.incW.p, s.h_incW.data(), s.nInc * sizeof(float4), cudaMemcpyHostToDevice, s.stream));
    CK(cudaMemcpyAsync(s.incIdx.p, s.h_incIdx.data(), s.nInc * sizeof(uint32_t), cudaMemcpyHostToDevice, s.stream));
    std::vector<int4> hvtx(N);
    for (int64_t i = 0; i < N; i++)
        hvtx[i] = make_int4(s.h_incOff[i], s.h_nbOff[i], (s.h_incOff[i + 1] - s.h_incOff[i]) | ((s.h_nbOff[i + 1] - s.h_nbOff[i]) << 16), hinfo[i]);
    CK(cudaMemcpyAsync(s.vtx.p, hvtx.data(), N * sizeof(int4), cudaMemcpyHostToDevice, s.stream));
    CK(cudaMemcpyAsync(s.nbOff.p, s.h_nbOff.data(), (N + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, s.stream));
    CK(cudaMemcpyAsync(s.nbList.p, s.h_nbList.data(), s.nNb * sizeof(int32_t), cudaMemcpyHostToDevice, s.stream));
    CK(cudaMemcpyAsync(s.sorted2orig.p, s.sortedToOrig.data(), N * sizeof(int32_t), cudaMemcpyHostToDevice, s.stream));
    CK(cudaMemcpyAsync(s.orig2sorted.p, o2s.data(), N * sizeof(int32_t), cudaMemcpyHostToDevice, s.stream));
    CK(cudaStreamSynchronize(s.stream));
    {   // tile streams of the colour sweep: small colours get smaller tiles, so that a colour still spreads over the SMs
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sms > 0) s.numSMs = sms;
        int64_t largest = 0;
        for (size_t g = 0; g + 1 < s.groupStart.size(); g++) largest = std::max<int64_t>(largest, s.groupStart[g + 1] - s.groupStart[g]);
        if (const char* e = getenv("GAIA_B200_LEGACY_SWEEP")) s.legacySweep = atoi(e) != 0;      // (validation of the tile sweep only)
        if (const char* e = getenv("GAIA_B200_TILE_V")) s.tileVOverride = atoi(e);
        int tileV = kTileV;
        while (tileV > 8 && largest < (int64_t)s.numSMs * 3 * tileV) tileV >>= 1;
        if (!build_tile_streams(s, s.tileVOverride > 0 ? s.tileVOverride : tileV)) return cudaErrorInvalidValue;
        CK(sweep_upload(s));
    }
    std::vector<uint32_t>().swap(s.h_incIdx);
    std::vector<float>().swap(s.h_incW);
    std::vector<int32_t>().swap(s.h_nbList);
    // step constants
    StepParamsDev& sp = s.stepParams;
    sp.dt = s.dt; sp.dtSqrReciprocal = s.dtSqrReciprocal;
    for (int d = 0; d < 3; d++) { sp.gravity[d] = s.phys.gravity[d]; sp.boundsMin[d] = s.phys.worldBoundsMin[d]; sp.boundsMax[d] = s.phys.worldBoundsMax[d]; }
    sp.kBoundary = s.phys.boundaryCollisionStiffness; sp.muBoundary = s.phys.boundaryFrictionDynamic; sp.epsVBoundary = s.phys.boundaryFrictionEpsV;
    sp.stepSize = s.phys.stepSize; sp.stepSizeGD = s.phys.stepSizeGD; sp.kCollision = s.phys.collisionStiffness;
    sp.usePlaneGround = s.phys.usePlaneGround; sp.associateGravityWithInertia = s.phys.associateGravityWithInertia;
    s.pinnedIOBytes = (size_t)(3 * N + 16) * sizeof(float);
    CK(cudaMallocHost((void**)&s.pinnedIO, s.pinnedIOBytes));
    s.meshParamsDirty = true;
    sync_mesh_params(s);
    return cudaSuccess;
}


static int g_lanes = 4;     // sub-warp width of the colour solve (GAIA_B200_LANES = 2|4|8)
static int g_variant = 1;   // prefetch depth of the incidence stream (GAIA_B200_VARIANT = 0|1|2)

template <int L, bool HC, int PF, int ROUNDS, bool HALO = false>
static void launch_one(Scene& s, int vb, int ve, int applyFriction, const ContactView& cv, const HaloView* hvp = nullptr) {
    HaloView hv;
    if (hvp) hv = *hvp; else std::memset(&hv, 0, sizeof(hv));
    const int n = ve - vb;
    const int vpb = (kBlock / L) * ROUNDS;
    const size_t smem = (size_t)(kBlock / L) * kNbStride * sizeof(float4) + (size_t)10 * vpb * sizeof(float);
    static bool attrSet = false;
    if (!attrSet) { cudaFuncSetAttribute(k_solve_colour<L, HC, PF, ROUNDS, HALO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); attrSet = true; }
    k_solve_colour<L, HC, PF, ROUNDS, HALO><<<nblk(n, vpb), kBlock, smem, s.stream>>>(
        vb, ve, s.pos.p, s.prev.p, s.inertia.p, s.vtx.p, s.incW.p, s.incIdx.p, s.nbList.p,
        s.meshParams.p, s.stepParams, applyFriction, cv, s.errFlags.p, hv);
}

template <bool HC>
static void launch_solve(Scene& s, int L, int variant, int vb, int ve, int applyFriction, const ContactView& cv, const HaloView* hv = nullptr) {
    // rounds per block: big colours walk L rounds per block (full-width tail); smaller ones fewer rounds, so that the grid
    // still fills the 148 SMs x 4 resident blocks for a few waves
    const int64_t n = ve - vb;
    const int rounds = n >= (int64_t)148 * 8 * kBlock ? L : (n >= (int64_t)148 * 4 * kBlock ? 2 : 1);
    if (hv) {   // fused halo exchange: the default shape only (4 lanes, prefetch depth 1)
        const int r4 = n >= (int64_t)148 * 8 * kBlock ? 4 : (n >= (int64_t)148 * 4 * kBlock ? 2 : 1);
        if (r4 == 4) launch_one<4, HC, 1, 4, true>(s, vb, ve, applyFriction, cv, hv);
        else if (r4 == 2) launch_one<4, HC, 1, 2, true>(s, vb, ve, applyFriction, cv, hv);
        else launch_one<4, HC, 1, 1, true>(s, vb, ve, applyFriction, cv, hv);
        return;
    }
#define BYVAR(LL) if (rounds == LL) { if (variant == 0) launch_one<LL, HC, 0, LL>(s, vb, ve, applyFriction, cv); else launch_one<LL, HC, 1, LL>(s, vb, ve, applyFriction, cv); } \
                  else if (rounds == 2) { if (variant == 0) launch_one<LL, HC, 0, 2>(s, vb, ve, applyFriction, cv); else launch_one<LL, HC, 1, 2>(s, vb, ve, applyFriction, cv); } \
                  else { if (variant == 0) launch_one<LL, HC, 0, 1>(s, vb, ve, applyFriction, cv); else launch_one<LL, HC, 1, 1>(s, vb, ve, applyFriction, cv); }
    switch (L) {
        case 8: BYVAR(8); break;
        default: BYVAR(4); break;
    }
#undef BYVAR
}

static cudaError_t enqueue_iterations(Scene& s, int nIters, bool allowIntermediateCD, bool captured = false) {
    static bool envRead = false, lanesFromEnv = false;
    if (!envRead) {
        const char* e = getenv("GAIA_B200_LANES"); if (e) { g_lanes = atoi(e); lanesFromEnv = true; }
        e = getenv("GAIA_B200_VARIANT"); if (e) g_variant = atoi(e);
        envRead = true;
    }
    const int G = (int)s.groupStart.size() - 1;
    const int N = (int)s.nVerts;
    const bool halo = halo_active(s);
    static int fusedHalo = -1;
    if (fusedHalo < 0) { const char* e = getenv("GAIA_B200_HALO_FUSED"); fusedHalo = e ? atoi(e) : 1; }
    HaloView hv;
    // (the accelerator needs full refreshes; with contacts the per-colour pre-pass reads ghost positions BEFORE the sweep,
    //  so the exchange of the previous colour must be complete when it starts: the separate push + wait launch is used)
    const bool fused = halo && fusedHalo && !s.phys.useAccelerator && !s.cs && halo_view(s, hv);
    static int haloDebug = -1;      // timing experiments only: 1 = do not wait, 2 = do not store to peers (results are wrong)
    if (haloDebug < 0) { const char* e = getenv("GAIA_B200_HALO_DEBUG"); haloDebug = e ? atoi(e) : 0; }
    if (fused) hv.debug = haloDebug;
    static uint32_t* dbgCtl = nullptr;
    bool forceHalo = false;
    if (!halo && getenv("GAIA_B200_FORCE_HALO_KERNEL")) {      // single-GPU timing of the HALO kernel variant (no peers)
        static int32_t* dbgPrefix = nullptr;
        if (!dbgCtl) { cudaMalloc((void**)&dbgCtl, 16); cudaMemset(dbgCtl, 0, 16); cudaMalloc((void**)&dbgPrefix, (N + 1) * 4); cudaMemset(dbgPrefix, 0, (N + 1) * 4); }
        std::memset(&hv, 0, sizeof(hv)); hv.ctl = dbgCtl; hv.boundaryPrefix = dbgPrefix; forceHalo = true;
    }
    ContactView cv; std::memset(&cv, 0, sizeof(cv));
    const uint8_t* colliding = nullptr;
    // (captured: the contact variant runs every substep, with device-sized pre-passes; empty tables when nothing touches)
    bool hasContacts = collision_contact_view(s, cv, &colliding, captured);
    // per-iteration accounting (active, non-fixed vertices)
    int64_t solvedPerIter = 0, bytesPerIter = 0;
    for (auto& m : s.meshes) {
        if (!m.activeForMaterialSolve) continue;
        solvedPerIter += m.nSolved;
        int64_t nb = 0; for (auto c : m.nbrDistinctPerGroup) nb += c;
        bytesPerIter += 60 * m.nSolved + 56 * m.nIncSolved + 12 * nb;
    }
    float omega = 1.f;
    if (s.phys.useAccelerator) {
        k_accelerate<<<nblk(N), kBlock, 0, s.stream>>>(N, s.pos.p, s.accPrevPrev.p, s.accPrevIter.p, s.info.p, s.meshParams.p, nullptr, 1.f, 1);
        s.stats.kernelLaunches++;
    }
    for (int it = 0; it < nIters; it++) {
        const int applyFriction = it >= s.phys.frictionStartIter;
        if (s.phys.useAccelerator) {
            const int order = it + 1;
            const float rho = s.phys.acceleratorRho;
            if (order == 1) omega = 1.f;
            else if (order == 2) omega = 2 / (2 - rho * rho);
            else omega = 4.f / (4 - rho * rho * omega);
        }
        for (int g = 0; g < G; g++) {
            const int vb = s.groupStart[g], ve = s.groupStart[g + 1];
            if (ve <= vb && !halo) continue;
            if (hasContacts) { CK(collision_update_infos(s, g, captured)); }
            cudaEvent_t pe0 = nullptr, pe1 = nullptr;
            if (s.profiling) {      // a pair of events per launch, read back at the end of the loop: no host sync between launches
                if (s.profUsed + 2 > s.profEv.size()) { s.profEv.resize(s.profUsed + 2, nullptr); }
                for (size_t q = s.profUsed; q < s.profUsed + 2; q++) if (!s.profEv[q]) CK(cudaEventCreate(&s.profEv[q]));
                pe0 = s.profEv[s.profUsed]; pe1 = s.profEv[s.profUsed + 1]; s.profUsed += 2;
                cudaEventRecord(pe0, s.stream);
            }
            hv.epochOffset = it * G + g;
            if (ve > vb && !s.legacySweep) {
                CK(sweep_launch(s, g, s.solveLanes == 8 ? 8 : 4, applyFriction, hasContacts, cv, fused ? &hv : nullptr));
            } else if (ve > vb) {
                const int lanes = (s.solveLanes && !lanesFromEnv) ? s.solveLanes : g_lanes;
                if (hasContacts) launch_solve<true>(s, lanes, g_variant, vb, ve, applyFriction, cv, (fused || forceHalo) ? &hv : nullptr);
                else launch_solve<false>(s, lanes, g_variant, vb, ve, applyFriction, cv, (fused || forceHalo) ? &hv : nullptr);
            }
            if (s.profiling) cudaEventRecord(pe1, s.stream);
            s.stats.kernelLaunches++; s.stats.colourSweeps++;
            if (fused) { if (ve > vb) halo_count_fused(s, g); }    // pushed by the solve launch itself, published by the next one
            else if (halo) {
                // with contacts the pre-pass of this colour read ghost positions of EVERY colour: a neighbour that is ahead
                // must not overwrite them before this rank has finished it -> barrier among neighbours, then push + wait
                if (s.cs) CK(halo_sync(s));
                CK(halo_push(s, g));
            }
        }
        s.stats.vertexSolves += solvedPerIter;
        s.stats.algorithmicBytes += bytesPerIter + (captured ? 0 : 92 * (s.stats.numContactsDCD + s.stats.numContactsCCD));
        if (s.phys.useAccelerator) {
            k_accelerate<<<nblk(N), kBlock, 0, s.stream>>>(N, s.pos.p, s.accPrevPrev.p, s.accPrevIter.p, s.info.p, s.meshParams.p, colliding, omega, 0);
            s.stats.kernelLaunches++;
            if (halo) { CK(halo_sync(s)); CK(halo_push(s, -1)); }    // the extrapolated positions replace the ghosts' copies
        }
        const int ici = s.phys.intermediateCollisionIterations;
        if (allowIntermediateCD && ici > 0 && it % ici == ici - 1) {
            CK(collision_dcd(s, true));
            CK(collision_ccd(s));
            CK(collision_prepare(s));
            hasContacts = collision_contact_view(s, cv, &colliding);
        }
    }
    if (fused) { k_halo_advance<<<1, 32, 0, s.stream>>>(hv, nIters * G); s.stats.kernelLaunches++; }
    if (s.profiling && s.profUsed) {
        CK(cudaStreamSynchronize(s.stream));
        for (size_t q = 0; q + 1 < s.profUsed; q += 2) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, s.profEv[q], s.profEv[q + 1]) == cudaSuccess) { s.solveKernelMs += ms; s.solveKernelLaunches++; }
        }
        s.profUsed = 0;
    }
    return cudaGetLastError();
}

// The colour loop of one substep is launch-bound for small colours (hundreds of ~10 us kernels): when nothing in it needs
// the host (no intermediate detection, no per-kernel profiling, no separate halo launches) it is captured once into a CUDA
// graph and replayed. With the collision pipeline on, the contact pre-passes are launched over the capacity of the contact
// list and read this substep's counts on the device (collision_graphable: small scenes only).
static cudaError_t solve_iterations(Scene& s, int nIters, bool allowIntermediateCD) {
    static int useGraphs = -1;
    if (useGraphs < 0) { const char* e = getenv("GAIA_B200_GRAPH"); useGraphs = e ? atoi(e) : 1; }
    const bool withContacts = s.cs != nullptr;
    const bool contactsOk = !withContacts || (collision_graphable(s) && !halo_active(s) && !s.phys.useAccelerator &&
                                              !(allowIntermediateCD && s.phys.intermediateCollisionIterations > 0));
    const bool graphable = useGraphs && !s.profiling && contactsOk && s.iterCalls > 0 && nIters > 0;
    s.iterCalls++;
    if (!graphable) return enqueue_iterations(s, nIters, allowIntermediateCD);
    if (!s.iterGraphExec || s.iterGraphIters != nIters) {
        if (s.iterGraphExec) { cudaGraphExecDestroy(s.iterGraphExec); s.iterGraphExec = nullptr; }
        const gaia_stats before = s.stats;
        CK(cudaStreamBeginCapture(s.stream, cudaStreamCaptureModeThreadLocal));
        cudaError_t e = enqueue_iterations(s, nIters, allowIntermediateCD, withContacts);
        cudaGraph_t g = nullptr;
        cudaError_t e2 = cudaStreamEndCapture(s.stream, &g);
        if (e != cudaSuccess) return e;
        CK(e2);
        CK(cudaGraphInstantiate(&s.iterGraphExec, g, 0));
        cudaGraphDestroy(g);
        s.iterGraphIters = nIters;
        s.iterDelta[0] = s.stats.vertexSolves - before.vertexSolves; s.iterDelta[1] = s.stats.colourSweeps - before.colourSweeps;
        s.iterDelta[2] = s.stats.kernelLaunches - before.kernelLaunches; s.iterDelta[3] = s.stats.algorithmicBytes - before.algorithmicBytes;
        s.stats = before;
    }
    CK(cudaGraphLaunch(s.iterGraphExec, s.stream));
    s.stats.vertexSolves += s.iterDelta[0]; s.stats.colourSweeps += s.iterDelta[1];
    s.stats.kernelLaunches += s.iterDelta[2]; s.stats.algorithmicBytes += s.iterDelta[3];
    if (withContacts) s.stats.algorithmicBytes += (int64_t)nIters * 92 * (s.stats.numContactsDCD + s.stats.numContactsCCD);
    return cudaSuccess;
}

static cudaError_t initial_step(Scene& s) {
    const int N = (int)s.nVerts;
    k_initial_step<<<nblk(N), kBlock, 0, s.stream>>>(N, s.pos.p, s.prev.p, s.inertia.p, s.vel.p, s.velPrev.p, s.accel.p, s.info.p, s.mass.p,
                                                   s.meshParams.p, s.meshSubsteps.p, s.stepParams);
    k_count_substeps<<<nblk((int64_t)s.meshes.size(), 128), 128, 0, s.stream>>>((int)s.meshes.size(), s.meshParams.p, s.meshSubsteps.p);
    s.stats.kernelLaunches += 2;
    return cudaGetLastError();
}

static cudaError_t apply_deformers(Scene& s) {
    for (Deformer* d : s.deformers) {
        const int n = (int)d->verts.size();
        if (!n) continue;
        double t = (double)s.curTime;
        if (t >= d->endTime && d->endTime > 0) t = d->endTime;
        if (d->type == 0) {
            const double theta = t * d->angularVelocity;
            const double ux = d->axis[0], uy = d->axis[1], uz = d->axis[2];
            const double c = cos(theta), sn = sin(theta);
            Rot9 R;
            R.r[0] = (float)(c + ux * ux * (1 - c)); R.r[1] = (float)(ux * uy * (1 - c) - uz * sn); R.r[2] = (float)(ux * uz * (1 - c) + uy * sn);
            R.r[3] = (float)(uy * ux * (1 - c) + uz * sn); R.r[4] = (float)(c + uy * uy * (1 - c)); R.r[5] = (float)(uy * uz * (1 - c) - ux * sn);
            R.r[6] = (float)(uz * ux * (1 - c) - uy * sn); R.r[7] = (float)(uz * uy * (1 - c) + ux * sn); R.r[8] = (float)(c + uz * uz * (1 - c));
            k_apply_rotator<<<nblk(n), kBlock, 0, s.stream>>>(n, d->sid.p, d->root.p, d->rootP.p, R, s.pos.p);
        } else {
            // speed * curTime + originalPos: Eigen casts the double scalar to float first
            const float tf = (float)t;
            k_apply_translator<<<nblk(n), kBlock, 0, s.stream>>>(n, d->sid.p, d->root.p, make_float3(d->speed[0] * tf, d->speed[1] * tf, d->speed[2] * tf), s.pos.p);
        }
        s.stats.kernelLaunches++;
    }
    return cudaGetLastError();
}

cudaError_t upload_deformers(Scene& s) {
    for (Deformer* d : s.deformers) {
        const HostMesh& m = s.meshes[d->mesh];
        const int n = (int)d->verts.size();
        std::vector<int32_t> sid(n);
        std::vector<float4> a(n), b(n);
        for (int i = 0; i < n; i++) {
            const int v = d->verts[i];
            sid[i] = m.sortedId[v];
            const float P[3] = {m.pos0[3 * (size_t)v], m.pos0[3 * (size_t)v + 1], m.pos0[3 * (size_t)v + 2]};
            if (d->type == 0) {
                const float OP[3] = {P[0] - d->origin[0], P[1] - d->origin[1], P[2] - d->origin[2]};
                const float dt_ = OP[0] * d->axis[0] + OP[1] * d->axis[1] + OP[2] * d->axis[2];
                const float root[3] = {dt_ * d->axis[0], dt_ * d->axis[1], dt_ * d->axis[2]};
                a[i] = make_float4(root[0], root[1], root[2], 0.f);
                b[i] = make_float4(P[0] - root[0], P[1] - root[1], P[2] - root[2], 0.f);
            } else {
                a[i] = make_float4(P[0], P[1], P[2], 0.f); b[i] = a[i];
            }
        }
        CK(d->sid.alloc(n)); CK(d->root.alloc(n)); CK(d->rootP.alloc(n));
        CK(cudaMemcpy(d->sid.p, sid.data(), n * sizeof(int32_t), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d->root.p, a.data(), n * sizeof(float4), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(d->rootP.p, b.data(), n * sizeof(float4), cudaMemcpyHostToDevice));
    }
    return cudaSuccess;
}

// after the initial step every ghost must hold its owner's new position: barrier (nobody still reads the old ghosts),
// push all boundary vertices, wait for the neighbours' pushes
static cudaError_t halo_refresh(Scene& s) {
    if (!halo_active(s)) return cudaSuccess;
    CK(halo_sync(s));
    CK(halo_push(s, -1));     // push + publish epoch + wait for the neighbours' pushes, one launch
    return cudaSuccess;
}

// per-phase device time with the reference's timer names (M/Timer/RunningTimeStatistics.h), only while profiling
template <class F>
static cudaError_t timed_phase(Scene& s, double& acc, F&& fn) {
    if (!s.profiling) return fn();
    cudaEventRecord(s.evP0, s.stream);
    cudaError_t e = fn();
    cudaEventRecord(s.evP1, s.stream);
    cudaEventSynchronize(s.evP1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, s.evP0, s.evP1);
    acc += ms;
    return e;
}

cudaError_t run_substeps(Scene& s, int nSubsteps) {
    sync_mesh_params(s);
    const int N = (int)s.nVerts;
    for (s.substep = 0; s.substep < nSubsteps; s.substep++) {
        s.curTime += s.dt;
        CK(timed_phase(s, s.stats.timeCsmpColDetectDCD, [&] { return collision_dcd(s, false); }));
        CK(timed_phase(s, s.stats.timeCsmpInitialStep, [&] { cudaError_t e = initial_step(s); if (e == cudaSuccess) e = apply_deformers(s); if (e == cudaSuccess) e = halo_refresh(s); return e; }));
        CK(timed_phase(s, s.stats.timeCsmpColDetectCCD, [&] { cudaError_t e = collision_ccd(s); if (e == cudaSuccess) e = collision_prepare(s); return e; }));
        {
            const double solveBefore = s.solveKernelMs;
            double whole = 0.0;
            CK(timed_phase(s, whole, [&] { return solve_iterations(s, s.phys.iterations, true); }));
            s.stats.timeCsmpMaterialSolve += whole;      // colour loop incl. contact pre-passes (and intermediate detection)
            (void)solveBefore;
        }
        CK(timed_phase(s, s.stats.timeCsmpUpdateVelocity, [&] {
            k_update_velocity<<<nblk(N), kBlock, 0, s.stream>>>(N, s.pos.p, s.prev.p, s.vel.p, s.info.p, s.meshParams.p, s.stepParams);
            return cudaGetLastError();
        }));
        s.stats.kernelLaunches++;
        s.stats.substeps++;
    }
    s.stats.frames++;
    return cudaGetLastError();
}

cudaError_t run_iterations_only(Scene& s, int nIters) {
    sync_mesh_params(s);
    s.substep = 0;
    s.curTime += s.dt;
    CK(collision_dcd(s, false));
    CK(initial_step(s));
    CK(apply_deformers(s));
    CK(halo_refresh(s));
    CK(collision_ccd(s));
    CK(collision_prepare(s));
    CK(solve_iterations(s, nIters, false));
    return cudaGetLastError();
}

// host <-> pinned staging copies of tens of MB: split over a few threads (one memcpy thread moves ~8 GB/s, which would make
// the staging copy, not PCIe, the cost of get/set on a 10 M-tet scene)
static void par_memcpy(void* dst, const void* src, size_t bytes) {
    const size_t kMin = (size_t)2 << 20;
    unsigned hw = std::thread::hardware_concurrency();
    int nt = (int)std::min<size_t>(std::min<size_t>(8, hw ? hw : 1), bytes / kMin);
    if (nt <= 1) { std::memcpy(dst, src, bytes); return; }
    std::vector<std::thread> th;
    const size_t per = ((bytes / nt) + 63) & ~(size_t)63;
    for (int i = 0; i < nt; i++) {
        const size_t b = per * i, e = std::min(bytes, b + per);
        if (b >= e) break;
        th.emplace_back([=] { std::memcpy((char*)dst + b, (const char*)src + b, e - b); });
    }
    for (auto& t : th) t.join();
}

static bool host_registered(const Scene& s, const void* p, size_t bytes) {
    for (auto& r : s.hostRegs) if ((const char*)p >= r.first && (const char*)p + bytes <= r.first + r.second) return true;
    return false;
}

cudaError_t gather_to_host(Scene& s, const float4* devSorted, int64_t origBegin, int64_t count, float* outXYZ) {
    if (!count) return cudaSuccess;
    float* stage = (float*)s.ioStage.p;
    k_gather_xyz<<<nblk(count), kBlock, 0, s.stream>>>(count, origBegin, devSorted, s.orig2sorted.p, stage);
    if (host_registered(s, outXYZ, 3 * count * sizeof(float))) {
        CK(cudaMemcpyAsync(outXYZ, stage, 3 * count * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
        CK(cudaStreamSynchronize(s.stream));
    } else {
        CK(cudaMemcpyAsync(s.pinnedIO, stage, 3 * count * sizeof(float), cudaMemcpyDeviceToHost, s.stream));
        CK(cudaStreamSynchronize(s.stream));
        par_memcpy(outXYZ, s.pinnedIO, 3 * count * sizeof(float));
    }
    s.stats.kernelLaunches++;
    return cudaSuccess;
}

cudaError_t scatter_from_host(Scene& s, float4* devSorted, int64_t origBegin, int64_t count, const float* inXYZ, bool) {
    if (!count) return cudaSuccess;
    float* stage = (float*)s.ioStage.p;
    if (host_registered(s, inXYZ, 3 * count * sizeof(float))) {
        CK(cudaMemcpyAsync(stage, inXYZ, 3 * count * sizeof(float), cudaMemcpyHostToDevice, s.stream));
    } else {
        par_memcpy(s.pinnedIO, inXYZ, 3 * count * sizeof(float));
        CK(cudaMemcpyAsync(stage, s.pinnedIO, 3 * count * sizeof(float), cudaMemcpyHostToDevice, s.stream));
    }
    k_scatter_xyz<<<nblk(count), kBlock, 0, s.stream>>>(count, origBegin, devSorted, s.orig2sorted.p, stage);
    CK(cudaStreamSynchronize(s.stream));
    s.stats.kernelLaunches++;
    return cudaSuccess;
}

int check_device_errors(Scene& s) {
    int32_t h[8];
    if (cudaMemcpyAsync(h, s.errFlags.p, sizeof(h), cudaMemcpyDeviceToHost, s.stream) != cudaSuccess) return GAIA_ERR_CUDA;
    if (cudaStreamSynchronize(s.stream) != cudaSuccess) return GAIA_ERR_CUDA;
    if (h[0]) { s.err = "NaN met in the vertex solve"; return GAIA_ERR_NAN; }
    if (h[2]) { s.err = "contact list overflow"; return GAIA_ERR_OVERFLOW; }
    if (h[3]) { s.err = "tetrahedral traversal stack overflow"; return GAIA_ERR_OVERFLOW; }
    if (h[4]) { s.err = "halo exchange timed out waiting for a neighbour partition"; return GAIA_ERR_TIMEOUT; }
    if (h[5]) { s.err = "colour sweep: a bulk copy never completed"; return GAIA_ERR_TIMEOUT; }
    return GAIA_OK;
}

}  // namespace gb
