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
// This file: the step schedule (initial step, deformers, colour loop, velocity update), the elementwise kernels and the
// host <-> device transfers. The per-colour vertex solve itself lives in sweep.cu.
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
    for (int k = 0; k < 2; k++) {
        if (surfPinned[k]) cudaFreeHost(surfPinned[k]);
        if (surfGathered[k]) cudaEventDestroy(surfGathered[k]);
        if (surfCopied[k]) cudaEventDestroy(surfCopied[k]);
    }
    if (ioStream) cudaStreamDestroy(ioStream);
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

cudaError_t sync_mesh_params(Scene& s) {
    if (!s.meshParamsDirty) return cudaSuccess;
    s.h_meshParams.resize(s.meshes.size());
    for (size_t i = 0; i < s.meshes.size(); i++) s.h_meshParams[i] = make_mesh_params(s.meshes[i]);
    CK(cudaMemcpyAsync(s.meshParams.p, s.h_meshParams.data(), s.h_meshParams.size() * sizeof(MeshParamsDev), cudaMemcpyHostToDevice, s.stream));
    CK(cudaStreamSynchronize(s.stream));
    s.meshParamsDirty = false;
    if (s.iterGraphExec) { cudaGraphExecDestroy(s.iterGraphExec); s.iterGraphExec = nullptr; }   // per-iteration accounting changed
    return cudaSuccess;
}

cudaError_t device_upload(Scene& s) {
    const int64_t N = s.nVerts;
    CK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&s.evA)); CK(cudaEventCreate(&s.evB)); CK(cudaEventCreate(&s.evP0)); CK(cudaEventCreate(&s.evP1));
    CK(s.pos.alloc(N + 1)); CK(s.prev.alloc(N + 1)); CK(s.inertia.alloc(N + 1)); CK(s.vel.alloc(N + 1));
    CK(s.mass.alloc(N)); CK(s.info.alloc(N));
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
    CK(cudaMemcpyAsync(s.sorted2orig.p, s.sortedToOrig.data(), N * sizeof(int32_t), cudaMemcpyHostToDevice, s.stream));
    CK(cudaMemcpyAsync(s.orig2sorted.p, o2s.data(), N * sizeof(int32_t), cudaMemcpyHostToDevice, s.stream));
    CK(cudaStreamSynchronize(s.stream));
    {   // tile streams of the colour sweep: small colours get smaller tiles, so that a colour still spreads over the SMs
        int dev = 0, sms = 148;
        cudaGetDevice(&dev);
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sms > 0) s.numSMs = sms;
        int64_t largest = 0;
        for (size_t g = 0; g + 1 < s.groupStart.size(); g++) largest = std::max<int64_t>(largest, s.groupStart[g + 1] - s.groupStart[g]);
        int tileV = kTileV;
        while (tileV > 8 && largest < (int64_t)s.numSMs * 3 * tileV) tileV >>= 1;
        if (!build_tile_streams(s, tileV)) return cudaErrorInvalidValue;
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
    sp.useBowlGround = s.phys.useBowlGround; sp.bowlCap = s.bowl.cap; sp.bowlRadius = s.bowl.radius;
    for (int d = 0; d < 3; d++) sp.bowlCenter[d] = s.bowl.center[d];
    if (s.phys.useBowlGround) {      // the bowl term lives in the contact variant of the sweep only: it needs (empty) contact tables
        std::vector<int32_t> none(N, -1);
        CK(s.bowlNoContact.alloc(N)); CK(cudaMemcpy(s.bowlNoContact.p, none.data(), N * sizeof(int32_t), cudaMemcpyHostToDevice));
        CK(s.bowlZeroOff.alloc(N + 1)); CK(cudaMemset(s.bowlZeroOff.p, 0, (N + 1) * sizeof(int32_t)));
    }
    s.pinnedIOBytes = (size_t)(3 * N + 16) * sizeof(float);
    CK(cudaMallocHost((void**)&s.pinnedIO, s.pinnedIOBytes));
    s.meshParamsDirty = true;
    return sync_mesh_params(s);
}


static cudaError_t enqueue_iterations(Scene& s, int nIters, bool allowIntermediateCD, bool captured = false, bool graphShape = false) {
    const int G = (int)s.groupStart.size() - 1;
    const int N = (int)s.nVerts;
    const bool halo = halo_active(s);
    HaloView hv;
    std::memset(&hv, 0, sizeof(hv));
    // (the accelerator needs full refreshes; with contacts the per-colour pre-pass reads ghost positions BEFORE the sweep,
    //  so the exchange of the previous colour must be complete when it starts: the separate push + wait launch is used.
    //  A scene WITH a collision pipeline may still use the fused exchange while no rank holds an active contact - there
    //  is no pre-pass then - which the ranks agree on after every detection: halo_vote)
    const bool lineSearch = s.lineSearch.enabled;      // (its kernel does not push: separate halo launches)
    const bool bowl = s.phys.useBowlGround != 0;      // the bowl term is compiled into the contact variant of the sweep only (sweep.cu: vertex_tail)
    const bool canFuse = halo && !s.phys.useAccelerator && !lineSearch && !bowl && halo_view(s, hv);
    bool fused = canFuse && !s.cs;
    ContactView cv; std::memset(&cv, 0, sizeof(cv));
    const uint8_t* colliding = nullptr;
    // (captured: the contact variant runs every substep, with device-sized pre-passes; empty tables when nothing touches)
    bool hasContacts = collision_contact_view(s, cv, &colliding, captured);
    // the contact variant with empty tables when the bowl is on and nothing touches
    auto bowl_tables = [&]() { if (bowl && !hasContacts) { std::memset(&cv, 0, sizeof(cv)); cv.contactOfVertex = s.bowlNoContact.p; cv.relOff = s.bowlZeroOff.p; } };
    bowl_tables();
    auto agree = [&]() -> cudaError_t {
        if (!canFuse || !s.cs) return cudaSuccess;
        bool any = true;
        if (halo_can_vote(s)) CK(halo_vote(s, hasContacts, &any));
        fused = !any;
        return cudaSuccess;
    };
    CK(agree());
    int batch = 0;        // colour steps of the running fused batch (the epoch base on the device advances when it is closed)
    auto close_batch = [&]() -> cudaError_t {
        if (!batch) return cudaSuccess;
        k_halo_advance<<<1, 32, 0, s.stream>>>(hv, batch); s.stats.kernelLaunches++;
        batch = 0;
        // a detection (or the next substep's) reads ghost positions: the neighbours' last pushes must have landed
        if (s.cs) CK(halo_sync(s));
        return cudaGetLastError();
    };
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
    std::vector<int> pairSteps;      // profiling level 1: colour steps covered by each event pair
    auto prof_begin = [&](int steps) -> cudaError_t {      // a pair of events per launch, read back at the end of the loop: no host sync between launches
        if (s.profiling != 1) return cudaSuccess;
        if (s.profUsed + 2 > s.profEv.size()) { s.profEv.resize(s.profUsed + 2, nullptr); }
        for (size_t q = s.profUsed; q < s.profUsed + 2; q++) if (!s.profEv[q]) CK(cudaEventCreate(&s.profEv[q]));
        pairSteps.push_back(steps);
        return cudaEventRecord(s.profEv[s.profUsed], s.stream);
    };
    auto prof_end = [&]() { if (s.profiling == 1) { cudaEventRecord(s.profEv[s.profUsed + 1], s.stream); s.profUsed += 2; } };
    const int ici = s.phys.intermediateCollisionIterations;
    // (measured on the 10 M-tet beam, whole and cut in two: a replayed graph of per-colour launches is 1.5 - 6 % faster than the
    //  multi-colour launch, which in turn is 3 % faster than per-colour launches issued by the host. So: loops that replay
    //  from a graph keep one launch per colour; loops the host has to issue - intermediate detection cuts them into
    //  stretches - use one multi-colour launch per stretch)
    const bool useMulti = s.multiStepMode != 0 && !s.phys.useAccelerator && !lineSearch && !bowl && G > 0 && !(captured || graphShape) && sweep_multi_available(s);
    for (int it = 0; it < nIters; it++) {
        const int applyFriction = it >= s.phys.frictionStartIter;
        // Contact-free stretch of iterations (up to the next intermediate detection): ONE launch sweeps all its colours, with
        // a grid barrier between colours instead of a launch boundary (sweep.cu, MULTI)
        if (useMulti && !hasContacts && (!halo || fused)) {
            int nIt = nIters - it;
            if (allowIntermediateCD && ici > 0) nIt = std::min(nIt, ici - (it % ici));
            const int nSteps = nIt * G;
            CK(prof_begin(nSteps));
            hv.epochOffset = batch;
            if (fused) batch += nSteps;
            CK(sweep_launch_multi(s, 0, nSteps, std::max(0, s.phys.frictionStartIter - it) * G, fused ? &hv : nullptr));
            prof_end();
            s.stats.kernelLaunches++; s.stats.colourSweeps += nSteps;
            if (fused) for (int q = 0; q < nIt; q++) for (int g = 0; g < G; g++) if (s.groupStart[g + 1] > s.groupStart[g]) halo_count_fused(s, g);
            s.stats.vertexSolves += solvedPerIter * nIt;
            s.stats.algorithmicBytes += bytesPerIter * nIt;
            it += nIt - 1;
            if (allowIntermediateCD && ici > 0 && it % ici == ici - 1) {
                CK(close_batch());
                CK(collision_dcd(s, true));
                CK(collision_ccd(s));
                CK(collision_prepare(s));
                hasContacts = collision_contact_view(s, cv, &colliding);
                bowl_tables();
                CK(agree());
            }
            continue;
        }
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
            CK(prof_begin(1));
            hv.epochOffset = batch;
            if (fused) batch++;
            if (ve > vb) CK(lineSearch ? sweep_launch_linesearch(s, g, applyFriction, hasContacts || bowl, cv)
                                       : sweep_launch(s, g, applyFriction, hasContacts || bowl, cv, fused ? &hv : nullptr));
            prof_end();
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
        if (allowIntermediateCD && ici > 0 && it % ici == ici - 1) {
            CK(close_batch());
            CK(collision_dcd(s, true));
            CK(collision_ccd(s));
            CK(collision_prepare(s));
            hasContacts = collision_contact_view(s, cv, &colliding);
            bowl_tables();
            CK(agree());
        }
    }
    CK(close_batch());
    if (s.profiling == 1 && s.profUsed) {
        CK(cudaStreamSynchronize(s.stream));
        for (size_t q = 0; q + 1 < s.profUsed; q += 2) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, s.profEv[q], s.profEv[q + 1]) == cudaSuccess) { s.solveKernelMs += ms; s.solveKernelLaunches += pairSteps[q / 2]; }
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
    const bool withContacts = s.cs != nullptr;
    const bool contactsOk = !withContacts || (collision_graphable(s) && !halo_active(s) && !s.phys.useAccelerator &&
                                              !(allowIntermediateCD && s.phys.intermediateCollisionIterations > 0));
    // (mode 2 of gaia_b200_set_multi_colour_launch: never replay a graph of per-colour launches, always issue multi-colour launches)
    const bool graphable = s.profiling != 1 && contactsOk && s.iterCalls > 0 && nIters > 0 && s.multiStepMode != 2;
    s.iterCalls++;
    // (graphShape: this loop is replayed from a graph whenever profiling / the first call do not prevent it - those calls take
    //  the same launch shape, so that what the profiling pass times is what the timed pass runs)
    if (!graphable) return enqueue_iterations(s, nIters, allowIntermediateCD, false, contactsOk && nIters > 0 && s.multiStepMode != 2);
    if (!s.iterGraphExec || s.iterGraphIters != nIters) {
        if (s.iterGraphExec) { cudaGraphExecDestroy(s.iterGraphExec); s.iterGraphExec = nullptr; }
        const gaia_stats before = s.stats;
        CK(cudaStreamBeginCapture(s.stream, cudaStreamCaptureModeThreadLocal));
        cudaError_t e = enqueue_iterations(s, nIters, allowIntermediateCD, withContacts, true);
        cudaGraph_t g = nullptr;
        cudaError_t e2 = cudaStreamEndCapture(s.stream, &g);
        if (e != cudaSuccess) { if (g) cudaGraphDestroy(g); return e; }
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
    CK(sweep_symplectic_guess(s));      // initializationType 5: k_initial_step left those meshes where they were
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
    CK(sync_mesh_params(s));
    const int N = (int)s.nVerts;
    for (s.substep = 0; s.substep < nSubsteps; s.substep++) {
        s.curTime += s.dt;
        CK(timed_phase(s, s.stats.timeCsmpColDetectDCD, [&] { return collision_dcd(s, false); }));
        CK(timed_phase(s, s.stats.timeCsmpInitialStep, [&] { cudaError_t e = initial_step(s); if (e == cudaSuccess) e = apply_deformers(s); if (e == cudaSuccess) e = halo_refresh(s); return e; }));
        CK(timed_phase(s, s.stats.timeCsmpColDetectCCD, [&] { cudaError_t e = collision_ccd(s); if (e == cudaSuccess) e = collision_prepare(s); return e; }));
        {
            double whole = 0.0;
            CK(timed_phase(s, whole, [&] { return solve_iterations(s, s.phys.iterations, true); }));
            s.stats.timeCsmpMaterialSolve += whole;      // colour loop incl. contact pre-passes (and intermediate detection)
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
    CK(sync_mesh_params(s));
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
    if (h[5]) { s.err = "colour sweep: a bulk copy or the barrier between colours never completed"; return GAIA_ERR_TIMEOUT; }
    return GAIA_OK;
}

}  // namespace gb
