// C ABI of libgaia_b200.so (include/gaia_b200.h): argument checking, error text, call ordering.
// No CPU path exists behind these entry points: create() fails without a CUDA device.
#include "scene_impl.h"
#include <cmath>
#include <cstring>
#include <string>

using namespace gb;

struct gaia_handle { Scene s; };

static thread_local std::string g_createErr;

static int fail(Scene& s, int code, const std::string& msg) { s.err = msg; return code; }
static int cuda_fail(Scene& s, cudaError_t e) { s.err = std::string("CUDA: ") + cudaGetErrorString(e); return GAIA_ERR_CUDA; }
#define REQUIRE(h) if (!(h)) return GAIA_ERR_INVALID_ARG
#define FINALIZED(h) if (!(h)->s.finalized) return fail((h)->s, GAIA_ERR_STATE, "scene not finalized")

extern "C" {

int gaia_b200_abi_version(void) { return GAIA_B200_ABI_VERSION; }
const char* gaia_b200_last_error(const gaia_handle* h) { return h ? h->s.err.c_str() : g_createErr.c_str(); }

void gaia_b200_default_physics_params(gaia_physics_params* p) {
    std::memset(p, 0, sizeof(*p));
    p->gravity[1] = -10.f; p->timeStep = 0.01666666666666666666f; p->numSubsteps = 20; p->iterations = 1;
    p->usePlaneGround = 1;
    for (int d = 0; d < 3; d++) { p->worldBoundsMin[d] = -1e7f; p->worldBoundsMax[d] = 1e7f; }
    p->boundaryFrictionDynamic = 0.1f;
    p->dcdTetMeshSceneBVHRebuildSteps = 16; p->dcdSurfaceSceneBVHRebuildSteps = 3; p->ccdBVHRebuildSteps = 7;
    p->stepSize = 1.0f; p->stepSizeGD = 1e-3f; p->associateGravityWithInertia = 1;
    p->collisionStiffness = 1e5f; p->intermediateCollisionIterations = -1;
    p->boundaryCollisionStiffness = 1e5f; p->boundaryFrictionEpsV = 1.0f; p->frictionStartIter = 0;
    p->useAccelerator = 0; p->acceleratorRho = 0.93f; p->collisionOffHeight = 1e7f;
}
void gaia_b200_default_collision_params(gaia_collision_params* c) {
    std::memset(c, 0, sizeof(*c));
    c->allowCCD = 1; c->allowDCD = 1; c->checkFeasibleRegion = 1; c->checkTetTraverse = 1; c->handleSelfCollision = 1;
    c->stopTraversingAfterPassingQueryPoint = 1; c->tetrahedralTraverseForNonSelfIntersection = 1; c->useStaticTraverse = 1;
    c->shiftQueryPointToCenter = 1; c->centerShiftLevel = 0.01f; c->maxNumberOfBVHQuery = 500;
    c->feasibleRegionEpsilon = 1e-2f; c->rayTriIntersectionEPSILON = 1e-10f; c->maxSearchDistanceMultiplier = 1.8f;
}
void gaia_b200_default_object_params(gaia_object_params* o) {
    std::memset(o, 0, sizeof(*o));
    for (int d = 0; d < 3; d++) o->scale[d] = 1.f;
    o->density = 1.f; o->maxVelocityMagnitude = -1.f; o->frameToAppear = -1;
    o->frictionDynamic = 0.1f; o->frictionEpsV = 0.01f; o->exponentialVelDamping = 1.0f; o->initRatio_g = 0.5f;
    o->miu = 1e4f; o->lmbd = 1e5f;
}

int gaia_b200_create(const gaia_physics_params* phys, const gaia_collision_params* col, int device, gaia_handle** out) {
    if (!phys || !col || !out) { g_createErr = "null argument"; return GAIA_ERR_INVALID_ARG; }
    *out = nullptr;
    int nDev = 0;
    cudaError_t e = cudaGetDeviceCount(&nDev);
    if (e != cudaSuccess || nDev <= 0) {
        g_createErr = std::string("no CUDA device available (") + cudaGetErrorString(e) + "); gaia_b200 has no CPU path";
        return GAIA_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= nDev) { g_createErr = "device ordinal out of range"; return GAIA_ERR_INVALID_ARG; }
    if (phys->numSubsteps <= 0 || phys->iterations < 0 || !(phys->timeStep > 0)) { g_createErr = "bad step parameters"; return GAIA_ERR_INVALID_ARG; }
    if ((e = cudaSetDevice(device)) != cudaSuccess) { g_createErr = cudaGetErrorString(e); return GAIA_ERR_CUDA; }
    gaia_handle* h = new gaia_handle();
    Scene& s = h->s;
    s.phys = *phys; s.col = *col; s.device = device;
    s.dt = s.phys.timeStep / s.phys.numSubsteps;                 // BasePhysicFramework::timeStepEqualize
    s.dtSqrReciprocal = 1.f / (s.dt * s.dt);
    std::memset(&s.stats, 0, sizeof(s.stats));
    *out = h;
    return GAIA_OK;
}

void gaia_b200_destroy(gaia_handle* h) {
    if (!h) return;
    cudaSetDevice(h->s.device);
    if (h->s.stream) cudaStreamSynchronize(h->s.stream);
    delete h;
}

int gaia_b200_add_tet_mesh(gaia_handle* h, int32_t nV, int32_t nT, const float* xyz, const int32_t* tets, const gaia_object_params* obj,
                           int32_t nColors, const int32_t* color_offsets, const int32_t* color_verts, int32_t nFixed, const int32_t* fixedPoints) {
    REQUIRE(h);
    Scene& s = h->s;
    if (s.finalized) return fail(s, GAIA_ERR_STATE, "add_tet_mesh after finalize");
    if (nV <= 0 || nT <= 0 || !xyz || !tets || !obj || nColors <= 0 || !color_offsets || !color_verts || nFixed < 0 || (nFixed && !fixedPoints))
        return fail(s, GAIA_ERR_INVALID_ARG, "add_tet_mesh: bad argument");
    if (s.meshes.size() >= (1u << 24) - 1) return fail(s, GAIA_ERR_INVALID_ARG, "too many meshes");
    for (int64_t i = 0; i < 4 * (int64_t)nT; i++) if (tets[i] < 0 || tets[i] >= nV) return fail(s, GAIA_ERR_INVALID_ARG, "tet index out of range");
    for (int i = 0; i < nFixed; i++) if (fixedPoints[i] < 0 || fixedPoints[i] >= nV) return fail(s, GAIA_ERR_INVALID_ARG, "fixed point out of range");
    s.meshes.emplace_back();
    HostMesh& m = s.meshes.back();
    m.id = (int)s.meshes.size() - 1;
    gaia_object_params o = *obj;
    o.exponentialVelDamping = std::pow(o.exponentialVelDamping, s.dt);   // VBDPhysics::timeStepEqualize
    o.constantVelDamping = o.constantVelDamping * s.dt;
    host_mesh_init(m, nV, nT, xyz, tets, o, nColors, color_offsets, color_verts, nFixed, fixedPoints, s.partitionMode);
    return m.id;
}

static int add_deformer(gaia_handle* h, int type, int mesh, const int32_t* verts, int32_t n, const float* origin, const float* axis,
                        const float* speed, double w, double endTime) {
    REQUIRE(h);
    Scene& s = h->s;
    if (s.finalized) return fail(s, GAIA_ERR_STATE, "deformers must be added before finalize");
    if (mesh < 0 || mesh >= (int)s.meshes.size() || n < 0 || (n && !verts)) return fail(s, GAIA_ERR_INVALID_ARG, "bad deformer argument");
    HostMesh& m = s.meshes[mesh];
    for (int i = 0; i < n; i++) if (verts[i] < 0 || verts[i] >= m.nV) return fail(s, GAIA_ERR_INVALID_ARG, "deformer vertex out of range");
    Deformer* d = new Deformer();
    d->type = type; d->mesh = mesh; d->verts.assign(verts, verts + n); d->angularVelocity = w; d->endTime = endTime;
    for (int k = 0; k < 3; k++) { if (origin) d->origin[k] = origin[k]; if (axis) d->axis[k] = axis[k]; if (speed) d->speed[k] = speed[k]; }
    for (int i = 0; i < n; i++)                              // setFixed (VBDDeformer.h:20-43)
        if (!m.fixedMask[verts[i]]) { m.fixedMask[verts[i]] = 1; m.fixedPoints.push_back(verts[i]); }
    s.deformers.push_back(d);
    return GAIA_OK;
}
int gaia_b200_add_rotator(gaia_handle* h, int mesh, const int32_t* verts, int32_t n, const float origin[3], const float axis[3],
                          double angularVelocity, double rotationEndTime) {
    if (!origin || !axis) return GAIA_ERR_INVALID_ARG;
    return add_deformer(h, 0, mesh, verts, n, origin, axis, nullptr, angularVelocity, rotationEndTime);
}
int gaia_b200_add_translator(gaia_handle* h, int mesh, const int32_t* verts, int32_t n, const float speed[3], double deformationEndTime) {
    if (!speed) return GAIA_ERR_INVALID_ARG;
    return add_deformer(h, 1, mesh, verts, n, nullptr, nullptr, speed, 0.0, deformationEndTime);
}

int gaia_b200_finalize(gaia_handle* h) {
    REQUIRE(h);
    Scene& s = h->s;
    if (s.finalized) return fail(s, GAIA_ERR_STATE, "already finalized");
    if (s.meshes.empty()) return fail(s, GAIA_ERR_STATE, "no meshes");
    cudaSetDevice(s.device);
    if (!host_finalize(s)) return GAIA_ERR_INVALID_ARG;
    cudaError_t e = device_upload(s);
    if (e != cudaSuccess) return cuda_fail(s, e);
    e = upload_deformers(s);
    if (e != cudaSuccess) return cuda_fail(s, e);
    e = collision_init(s);
    if (e != cudaSuccess) return cuda_fail(s, e);
    for (auto& m : s.meshes)
        if (m.obj.initializationType == 5 && !s.symScratch.p) {      // symplectic-Euler initial guess: per-tet tables + a Jacobi target
            e = linesearch_upload(s);
            if (e == cudaSuccess) e = s.symScratch.alloc((size_t)s.nVerts);
            if (e != cudaSuccess) return cuda_fail(s, e);
        }
    s.finalized = true;
    return GAIA_OK;
}

static void enable_models(Scene& s) {   // VBDPhysics::enableModels (VBDPhysics.cpp:685-709): frameToAppear <= frameId
    for (auto& m : s.meshes)
        if (m.obj.frameToAppear <= s.frameId && (!m.activeForCollision || !m.activeForMaterialSolve)) {
            m.activeForCollision = true; m.activeForMaterialSolve = true; s.meshParamsDirty = true;
        }
}

int gaia_b200_run_frames(gaia_handle* h, int32_t nFrames) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    cudaSetDevice(s.device);
    for (int f = 0; f < nFrames; f++) {
        enable_models(s);
        cudaError_t e = run_substeps(s, s.phys.numSubsteps);
        if (e != cudaSuccess) return cuda_fail(s, e);
        ++s.frameId;
    }
    return GAIA_OK;
}

int gaia_b200_run_iterations(gaia_handle* h, int32_t nIters) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    cudaSetDevice(s.device);
    enable_models(s);
    cudaError_t e = run_iterations_only(s, nIters);
    if (e != cudaSuccess) return cuda_fail(s, e);
    return GAIA_OK;
}

int gaia_b200_synchronize(gaia_handle* h) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    cudaSetDevice(s.device);
    cudaError_t e = cudaStreamSynchronize(s.stream);
    if (e != cudaSuccess) return cuda_fail(s, e);
    return check_device_errors(s);
}
int gaia_b200_frame_id(const gaia_handle* h) { return h ? h->s.frameId : GAIA_ERR_INVALID_ARG; }
int gaia_b200_set_time(gaia_handle* h, int32_t frameId, double curTime) {
    REQUIRE(h);
    if (frameId < 0) return fail(h->s, GAIA_ERR_INVALID_ARG, "negative frame id");
    h->s.frameId = frameId; h->s.curTime = (float)curTime;
    collision_invalidate(h->s);
    return GAIA_OK;
}
double gaia_b200_cur_time(const gaia_handle* h) { return h ? (double)h->s.curTime : 0.0; }
int gaia_b200_num_meshes(const gaia_handle* h) { return h ? (int)h->s.meshes.size() : GAIA_ERR_INVALID_ARG; }
int gaia_b200_mesh_sizes(const gaia_handle* h, int mesh, int32_t* nV, int32_t* nT, int32_t* nSurfV, int32_t* nSurfF) {
    REQUIRE(h);
    if (mesh < 0 || mesh >= (int)h->s.meshes.size()) return GAIA_ERR_INVALID_ARG;
    const HostMesh& m = h->s.meshes[mesh];
    if (nV) *nV = m.nV; if (nT) *nT = m.nT; if (nSurfV) *nSurfV = m.nSurfV(); if (nSurfF) *nSurfF = m.nSurfF();
    return GAIA_OK;
}

static int get_field(gaia_handle* h, int mesh, const float4* buf, float* xyz) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    if (mesh < 0 || mesh >= (int)s.meshes.size() || !xyz) return fail(s, GAIA_ERR_INVALID_ARG, "bad mesh id / null output");
    cudaSetDevice(s.device);
    cudaError_t e = gather_to_host(s, buf, s.meshes[mesh].vertBase, s.meshes[mesh].nV, xyz);
    if (e != cudaSuccess) return cuda_fail(s, e);
    return check_device_errors(s);
}
int gaia_b200_get_positions(gaia_handle* h, int mesh, float* xyz) { return get_field(h, mesh, h ? h->s.pos.p : nullptr, xyz); }
int gaia_b200_get_prev_positions(gaia_handle* h, int mesh, float* xyz) { return get_field(h, mesh, h ? h->s.prev.p : nullptr, xyz); }
int gaia_b200_get_velocities(gaia_handle* h, int mesh, float* xyz) { return get_field(h, mesh, h ? h->s.vel.p : nullptr, xyz); }
int gaia_b200_get_inertia(gaia_handle* h, int mesh, float* xyz) { return get_field(h, mesh, h ? h->s.inertia.p : nullptr, xyz); }

int gaia_b200_set_state(gaia_handle* h, int mesh, const float* pos, const float* vel) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    if (mesh < 0 || mesh >= (int)s.meshes.size()) return fail(s, GAIA_ERR_INVALID_ARG, "bad mesh id");
    cudaSetDevice(s.device);
    const HostMesh& m = s.meshes[mesh];
    cudaError_t e;
    if (pos && (e = scatter_from_host(s, s.pos.p, m.vertBase, m.nV, pos, false)) != cudaSuccess) return cuda_fail(s, e);
    if (vel && (e = scatter_from_host(s, s.vel.p, m.vertBase, m.nV, vel, false)) != cudaSuccess) return cuda_fail(s, e);
    return GAIA_OK;
}

// ---- surface output: gathered on the device (M/IO/FileIO.cpp:52-83 writes surface vertices in ascending tet-vertex id)
static __global__ void k_gather_surface(int64_t n, const float4* __restrict__ pos, const int32_t* __restrict__ surfSid, float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 v = pos[surfSid[i]];
    out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z;
}
static cudaError_t surface_io_init(Scene& s) {
    if (s.ioStream) return cudaSuccess;
    const int64_t nS = s.nSurfVerts;
    cudaError_t e;
    if ((e = cudaStreamCreateWithFlags(&s.ioStream, cudaStreamNonBlocking)) != cudaSuccess) return e;
    std::vector<int32_t> sid((size_t)std::max<int64_t>(nS, 1));
    for (auto& m : s.meshes) for (int i = 0; i < m.nSurfV(); i++) sid[m.surfVBase + i] = m.sortedId[m.surfaceVIds[i]];
    if ((e = s.surfSidAll.alloc(sid.size())) != cudaSuccess) return e;
    if ((e = cudaMemcpy(s.surfSidAll.p, sid.data(), sid.size() * sizeof(int32_t), cudaMemcpyHostToDevice)) != cudaSuccess) return e;
    for (int k = 0; k < 2; k++) {
        if ((e = s.surfStage[k].alloc(3 * sid.size())) != cudaSuccess) return e;
        if ((e = cudaMallocHost((void**)&s.surfPinned[k], 3 * sid.size() * sizeof(float))) != cudaSuccess) return e;
        if ((e = cudaEventCreateWithFlags(&s.surfGathered[k], cudaEventDisableTiming)) != cudaSuccess) return e;
        if ((e = cudaEventCreateWithFlags(&s.surfCopied[k], cudaEventDisableTiming)) != cudaSuccess) return e;
    }
    return cudaSuccess;
}
static cudaError_t surface_snapshot(Scene& s, int slot, int64_t first, int64_t count) {
    cudaError_t e = surface_io_init(s);
    if (e != cudaSuccess) return e;
    if ((e = cudaStreamWaitEvent(s.stream, s.surfCopied[slot], 0)) != cudaSuccess) return e;      // the slot's previous copy has left the staging buffer
    if (count > 0) {
        k_gather_surface<<<(unsigned)((count + 255) / 256), 256, 0, s.stream>>>(count, s.pos.p, s.surfSidAll.p + first, s.surfStage[slot].p + 3 * first);
        s.stats.kernelLaunches++;
    }
    if ((e = cudaEventRecord(s.surfGathered[slot], s.stream)) != cudaSuccess) return e;
    if ((e = cudaStreamWaitEvent(s.ioStream, s.surfGathered[slot], 0)) != cudaSuccess) return e;
    if (count > 0 && (e = cudaMemcpyAsync(s.surfPinned[slot] + 3 * first, s.surfStage[slot].p + 3 * first, 3 * (size_t)count * sizeof(float),
                                          cudaMemcpyDeviceToHost, s.ioStream)) != cudaSuccess) return e;
    return cudaEventRecord(s.surfCopied[slot], s.ioStream);
}

int gaia_b200_get_surface_positions(gaia_handle* h, int mesh, float* xyz) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    if (mesh < 0 || mesh >= (int)s.meshes.size() || !xyz) return fail(s, GAIA_ERR_INVALID_ARG, "bad mesh id / null output");
    cudaSetDevice(s.device);
    const HostMesh& m = s.meshes[mesh];
    cudaError_t e = surface_snapshot(s, 0, m.surfVBase, m.nSurfV());
    if (e == cudaSuccess) e = cudaEventSynchronize(s.surfCopied[0]);
    if (e != cudaSuccess) return cuda_fail(s, e);
    std::memcpy(xyz, s.surfPinned[0] + 3 * m.surfVBase, 3 * (size_t)m.nSurfV() * sizeof(float));
    return GAIA_OK;
}
int64_t gaia_b200_total_surface_vertices(const gaia_handle* h) { return h ? h->s.nSurfVerts : GAIA_ERR_INVALID_ARG; }
int gaia_b200_surface_snapshot_begin(gaia_handle* h, int slot) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    if (slot < 0 || slot > 1) return fail(s, GAIA_ERR_INVALID_ARG, "snapshot slot: 0 or 1");
    cudaSetDevice(s.device);
    const cudaError_t e = surface_snapshot(s, slot, 0, s.nSurfVerts);
    return e == cudaSuccess ? GAIA_OK : cuda_fail(s, e);
}
int gaia_b200_surface_snapshot_wait(gaia_handle* h, int slot, const float** xyz, int64_t* nFloats) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    if (slot < 0 || slot > 1 || !xyz || !s.ioStream) return fail(s, GAIA_ERR_INVALID_ARG, "snapshot slot not begun / null output");
    cudaSetDevice(s.device);      // (may be a writer thread: the current device is per thread)
    const cudaError_t e = cudaEventSynchronize(s.surfCopied[slot]);
    if (e != cudaSuccess) return cuda_fail(s, e);
    *xyz = s.surfPinned[slot];
    if (nFloats) *nFloats = 3 * s.nSurfVerts;
    return GAIA_OK;
}

int64_t gaia_b200_total_vertices(const gaia_handle* h) { return h ? h->s.nVerts : GAIA_ERR_INVALID_ARG; }
int gaia_b200_get_all_positions(gaia_handle* h, float* xyz) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    if (!xyz) return fail(s, GAIA_ERR_INVALID_ARG, "null output");
    cudaSetDevice(s.device);
    cudaError_t e = gather_to_host(s, s.pos.p, 0, s.nVerts, xyz);
    if (e != cudaSuccess) return cuda_fail(s, e);
    return check_device_errors(s);
}
int gaia_b200_set_all_state(gaia_handle* h, const float* pos, const float* vel) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    cudaSetDevice(s.device);
    cudaError_t e;
    if (pos && (e = scatter_from_host(s, s.pos.p, 0, s.nVerts, pos, false)) != cudaSuccess) return cuda_fail(s, e);
    if (vel && (e = scatter_from_host(s, s.vel.p, 0, s.nVerts, vel, false)) != cudaSuccess) return cuda_fail(s, e);
    return GAIA_OK;
}

int gaia_b200_register_host_buffer(gaia_handle* h, void* ptr, int64_t bytes) {
    REQUIRE(h);
    Scene& s = h->s;
    if (!ptr || bytes <= 0) return fail(s, GAIA_ERR_INVALID_ARG, "bad host buffer");
    cudaSetDevice(s.device);
    cudaError_t e = cudaHostRegister(ptr, (size_t)bytes, cudaHostRegisterDefault);
    if (e != cudaSuccess) return cuda_fail(s, e);
    s.hostRegs.push_back({(char*)ptr, (size_t)bytes});
    return GAIA_OK;
}
int gaia_b200_unregister_host_buffer(gaia_handle* h, void* ptr) {
    REQUIRE(h);
    Scene& s = h->s;
    for (size_t i = 0; i < s.hostRegs.size(); i++)
        if (s.hostRegs[i].first == (char*)ptr) {
            if (s.stream) cudaStreamSynchronize(s.stream);
            cudaHostUnregister(ptr);
            s.hostRegs.erase(s.hostRegs.begin() + i);
            return GAIA_OK;
        }
    return fail(s, GAIA_ERR_INVALID_ARG, "buffer was not registered");
}

int gaia_b200_get_world_bounds(const gaia_handle* h, float* mn, float* mx) {
    REQUIRE(h);
    for (int d = 0; d < 3; d++) { mn[d] = h->s.phys.worldBoundsMin[d]; mx[d] = h->s.phys.worldBoundsMax[d]; }
    return GAIA_OK;
}
int gaia_b200_num_parallel_groups(const gaia_handle* h) { return h ? (int)h->s.groups.size() : GAIA_ERR_INVALID_ARG; }
int gaia_b200_get_parallel_group(const gaia_handle* h, int group, int32_t* out, int32_t* nPairs) {
    REQUIRE(h);
    if (group < 0 || group >= (int)h->s.groups.size() || !nPairs) return GAIA_ERR_INVALID_ARG;
    const auto& g = h->s.groups[group];
    *nPairs = (int32_t)g.size() / 2;
    if (out) std::memcpy(out, g.data(), g.size() * sizeof(int32_t));
    return GAIA_OK;
}
#define MESH_ARRAY_GETTER(name, field, T)                                                  \
    int name(const gaia_handle* h, int mesh, T* out) {                                     \
        REQUIRE(h);                                                                        \
        if (mesh < 0 || mesh >= (int)h->s.meshes.size() || !out) return GAIA_ERR_INVALID_ARG; \
        const auto& v = h->s.meshes[mesh].field;                                           \
        std::memcpy(out, v.data(), v.size() * sizeof(T));                                  \
        return GAIA_OK;                                                                    \
    }
MESH_ARRAY_GETTER(gaia_b200_get_surface_vertex_ids, surfaceVIds, int32_t)
MESH_ARRAY_GETTER(gaia_b200_get_surface_faces, surfFaces, int32_t)
MESH_ARRAY_GETTER(gaia_b200_get_vertex_masses, mass, float)

int gaia_b200_detect_collisions(gaia_handle* h, int doDCD, int doCCD) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    cudaSetDevice(s.device);
    { const cudaError_t e0 = sync_mesh_params(s); if (e0 != cudaSuccess) return cuda_fail(s, e0); }
    const gaia_collision_params saved = s.col;
    s.col.allowDCD = saved.allowDCD && doDCD; s.col.allowCCD = saved.allowCCD && doCCD;
    s.substep = 0;
    cudaError_t e = collision_dcd(s, false);
    if (e == cudaSuccess) e = collision_ccd(s);
    if (e == cudaSuccess) e = collision_prepare(s);
    s.col = saved;
    if (e != cudaSuccess) return cuda_fail(s, e);
    return check_device_errors(s);
}
int gaia_b200_get_contacts(gaia_handle* h, gaia_contact* out, int64_t capacity, int64_t* n) {
    REQUIRE(h); FINALIZED(h);
    if (!n) return GAIA_ERR_INVALID_ARG;
    cudaSetDevice(h->s.device);
    *n = collision_fetch_contacts(h->s, out, capacity);
    return *n < 0 ? GAIA_ERR_CUDA : GAIA_OK;
}
int gaia_b200_get_bvh_leaf_order(gaia_handle* h, int which, int32_t* out, int64_t capacity, int64_t* n) {
    REQUIRE(h); FINALIZED(h);
    if (!n) return GAIA_ERR_INVALID_ARG;
    cudaSetDevice(h->s.device);
    *n = collision_leaf_order(h->s, which, out, capacity);
    return *n < 0 ? GAIA_ERR_INVALID_ARG : GAIA_OK;
}

int gaia_b200_get_active_list(gaia_handle* h, int group, int32_t* out, int64_t capacity, int64_t* n) {
    REQUIRE(h); FINALIZED(h);
    if (!n) return GAIA_ERR_INVALID_ARG;
    if (group < 0 || group >= (int)h->s.groupStart.size() - 1) return fail(h->s, GAIA_ERR_INVALID_ARG, "bad parallel group");
    cudaSetDevice(h->s.device);
    *n = collision_fetch_active_list(h->s, group, out, capacity);
    return *n < 0 ? GAIA_ERR_CUDA : GAIA_OK;
}
int gaia_b200_get_relations(gaia_handle* h, int mesh, int32_t* out, int64_t capacity, int64_t* n) {
    REQUIRE(h); FINALIZED(h);
    if (!n) return GAIA_ERR_INVALID_ARG;
    if (mesh < 0 || mesh >= (int)h->s.meshes.size()) return fail(h->s, GAIA_ERR_INVALID_ARG, "bad mesh id");
    cudaSetDevice(h->s.device);
    *n = collision_fetch_relations(h->s, mesh, out, capacity);
    return *n < 0 ? GAIA_ERR_CUDA : GAIA_OK;
}

int gaia_b200_enable_mesh(gaia_handle* h, int mesh, int activeForCollision, int activeForMaterialSolve) {
    REQUIRE(h);
    Scene& s = h->s;
    if (mesh < 0 || mesh >= (int)s.meshes.size()) return fail(s, GAIA_ERR_INVALID_ARG, "bad mesh id");
    s.meshes[mesh].activeForCollision = activeForCollision != 0;
    s.meshes[mesh].activeForMaterialSolve = activeForMaterialSolve != 0;
    s.meshParamsDirty = true;
    return GAIA_OK;
}

int gaia_b200_set_partition_mode(gaia_handle* h, int32_t myRank, int32_t worldSize) {
    REQUIRE(h);
    Scene& s = h->s;
    if (!s.meshes.empty()) return fail(s, GAIA_ERR_STATE, "set_partition_mode must precede add_tet_mesh");
    if (worldSize < 1 || myRank < 0 || myRank >= worldSize) return fail(s, GAIA_ERR_INVALID_ARG, "bad rank / world size");
    s.partitionMode = true; s.myRank = myRank; s.worldSize = worldSize;
    return GAIA_OK;
}
int gaia_b200_set_ghosts(gaia_handle* h, int mesh, const int32_t* verts, int32_t n) {
    REQUIRE(h);
    Scene& s = h->s;
    if (s.finalized) return fail(s, GAIA_ERR_STATE, "set_ghosts after finalize");
    if (mesh < 0 || mesh >= (int)s.meshes.size() || n < 0 || (n && !verts)) return fail(s, GAIA_ERR_INVALID_ARG, "bad argument");
    HostMesh& m = s.meshes[mesh];
    for (int i = 0; i < n; i++) {
        if (verts[i] < 0 || verts[i] >= m.nV) return fail(s, GAIA_ERR_INVALID_ARG, "ghost vertex out of range");
        m.ghostMask[verts[i]] = 1;
    }
    return GAIA_OK;
}
int gaia_b200_get_sorted_ids(gaia_handle* h, int mesh, const int32_t* verts, int32_t n, int32_t* out) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    if (mesh < 0 || mesh >= (int)s.meshes.size() || n < 0 || (n && (!verts || !out))) return fail(s, GAIA_ERR_INVALID_ARG, "bad argument");
    const HostMesh& m = s.meshes[mesh];
    for (int i = 0; i < n; i++) {
        if (verts[i] < 0 || verts[i] >= m.nV) return fail(s, GAIA_ERR_INVALID_ARG, "vertex out of range");
        out[i] = m.sortedId[verts[i]];
    }
    return GAIA_OK;
}
int gaia_b200_halo_ipc_handles(gaia_handle* h, void* out128) {
    REQUIRE(h); FINALIZED(h);
    if (!out128) return GAIA_ERR_INVALID_ARG;
    cudaSetDevice(h->s.device);
    cudaError_t e = halo_ipc_handles(h->s, out128);
    return e == cudaSuccess ? GAIA_OK : cuda_fail(h->s, e);
}
int gaia_b200_halo_add_peer(gaia_handle* h, int32_t peerRank, const void* peerHandles128, int32_t nSend, const int32_t* sendMeshVerts,
                            const int32_t* remoteSlots) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    if (!peerHandles128 || nSend < 0 || (nSend && (!sendMeshVerts || !remoteSlots)) || peerRank < 0 || peerRank >= s.worldSize || peerRank == s.myRank)
        return fail(s, GAIA_ERR_INVALID_ARG, "bad halo peer argument");
    for (int i = 0; i < nSend; i++) {
        const int mm = sendMeshVerts[2 * i], v = sendMeshVerts[2 * i + 1];
        if (mm < 0 || mm >= (int)s.meshes.size() || v < 0 || v >= s.meshes[mm].nV) return fail(s, GAIA_ERR_INVALID_ARG, "halo send vertex out of range");
    }
    cudaSetDevice(s.device);
    cudaError_t e = halo_add_peer(s, peerRank, peerHandles128, nSend, sendMeshVerts, remoteSlots);
    return e == cudaSuccess ? GAIA_OK : cuda_fail(s, e);
}
int gaia_b200_halo_local_pointers(gaia_handle* h, void* out16) {
    REQUIRE(h); FINALIZED(h);
    if (!out16) return GAIA_ERR_INVALID_ARG;
    cudaSetDevice(h->s.device);
    const cudaError_t e = halo_local_pointers(h->s, out16);
    return e == cudaSuccess ? GAIA_OK : cuda_fail(h->s, e);
}
int gaia_b200_halo_add_peer_pointers(gaia_handle* h, int32_t peerRank, int32_t peerDevice, const void* peerPointers16, int32_t nSend,
                                     const int32_t* sendMeshVerts, const int32_t* remoteSlots) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    if (!peerPointers16 || nSend < 0 || (nSend && (!sendMeshVerts || !remoteSlots)) || peerRank < 0 || peerRank >= s.worldSize || peerRank == s.myRank)
        return fail(s, GAIA_ERR_INVALID_ARG, "bad halo peer argument");
    for (int i = 0; i < nSend; i++) {
        const int mm = sendMeshVerts[2 * i], v = sendMeshVerts[2 * i + 1];
        if (mm < 0 || mm >= (int)s.meshes.size() || v < 0 || v >= s.meshes[mm].nV) return fail(s, GAIA_ERR_INVALID_ARG, "halo send vertex out of range");
    }
    cudaSetDevice(s.device);
    const cudaError_t e = halo_add_peer(s, peerRank, nullptr, nSend, sendMeshVerts, remoteSlots, peerPointers16, peerDevice);
    return e == cudaSuccess ? GAIA_OK : cuda_fail(s, e);
}
int gaia_b200_halo_set_world(gaia_handle* h, int32_t worldSize, const void* handles) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    if (!handles || worldSize != s.worldSize) return fail(s, GAIA_ERR_INVALID_ARG, "halo_set_world: one 128-byte record per rank of the partition mode");
    cudaSetDevice(s.device);
    const cudaError_t e = halo_set_world(s, worldSize, handles);
    return e == cudaSuccess ? GAIA_OK : cuda_fail(s, e);
}
int gaia_b200_halo_finalize(gaia_handle* h) {
    REQUIRE(h); FINALIZED(h);
    cudaError_t e = halo_finalize(h->s);
    return e == cudaSuccess ? GAIA_OK : cuda_fail(h->s, e);
}

int gaia_b200_tet_surface(int32_t nV, int32_t nT, const int32_t* tets, int32_t* nSurfV, int32_t* surfaceVIds, int32_t* nSurfF, int32_t* surfaceFaces) {
    if (nV <= 0 || nT <= 0 || !tets) return GAIA_ERR_INVALID_ARG;
    for (int64_t i = 0; i < 4 * (int64_t)nT; i++) if (tets[i] < 0 || tets[i] >= nV) return GAIA_ERR_INVALID_ARG;
    HostMesh m;
    std::vector<float> zeros(3 * (size_t)nV, 0.f);
    for (int v = 0; v < nV; v++) { zeros[3 * (size_t)v] = (float)v; }      // positions do not matter for the topology
    gaia_object_params obj;
    gaia_b200_default_object_params(&obj);
    const int32_t offs[2] = {0, 0};
    host_mesh_init(m, nV, nT, zeros.data(), tets, obj, 0, offs, nullptr, 0, nullptr, true);
    if (nSurfV) *nSurfV = m.nSurfV();
    if (nSurfF) *nSurfF = m.nSurfF();
    if (surfaceVIds) std::copy(m.surfaceVIds.begin(), m.surfaceVIds.end(), surfaceVIds);
    if (surfaceFaces) std::copy(m.surfFaces.begin(), m.surfFaces.end(), surfaceFaces);
    return GAIA_OK;
}

int gaia_b200_get_stats(gaia_handle* h, gaia_stats* out) {
    REQUIRE(h);
    if (!out) return GAIA_ERR_INVALID_ARG;
    *out = h->s.stats;
    out->solveKernelMs = h->s.solveKernelMs;
    return GAIA_OK;
}
int gaia_b200_reset_stats(gaia_handle* h) {
    REQUIRE(h);
    std::memset(&h->s.stats, 0, sizeof(gaia_stats));
    h->s.solveKernelMs = 0.0; h->s.solveKernelLaunches = 0;
    return GAIA_OK;
}
int gaia_b200_set_profiling(gaia_handle* h, int enabled) { REQUIRE(h); h->s.profiling = enabled < 0 ? 0 : (enabled > 2 ? 1 : enabled); return GAIA_OK; }
int gaia_b200_set_solver_lanes(gaia_handle* h, int lanes) {
    REQUIRE(h);
    if (lanes != 0 && lanes != 4 && lanes != 8) return fail(h->s, GAIA_ERR_INVALID_ARG, "solver lanes: 0 (default), 4 or 8");
    if (h->s.iterGraphExec) { cudaGraphExecDestroy(h->s.iterGraphExec); h->s.iterGraphExec = nullptr; }
    h->s.solveLanes = lanes;
    return GAIA_OK;
}
int gaia_b200_set_line_search(gaia_handle* h, int32_t enabled, float tau, float tauGD, int32_t maxIters) {
    REQUIRE(h); FINALIZED(h);
    Scene& s = h->s;
    if (enabled && s.phys.useBowlGround) return fail(s, GAIA_ERR_UNSUPPORTED, "line search with the bowl ground: its merit energy has no bowl term");
    if (enabled && (!(tau > 0.f && tau < 1.f) || !(tauGD > 0.f && tauGD < 1.f) || maxIters < 0 || maxIters > 64))
        return fail(s, GAIA_ERR_INVALID_ARG, "line search: 0 < tau, tauGD < 1 and 0 <= maxIters <= 64");
    cudaSetDevice(s.device);
    if (s.iterGraphExec) { cudaGraphExecDestroy(s.iterGraphExec); s.iterGraphExec = nullptr; }
    if (enabled) {
        const cudaError_t e = linesearch_upload(s);
        if (e != cudaSuccess) return cuda_fail(s, e);
        s.lineSearch.tau = tau; s.lineSearch.tauGD = tauGD; s.lineSearch.maxIters = maxIters;
    }
    s.lineSearch.enabled = enabled != 0;
    return GAIA_OK;
}
// BasePhysicsParams::bowlCenter / bowlRadius / bowlCap (M/Parameters/PhysicsParameters.h:18, 40-45) for a scene created with useBowlGround
int gaia_b200_set_bowl_ground(gaia_handle* h, const float center[3], float radius, int32_t cap) {
    REQUIRE(h);
    Scene& s = h->s;
    if (!center || !(radius > 0.f)) return fail(s, GAIA_ERR_INVALID_ARG, "bowl ground: a centre and a positive radius");
    for (int d = 0; d < 3; d++) { s.bowl.center[d] = center[d]; s.stepParams.bowlCenter[d] = center[d]; }
    s.bowl.radius = radius; s.bowl.cap = cap != 0;
    s.stepParams.bowlRadius = radius; s.stepParams.bowlCap = cap != 0;
    if (s.iterGraphExec) { cudaSetDevice(s.device); cudaGraphExecDestroy(s.iterGraphExec); s.iterGraphExec = nullptr; }   // the captured launches hold the old values
    return GAIA_OK;
}
int64_t gaia_b200_line_search_shrinks(gaia_handle* h) {
    if (!h || !h->s.lsShrinks.p) return 0;
    unsigned long long v = 0;
    cudaSetDevice(h->s.device);
    if (cudaStreamSynchronize(h->s.stream) != cudaSuccess || cudaMemcpy(&v, h->s.lsShrinks.p, sizeof(v), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return (int64_t)v;
}
int gaia_b200_set_lazy_refit_margin(gaia_handle* h, float edges) {
    REQUIRE(h);
    if (h->s.finalized) return fail(h->s, GAIA_ERR_STATE, "set_lazy_refit_margin after finalize");
    if (!(edges <= 4.f)) return fail(h->s, GAIA_ERR_INVALID_ARG, "lazy refit margin: 0 .. 4 typical tet edges, or negative for the adaptive default");
    h->s.lazyRefitMargin = edges;
    return GAIA_OK;
}
int gaia_b200_set_multi_colour_launch(gaia_handle* h, int enabled) {
    REQUIRE(h);
    if (h->s.iterGraphExec) { cudaGraphExecDestroy(h->s.iterGraphExec); h->s.iterGraphExec = nullptr; }
    if (enabled < 0 || enabled > 2) return fail(h->s, GAIA_ERR_INVALID_ARG, "multi-colour launch: 0 (off), 1 (default) or 2 (always)");
    h->s.multiStepMode = enabled;
    return GAIA_OK;
}
int gaia_b200_get_layout_bytes(gaia_handle* h, int64_t out[6]) {
    REQUIRE(h); FINALIZED(h);
    if (!out) return GAIA_ERR_INVALID_ARG;
    const Scene& s = h->s;
    out[0] = (int64_t)s.rec.n * 8; out[1] = (int64_t)s.nbw.n * 8; out[2] = (int64_t)s.nVerts * (16 + 16 + 16 + 16);
    out[3] = (int64_t)s.tiles.n * (int64_t)sizeof(TileDesc); out[4] = s.sweepPosReads; out[5] = s.nInc;
    return GAIA_OK;
}
void* gaia_b200_cuda_stream(gaia_handle* h) { return h ? (void*)h->s.stream : nullptr; }

}  // extern "C"
