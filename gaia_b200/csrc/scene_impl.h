// Scene object behind gaia_handle: host meshes + device buffers + stream/graph state.
#pragma once
#include "scene.h"
#include <cstdio>
#include <map>

namespace gb {

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;                 // owning: a copy would free the same allocation twice
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
    DevBuf& operator=(DevBuf&& o) noexcept { if (this != &o) { free(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; } return *this; }
    cudaError_t alloc(size_t count) {
        free();
        n = count;
        if (!count) return cudaSuccess;
        return cudaMalloc((void**)&p, count * sizeof(T));
    }
    void free() { if (p) cudaFree(p); p = nullptr; n = 0; }
    ~DevBuf() { free(); }
};

struct CollisionState;   // collision.cu
struct HaloState;        // halo.cu

// VBDDeformer.h: Rotator / Translator evaluated on the device
struct Deformer {
    int type = 0;                      // 0 rotator, 1 translator
    int mesh = 0;
    std::vector<int32_t> verts;        // original vertex ids
    float origin[3] = {0, 0, 0}, axis[3] = {0, 0, 0}, speed[3] = {0, 0, 0};
    double angularVelocity = 0, endTime = -1;
    DevBuf<int32_t> sid;               // sorted ids
    DevBuf<float4> root, rootP;        // rotator: root, RootP; translator: original position in root
};

// what the colour-solve kernel needs to see of the contact data (filled by collision.cu)
struct ContactView {
    const int32_t* contactOfVertex;   // sorted vertex -> contact slot or -1
    const int32_t* relOff;            // CSR over sorted vertices: contacts on faces this vertex belongs to
    const int2* rel;                  // (contact slot, vertex order 0..2)
    const float4* cF;                 // force.xyz, penetrationDepth
    const float4* cH0;                // h00,h01,h02,h11
    const float4* cH1;                // h12,h22,bary0,bary1
    const float* cB2;                 // bary2
};

// what the colour-solve kernel needs for the fused halo exchange (filled by halo.cu): boundary vertices (bit 30 of
// vtx.w) wait for the neighbours' epoch before they gather ghost positions and store their new position straight into
// the neighbours' ghost slots; the last block of the launch publishes the new epoch
static const int kMaxHaloPeers = 8;
static const int kHaloBoundaryBit = 1 << 30;
struct HaloPeerTable {                        // lives in device memory (indexed dynamically by the kernels)
    float4* peerPos[kMaxHaloPeers];
    uint32_t* peerFlag[kMaxHaloPeers];        // &neighbourFlags[myRank]
    const uint32_t* myFlag[kMaxHaloPeers];    // &myFlags[neighbourRank]
};
struct HaloView {
    int nPeers;
    const HaloPeerTable* tab;
    const int32_t* sendIdx;                   // per sorted vertex: first entry << 4 | entry count (boundary vertices only)
    const int2* sendEnt;                      // (peer index, remote slot)
    const int32_t* boundaryPrefix;            // [N+1] number of boundary vertices below a sorted id
    uint32_t* ctl;                            // [0] epoch, [1] block ticket
    int epochOffset;                          // colour launches of this batch before this one
    int debug;                                // 0 (kept for layout; bit 0 = do not wait, bit 1 = do not push: timing experiments only)
};

// One tile of the colour sweep (sweep.cu): up to kTileV consecutive vertices of one colour whose incidence records and
// neighbour entries are contiguous in the `rec` / `nbw` streams, so one cp.async.bulk per stream brings a tile into shared
// memory. Counts are padded to even (16-byte copies); begins are even.
static const int kSweepLanes = 4;      // lanes per vertex of the colour sweep (fixes the record order and the association of the sums)
static const int kTileV = 64, kTileRecCap = 1536, kTileNbCap = 896;
struct TileDesc { int32_t vBegin, nV, recBegin, recCount, nbBegin, nbCount, flags, pad; };

struct Scene {
    gaia_physics_params phys;
    gaia_collision_params col;
    float dt = 0.f, dtSqrReciprocal = 0.f;
    int device = 0;
    std::string err;
    bool finalized = false;
    int frameId = 0, substep = 0;

    std::vector<HostMesh> meshes;
    std::vector<std::vector<int32_t>> groups;   // [mesh,v,...] per merged parallel group
    std::vector<int32_t> groupStart;            // sorted-id range of each group
    std::vector<int32_t> sortedToOrig;          // sorted id -> global original index (vertBase + v)
    int64_t nVerts = 0, nTets = 0, nSurfVerts = 0, nSurfFaces = 0, nInc = 0;
    // host staging of the incidence streams (freed after upload)
    std::vector<int32_t> h_incOff, h_nbOff, h_nbList;
    std::vector<uint32_t> h_incIdx;   // packed local neighbour slots a | b<<8 | c<<16
    std::vector<float> h_incW;        // 4 floats per incidence: A*w'.xyz, det(B')

    // ---- device state (sorted order)
    cudaStream_t stream = nullptr;
    // frame output (capi.cu: surface snapshots): second stream, device staging and page-locked ring, two slots
    cudaStream_t ioStream = nullptr;
    DevBuf<int32_t> surfSidAll;        // sorted id of every surface vertex, meshes concatenated (writeAllToBinary order)
    DevBuf<float> surfStage[2];
    float* surfPinned[2] = {nullptr, nullptr};
    cudaEvent_t surfGathered[2] = {nullptr, nullptr}, surfCopied[2] = {nullptr, nullptr};
    DevBuf<float4> pos, prev, inertia, vel, velPrev, accel, accPrevPrev, accPrevIter;
    DevBuf<float> mass;
    DevBuf<int32_t> info;          // meshId | fixed<<31
    int64_t nNb = 0;
    // tile streams of the colour sweep (sweep.cu; built by build_tile_streams)
    std::vector<int4> h_vtx2;          // per vertex: {recLocal | nRec << 16, nbLocal | nNb << 16, bits(A*sum|m0|^2), meshId | flags}
    std::vector<uint2> h_rec;          // per incidence: {bits(det B'), slots a | b << 8 | c << 16}
    std::vector<int2> h_nbw;           // per (vertex, distinct neighbour): {sorted id, bits(sum of A*w' over the shared tets)}
    std::vector<TileDesc> h_tiles;
    std::vector<int32_t> groupTileStart;   // [G + 1]
    DevBuf<int2> groupTilesDev;            // [G] {first tile, end tile}: what the multi-colour launch steps through
    DevBuf<uint32_t> gridBar;              // its grid barrier: [0] arrivals, [1] exits
    int multiStep = -1;                    // -1 not probed yet, 0 unavailable, 1 available
    int multiStepMode = 1;                 // gaia_b200_set_multi_colour_launch: 0 never, 1 where the host issues the loop, 2 always
    int64_t sweepPosReads = 0;             // shared-memory position reads per sweep over all vertices (3 per restart, 1 per strip step)
    DevBuf<int4> vtx2;
    DevBuf<uint2> rec;
    DevBuf<int2> nbw;
    DevBuf<TileDesc> tiles;
    int numSMs = 148;
    uint32_t sweepAttrSet = 0;         // kernel variants whose shared-memory attribute is set on this scene's device
    DevBuf<int32_t> sorted2orig, orig2sorted;
    DevBuf<MeshParamsDev> meshParams;
    DevBuf<float4> ioStage;        // original-order staging for get/set
    DevBuf<int32_t> meshSubsteps;  // substeps each mesh has taken (hasVelocitiesPrev / hasApproxAcceleration)
    DevBuf<int32_t> errFlags;      // [0]=NaN count, [1]=solver-failed count, [2]=contact overflow, [3]=traversal overflow
    std::vector<MeshParamsDev> h_meshParams;
    bool meshParamsDirty = true;
    StepParamsDev stepParams;
    float* pinnedIO = nullptr;
    std::vector<std::pair<char*, size_t>> hostRegs;   // caller buffers page-locked with register_host_buffer
    size_t pinnedIOBytes = 0;

    CollisionState* cs = nullptr;
    HaloState* halo = nullptr;
    bool partitionMode = false;         // multi-GPU partition: colour order kept as given
    int myRank = 0, worldSize = 1;
    std::vector<Deformer*> deformers;
    float curTime = 0.f;                // BasePhysicFramework::curTime (FloatingType)

    // CUDA graph of the colour loop of one substep (solve.cu: solve_iterations)
    cudaGraphExec_t iterGraphExec = nullptr;
    int iterGraphIters = -1, iterCalls = 0;
    int solveLanes = 0;   // gaia_b200_set_solver_lanes: accepted for compatibility, the sweep always uses kSweepLanes
    int64_t iterDelta[4] = {0, 0, 0, 0};

    // statistics / profiling
    gaia_stats stats;
    // gaia_b200_set_line_search (sweep.cu: k_sweep_linesearch): plain per-tet tables, uploaded when the option is first switched on
    struct { bool enabled = false; float tau = 0.5f, tauGD = 0.1f; int maxIters = 16; } lineSearch;
    struct { float center[3] = {0.f, 0.f, 0.f}; float radius = 1.f; int cap = 0; } bowl;      // used when phys.useBowlGround (gaia_b200_set_bowl_ground)
    DevBuf<int32_t> bowlNoContact, bowlZeroOff;   // bowl ground without contacts: empty contact tables for the contact variant of the sweep ([N] = -1, [N + 1] = 0)
    DevBuf<float4> symScratch;          // initializationType 5: the Jacobi target of the symplectic-Euler initial guess (N; only if some mesh asks for it)
    DevBuf<int32_t> lsOff; DevBuf<int2> lsEnt; DevBuf<int4> lsTetSid; DevBuf<float4> lsTetDm; DevBuf<unsigned long long> lsShrinks;
    float lazyRefitMargin = -1.f;  // gaia_b200_set_lazy_refit_margin: margin of the DCD trees' boxes in typical tet edges; < 0: adaptive, from 0.2
    int profiling = 0;             // 0 off; 1 per-phase timers + an event pair around every colour-solve launch; 2 per-phase timers only (the colour loop stays a CUDA graph)
    cudaEvent_t evA = nullptr, evB = nullptr, evP0 = nullptr, evP1 = nullptr;
    std::vector<cudaEvent_t> profEv;   // event pairs around the colour-solve launches of one iteration loop (profiling only)
    size_t profUsed = 0;
    double solveKernelMs = 0.0;    // accumulated colour-solve kernel time while profiling
    int64_t solveKernelLaunches = 0;

    ~Scene();
};

// host_setup.cpp
void host_mesh_init(HostMesh& m, int32_t nV, int32_t nT, const float* xyz, const int32_t* tets, const gaia_object_params& obj,
                    int32_t nColors, const int32_t* colorOffsets, const int32_t* colorVerts, int32_t nFixed, const int32_t* fixed,
                    bool keepColourOrder);
bool host_finalize(Scene& s);
bool build_tile_streams(Scene& s, int tileV);

// sweep.cu: the per-colour vertex solve over tiles (TMA-fed persistent blocks)
cudaError_t sweep_launch(Scene& s, int group, int applyFriction, bool hasContacts, const ContactView& cv, const HaloView* hv);
cudaError_t linesearch_upload(Scene& s);
cudaError_t sweep_symplectic_guess(Scene& s);      // initializationType 5 (no-op unless symScratch was allocated at finalize)
cudaError_t sweep_launch_linesearch(Scene& s, int group, int applyFriction, bool hasContacts, const ContactView& cv);
bool sweep_multi_available(Scene& s);
cudaError_t sweep_launch_multi(Scene& s, int firstGroup, int nSteps, int frictionFromStep, const HaloView* hv);
cudaError_t sweep_upload(Scene& s);
cudaError_t sweep_mark_halo_tiles(Scene& s, const int32_t* boundaryPrefixDev);

// solve.cu
cudaError_t device_upload(Scene& s);
cudaError_t upload_deformers(Scene& s);
cudaError_t sync_mesh_params(Scene& s);
cudaError_t run_substeps(Scene& s, int nSubsteps);                 // one frame's worth of substeps
cudaError_t run_iterations_only(Scene& s, int nIters);
cudaError_t gather_to_host(Scene& s, const float4* devSorted, int64_t origBegin, int64_t count, float* outXYZ);
cudaError_t scatter_from_host(Scene& s, float4* devSorted, int64_t origBegin, int64_t count, const float* inXYZ, bool keepW);
int check_device_errors(Scene& s);

// halo.cu
cudaError_t halo_ipc_handles(Scene& s, void* out128);
cudaError_t halo_local_pointers(Scene& s, void* out16);
cudaError_t halo_add_peer(Scene& s, int peerRank, const void* handles128, int nSend, const int32_t* sendMeshVerts, const int32_t* remoteSlots,
                          const void* pointers16 = nullptr, int peerDevice = -1);
cudaError_t halo_finalize(Scene& s);
cudaError_t halo_set_world(Scene& s, int world, const void* handles);   // flag arrays of ALL ranks: agreement votes
bool halo_can_vote(const Scene& s);
cudaError_t halo_vote(Scene& s, bool mine, bool* any);                 // "does any rank say yes?" (host-blocking)
void halo_destroy(Scene& s);
bool halo_active(const Scene& s);
cudaError_t halo_push(Scene& s, int group);
cudaError_t halo_sync(Scene& s);
bool halo_view(Scene& s, HaloView& hv);                           // fused exchange available (halo_finalize done)
void halo_count_fused(Scene& s, int group);                       // statistics: bytes a fused colour launch pushes

// collision.cu
cudaError_t collision_init(Scene& s);
void collision_destroy(Scene& s);
void collision_invalidate(Scene& s);                              // force BVH rebuilds at the next detection
cudaError_t collision_dcd(Scene& s, bool intermediate);
cudaError_t collision_ccd(Scene& s);
cudaError_t collision_prepare(Scene& s);
cudaError_t collision_update_infos(Scene& s, int group, bool deviceSized = false);   // per-colour contact force/Hessian pre-pass
bool collision_graphable(const Scene& s);                         // contact lists small enough for device-sized pre-pass launches
int64_t collision_fetch_contacts(Scene& s, gaia_contact* out, int64_t cap);
int64_t collision_leaf_order(Scene& s, int which, int32_t* out, int64_t cap);
int64_t collision_fetch_active_list(Scene& s, int group, int32_t* out, int64_t cap);
int64_t collision_fetch_relations(Scene& s, int mesh, int32_t* out, int64_t cap);
bool collision_contact_view(Scene& s, ContactView& cv, const uint8_t** collidingMask, bool force = false);

}  // namespace gb
