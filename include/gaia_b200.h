/*
 * gaia_b200.h — C ABI of the B200-native Vertex Block Descent (VBD) step.
 *
 * This is the drop-in boundary for ONE hot path of AnkaChan/Gaia: the VBD step
 * (per-colour vertex solve + collision detection).  The reference has no FFI
 * layer; its seam is a set of C++ call sites inside `VBDPhysics`.  Every entry
 * point below names the reference interface it replaces (paths relative to
 * Simulator/Modules/ of the reference tree, "M/").
 *
 * Conventions
 *   - plain pointers and sizes only; no C++/torch types cross this boundary;
 *   - every function returns GAIA_OK (0) or a negative gaia_status; the text of
 *     the last error of a handle is available from gaia_b200_last_error();
 *   - one host thread per handle; work is enqueued on an internal CUDA stream
 *     and `*_get_*` calls synchronise before copying out;
 *   - positions / velocities cross the boundary as xyz-interleaved float32
 *     (the reference's Eigen 3xN column-major layout, M/TetMesh/TetMeshFEM.h:158)
 *     in the ORIGINAL vertex order of each mesh — the colour-sorted device
 *     order is internal;
 *   - there is no CPU fallback: without a usable CUDA device create() fails.
 */
#ifndef GAIA_B200_H
#define GAIA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GAIA_B200_ABI_VERSION 1
#define GAIA_COLOR_BALANCE 1
#define GAIA_COLOR_MCS 2

typedef enum gaia_status {
    GAIA_OK = 0,
    GAIA_ERR_INVALID_ARG = -1,
    GAIA_ERR_NO_DEVICE = -2,     /* no CUDA device / driver: there is no CPU path */
    GAIA_ERR_CUDA = -3,
    GAIA_ERR_STATE = -4,         /* call order violated (e.g. step before finalize) */
    GAIA_ERR_UNSUPPORTED = -5,   /* a reference feature outside the hot path, or a combination of options that is not built */
    GAIA_ERR_OVERFLOW = -6,      /* a device-side list overflowed its capacity */
    GAIA_ERR_NAN = -7,           /* NaN met in the solve (reference: std::exit(-1), M/VBD/VBDPhysics.cpp:2460) */
    GAIA_ERR_TIMEOUT = -8        /* a neighbour partition stopped answering the halo exchange */
} gaia_status;

/* PhysicsParams of Parameters.json.
 * Mirrors BasePhysicsParams (M/Parameters/PhysicsParameters.h:9-105) and
 * VBDPhysicsParameters (M/VBD/VBDPhysicsParameters.h:5-82); same names, same
 * defaults (gaia_b200_default_physics_params). Only fields the step reads. */
typedef struct gaia_physics_params {
    float gravity[3];                       /* (0,-10,0) */
    float timeStep;                         /* 0.0166666 */
    int32_t numSubsteps;                    /* 20 */
    int32_t iterations;                     /* 1 */
    int32_t usePlaneGround;                 /* 1 */
    int32_t useBowlGround;                  /* 0; 1: the inside of a sphere is the ground (gaia_b200_set_bowl_ground) instead of the world box */
    int32_t checkAndUpdateWorldBounds;      /* 0 */
    float worldBoundsMin[3];                /* -1e7 */
    float worldBoundsMax[3];                /* +1e7 */
    float boundaryFrictionDynamic;          /* 0.1 */
    int32_t dcdTetMeshSceneBVHRebuildSteps; /* 16 */
    int32_t dcdSurfaceSceneBVHRebuildSteps; /* 3 */
    int32_t ccdBVHRebuildSteps;             /* 7 */
    /* VBDPhysicsParameters */
    float stepSize;                         /* 1.0 */
    float stepSizeGD;                       /* 1e-3 */
    int32_t associateGravityWithInertia;    /* 1 */
    float collisionStiffness;               /* 1e5 */
    int32_t intermediateCollisionIterations;/* -1 */
    float boundaryCollisionStiffness;       /* 1e5 */
    float boundaryFrictionEpsV;             /* 1.0 */
    int32_t frictionStartIter;              /* 0 */
    int32_t useAccelerator;                 /* 0 */
    float acceleratorRho;                   /* 0.93 */
    float collisionOffHeight;               /* 1e7 */
} gaia_physics_params;

/* CollisionParams of Parameters.json.
 * Mirrors CollisionDetectionParamters (M/CollisionDetector/CollisionDetertionParameters.h:44-80). */
typedef struct gaia_collision_params {
    int32_t allowCCD;                                   /* 1 */
    int32_t allowDCD;                                   /* 1 */
    int32_t checkFeasibleRegion;                        /* 1 */
    int32_t checkTetTraverse;                           /* 1 */
    int32_t handleSelfCollision;                        /* 1 */
    int32_t stopTraversingAfterPassingQueryPoint;       /* 1 */
    int32_t tetrahedralTraverseForNonSelfIntersection;  /* 1 */
    int32_t useStaticTraverse;                          /* 1 */
    int32_t loopLessTraverse;                           /* 0 */
    int32_t shiftQueryPointToCenter;                    /* 1 */
    float centerShiftLevel;                             /* 0.01 */
    int32_t maxNumberOfBVHQuery;                        /* 500 */
    float feasibleRegionEpsilon;                        /* 1e-2 */
    float rayTriIntersectionEPSILON;                    /* 1e-10 */
    float maxSearchDistanceMultiplier;                  /* 1.8 */
} gaia_collision_params;

/* One entry of Models.json.
 * Mirrors ObjectParams (M/Materials/Materials.h:12-53), ObjectParamsVBD
 * (M/VBD/VBD_BaseMaterial.h:21-49) and ObjectParametersVBDNeoHookean
 * (M/VBD/VBD_NeoHookean.h:16-34). */
typedef struct gaia_object_params {
    float scale[3];                 /* (1,1,1) */
    float rotation[3];              /* axis*angle, (0,0,0) */
    float translation[3];           /* (0,0,0) */
    float initialVelocity[3];       /* (0,0,0) */
    float density;                  /* 1 */
    int32_t hasNoGravZone;          /* 0 */
    float noGravZoneThreshold;      /* 0 */
    float maxVelocityMagnitude;     /* -1 */
    int32_t frameToAppear;          /* -1 */
    /* ObjectParamsVBD */
    float frictionDynamic;          /* 0.1 */
    float frictionEpsV;             /* 0.01 */
    float exponentialVelDamping;    /* 1.0 (raised to dt at create, M/VBD/VBDPhysics.cpp:794-807) */
    float constantVelDamping;       /* 0.0 (multiplied by dt likewise) */
    int32_t initializationType;     /* 0..6, M/VBD/VBD_BaseMaterial.cpp:16-131 (5: forwardStepSimplecticEuler, :134-147) */
    float initRatio_g;              /* 0.5 */
    /* NeoHookean */
    float miu;                      /* 1e4 */
    float lmbd;                     /* 1e5 */
    float dampingHydrostatic;       /* 0 */
    float dampingDeviatoric;        /* 0 */
} gaia_object_params;

/* One detected vertex-face contact, for bit-exact parity of contact sets
 * (CollidingPointInfo + CollisionDetectionResult,
 *  M/CollisionDetector/CollisionDetertionParameters.h:166-222).
 * Lists are returned sorted by (queryMesh, queryVertex). */
typedef struct gaia_contact {
    int32_t queryMesh;      /* idTMQuery */
    int32_t queryVertex;    /* idVQuery (tet-mesh vertex id, original order) */
    int32_t hitMesh;        /* intersectedMeshId */
    int32_t hitTet;         /* intersectedElement (-1 for CCD contacts) */
    int32_t hitFace;        /* closestSurfaceFaceId (-1 when no shortest path was found) */
    int32_t fromCCD;        /* 0: DCD, 1: CCD */
    int32_t shortestPathFound;
    int32_t closestPointType; /* ClosestPointOnTriangleType as int */
} gaia_contact;

/* Per-phase accumulated device time (ms) with the reference's timer names
 * (M/Timer/RunningTimeStatistics.h:54-183) and the counters the metric needs. */
typedef struct gaia_stats {
    double timeCsmpInitialStep;
    double timeCsmpMaterialSolve;
    double timeCsmpUpdateVelocity;
    double timeCsmpUpdatingBVHDCD;
    double timeCsmpColDetectDCD;
    double timeCsmpUpdatingBVHCCD;
    double timeCsmpColDetectCCD;
    double timeCsmpFrame;
    double solveKernelMs;       /* device time of the colour-solve kernels alone (CUDA events, profiling mode) */
    int64_t vertexSolves;       /* non-fixed active vertices solved, summed over colour sweeps */
    int64_t colourSweeps;       /* colour kernels executed */
    int64_t substeps;
    int64_t frames;
    int64_t kernelLaunches;     /* kernels of this library enqueued */
    int64_t algorithmicBytes;   /* SURVEY §8(d) formula, summed over colour sweeps */
    int64_t numContactsDCD;     /* last detection */
    int64_t numContactsCCD;
} gaia_stats;

typedef struct gaia_handle gaia_handle;

/* library ------------------------------------------------------------------ */
int  gaia_b200_abi_version(void);
const char* gaia_b200_last_error(const gaia_handle* h);     /* h may be NULL: last create() error */
void gaia_b200_default_physics_params(gaia_physics_params* p);
void gaia_b200_default_collision_params(gaia_collision_params* p);
void gaia_b200_default_object_params(gaia_object_params* p);

/* scene construction -------------------------------------------------------
 * create:   VBDPhysics::loadRunningparameters + BasePhysicFramework::timeStepEqualize
 *           (M/VBD/VBDPhysics.cpp:90, M/Framework/BasePhysicsFramework.cpp:245-249).
 *           `device` is the CUDA ordinal (one process per GPU). */
int gaia_b200_create(const gaia_physics_params* phys, const gaia_collision_params* col,
                     int device, gaia_handle** out);
void gaia_b200_destroy(gaia_handle* h);

/* add_tet_mesh: VBDBaseTetMesh::initialize -> TetMeshFEM::initialize +
 *           TetMeshTopology::initialize (M/TetMesh/TetMeshFEM.cpp:332-443, 40-330).
 *           xyz: 3*nV floats as read from the .t file (transform is applied inside);
 *           tets: 4*nT vertex ids; colouring = the *.vertexColoring.json lists,
 *           flattened: colour c owns color_verts[color_offsets[c] .. color_offsets[c+1]).
 *           Returns the mesh id (>=0) or a negative status. */
int gaia_b200_add_tet_mesh(gaia_handle* h, int32_t nV, int32_t nT,
                           const float* xyz, const int32_t* tets,
                           const gaia_object_params* obj,
                           int32_t nColors, const int32_t* color_offsets, const int32_t* color_verts,
                           int32_t nFixed, const int32_t* fixedPoints);

/* Deformers[] of Parameters.json (M/VBD/VBDDeformer.h:45-222), evaluated on the device after every applyInitialStep
 * (VBDPhysics::applyDeformers, M/VBD/VBDPhysics.cpp:1572-1578). The selected vertices become fixed points (setFixed).
 *   rotator:    p <- root + R(axis, angularVelocity * min(curTime, rotationEndTime)) * (p0 - root)
 *   translator: p <- p0 + speed * min(curTime, deformationEndTime)
 * end time <= 0 means "never stops". Must be called before finalize. */
int gaia_b200_add_rotator(gaia_handle* h, int mesh, const int32_t* verts, int32_t n, const float origin[3],
                          const float axis[3], double angularVelocity, double rotationEndTime);
int gaia_b200_add_translator(gaia_handle* h, int mesh, const int32_t* verts, int32_t n, const float speed[3],
                             double deformationEndTime);

/* finalize: the rest of VBDPhysics::initialize (colour-group merge
 *           M/VBD/VBDPhysics.cpp:236-274, initializeGPU :563-651,
 *           initializeCollisionDetector BasePhysicsFramework.cpp:251-262). */
int gaia_b200_finalize(gaia_handle* h);

/* stepping ------------------------------------------------------------------
 * run_frames: BasePhysicFramework::simulate loop body (enableModels -> runStep -> ++frameId,
 *           BasePhysicsFramework.cpp:355-390) repeated nFrames times; runStep follows
 *           VBDPhysics::runStep_hybridCollisionHandling (M/VBD/VBDPhysics.cpp:1634-1707). */
int gaia_b200_run_frames(gaia_handle* h, int32_t nFrames);
/* run_iterations: debugging/parity entry: applyInitialStep once, then nIters colour sweeps of
 *           the material solve (no velocity update) — "positions after one iteration". The stored velocities are left as
 *           they were (the initial step's gravity kick lives in the inertia term only; the reference's computeInertia adds
 *           it to mVelocity, where updateVelocities would overwrite it): do not continue a trajectory from this state. */
int gaia_b200_run_iterations(gaia_handle* h, int32_t nIters);
int gaia_b200_synchronize(gaia_handle* h);
int gaia_b200_frame_id(const gaia_handle* h);
/* recovery (PhysicsState::initializePhysics, M/Framework/BasePhysicsFramework.h:177-200): frameId and curTime of a saved
 * state; positions / velocities go through set_state. The BVHs are rebuilt at the next detection
 * (BasePhysicFramework::recoverFromState, BasePhysicsFramework.cpp:264-282). */
int gaia_b200_set_time(gaia_handle* h, int32_t frameId, double curTime);
double gaia_b200_cur_time(const gaia_handle* h);

/* state access (VBDBaseTetMesh::syncToCPU / syncToGPU, M/VBD/VBD_BaseMaterial.h:120-128) */
int gaia_b200_num_meshes(const gaia_handle* h);
int gaia_b200_mesh_sizes(const gaia_handle* h, int mesh, int32_t* nV, int32_t* nT,
                         int32_t* nSurfV, int32_t* nSurfF);
int gaia_b200_get_positions(gaia_handle* h, int mesh, float* xyz);
int gaia_b200_get_prev_positions(gaia_handle* h, int mesh, float* xyz);
int gaia_b200_get_velocities(gaia_handle* h, int mesh, float* xyz);
int gaia_b200_get_inertia(gaia_handle* h, int mesh, float* xyz);
int gaia_b200_set_state(gaia_handle* h, int mesh, const float* pos, const float* vel);
/* surface positions in writeAllToBinary order (M/IO/FileIO.cpp:52-83): gathered on the device, only the surface crosses PCIe */
int gaia_b200_get_surface_positions(gaia_handle* h, int mesh, float* xyz);
/* Frame output without stalling the step (writeAllToBinary / writeAllToPLY, M/IO/FileIO.cpp:52-185, feed on exactly this
 * data). snapshot_begin enqueues, behind everything already enqueued, a device-side gather of ALL meshes' surface positions
 * (meshes concatenated, the file order) and its copy into page-locked ring slot `slot` (0 or 1) on a second stream, and
 * returns at once: the next frame can be enqueued while the copy runs. snapshot_wait blocks until that slot has landed and
 * returns the ring buffer (valid until the slot's next snapshot_begin) - it may be called from another host thread, e.g. a
 * writer thread. total_surface_vertices sizes the files. */
int64_t gaia_b200_total_surface_vertices(const gaia_handle* h);
int gaia_b200_surface_snapshot_begin(gaia_handle* h, int slot);
int gaia_b200_surface_snapshot_wait(gaia_handle* h, int slot, const float** xyz, int64_t* nFloats);
/* whole-scene transfers: all meshes concatenated in mesh order (one D2H / H2D) */
int64_t gaia_b200_total_vertices(const gaia_handle* h);
int gaia_b200_get_all_positions(gaia_handle* h, float* xyz);
int gaia_b200_set_all_state(gaia_handle* h, const float* pos, const float* vel);
/* Page-lock a caller buffer (cudaHostRegister) so that get / set calls whose pointer lies inside it DMA straight from / to
 * it instead of going through the library's pinned staging buffer; the reference does the same to its Eigen storage
 * (ManagedBuffer, 3P/CuMatrix/CuMatrix/Buffers/ManagedBuffer.h:72-76). The buffer must stay alive until it is unregistered
 * or the handle is destroyed. */
int gaia_b200_register_host_buffer(gaia_handle* h, void* ptr, int64_t bytes);
int gaia_b200_unregister_host_buffer(gaia_handle* h, void* ptr);

/* integer parity surfaces ---------------------------------------------------- */
int gaia_b200_get_world_bounds(const gaia_handle* h, float* min3, float* max3);
int gaia_b200_num_parallel_groups(const gaia_handle* h);
/* out: [mesh0,v0, mesh1,v1, ...] exactly as VBDPhysics::vertexParallelGroups[g] (M/VBD/VBDPhysics.h:219-224);
 * call with out==NULL to query the pair count */
int gaia_b200_get_parallel_group(const gaia_handle* h, int group, int32_t* out, int32_t* nPairs);
int gaia_b200_get_surface_vertex_ids(const gaia_handle* h, int mesh, int32_t* out);   /* surfaceVIds */
int gaia_b200_get_surface_faces(const gaia_handle* h, int mesh, int32_t* out);        /* surfaceFacesTetMeshVIds, 3*nSurfF */
int gaia_b200_get_vertex_masses(const gaia_handle* h, int mesh, float* out);
/* collision detection on the current state (VBDPhysics::dcd / ccd, M/VBD/VBDPhysics.cpp:1958-2116),
 * without stepping; fills the contact list that get_contacts returns */
int gaia_b200_detect_collisions(gaia_handle* h, int doDCD, int doCCD);
int gaia_b200_get_contacts(gaia_handle* h, gaia_contact* out, int64_t capacity, int64_t* n);
/* which: 0 = all-tets scene, 1 = surface-faces scene; out = primitive ids in Morton leaf order */
int gaia_b200_get_bvh_leaf_order(gaia_handle* h, int which, int32_t* out, int64_t capacity, int64_t* n);
/* Prepared collision data of the last step / run_iterations call (integer parity surfaces of SURVEY.md §8 row a15):
 *  - the active collision list of one parallel group, VBDPhysics::activeCollisionsEachParallelGroup[group] as filled by
 *    ActiveCollisionList::addToActiveCollisionList (M/VBD/ActiveCollisionList.cpp:43-110, called from
 *    prepareCollisionDataCPU, M/VBD/VBDPhysics.cpp:1709-1810): pairs (queryMesh, queryVertex) of every contact that
 *    involves a vertex of the group (its own vertex or one of the three face vertices), each contact once per group;
 *    returned sorted ascending (the reference's order is its serial loop order = ascending too);
 *  - the face-side relations of one mesh (VBDCollisionDetectionResult relations, VBDPhysics.cpp:1793-1806): rows
 *    (vertex, queryMesh, queryVertex, order of `vertex` in the contact's face), sorted ascending.
 * out == NULL queries the count (pairs / rows). */
int gaia_b200_get_active_list(gaia_handle* h, int group, int32_t* outPairs, int64_t capacity, int64_t* nPairs);
int gaia_b200_get_relations(gaia_handle* h, int mesh, int32_t* outRows, int64_t capacity, int64_t* nRows);

/* mesh activation (VBDPhysics::enableModels, M/VBD/VBDPhysics.cpp:685-709) */
int gaia_b200_enable_mesh(gaia_handle* h, int mesh, int activeForCollision, int activeForMaterialSolve);

/* multi-GPU (no reference counterpart: the reference is single-process) ------------------------------
 * One process per GPU holds one PARTITION of the scene: the tets incident to the vertices it owns, i.e. owned vertices
 * plus a one-ring ghost layer. The per-vertex VBD update is local, so owned vertices see their full stencil; after every
 * colour sweep each partition stores the new positions of its boundary vertices of that colour straight into the ghost
 * slots of its neighbours' position arrays (peer memory over NVLink, opened through CUDA IPC) and the neighbours wait on
 * a device-side epoch flag before the next colour.
 *   set_partition_mode: keep the colour order as given (all partitions must sweep the same global colour at the same
 *                       time; the per-mesh size-descending sort of M/TetMesh/TetMeshFEM.cpp:122-133 would differ per
 *                       partition). Call before add_tet_mesh.
 *   set_ghosts:         vertices of `mesh` owned by another partition: never solved here, overwritten by the halo.
 *   get_sorted_ids:     device slot of each vertex (what a neighbour needs to address this partition's ghosts).
 *   halo_ipc_handles:   128 bytes: cudaIpcMemHandle_t of the position array, then of the flag array.
 *   halo_add_peer:      neighbour `peerRank`; send list = nSend vertices of mesh 0.. given as (mesh, vertex) pairs in
 *                       `sendMeshVerts` [2*nSend] and the neighbour's slots `remoteSlots` [nSend] they are stored to.
 *   halo_set_world:     (optional, processes only) the 128-byte handle records of ALL ranks, rank 0 first. With a collision
 *                       pipeline in the scene the exchange fused into the colour sweep is valid only while NO rank holds an
 *                       active contact (the contact pre-pass of a colour reads ghost positions before the sweep); with the
 *                       world known the ranks agree on that after every detection (one tiny launch over the peers' flag
 *                       arrays) and fall back to the separate barrier + push launches otherwise. Without it: always the
 *                       separate launches when a collision pipeline exists.
 *   halo_finalize:      after all peers were added and every process has opened its handles. */
int gaia_b200_set_partition_mode(gaia_handle* h, int32_t myRank, int32_t worldSize);
int gaia_b200_set_ghosts(gaia_handle* h, int mesh, const int32_t* verts, int32_t n);
int gaia_b200_get_sorted_ids(gaia_handle* h, int mesh, const int32_t* verts, int32_t n, int32_t* out);
int gaia_b200_halo_ipc_handles(gaia_handle* h, void* out128);
int gaia_b200_halo_add_peer(gaia_handle* h, int32_t peerRank, const void* peerHandles128, int32_t nSend,
                            const int32_t* sendMeshVerts, const int32_t* remoteSlots);
int gaia_b200_halo_set_world(gaia_handle* h, int32_t worldSize, const void* handles /* 128 * worldSize bytes */);
int gaia_b200_halo_finalize(gaia_handle* h);
/* The same for partitions that live in ONE process (one host thread per GPU, e.g. gaia_vbd_b200 --gpus N): CUDA IPC handles
 * cannot be opened by the process that made them, so the neighbours' arrays are addressed directly (unified addressing +
 * cudaDeviceEnablePeerAccess). local_pointers: 16 bytes = the device pointers of the position array and of the flag array. */
int gaia_b200_halo_local_pointers(gaia_handle* h, void* out16);
int gaia_b200_halo_add_peer_pointers(gaia_handle* h, int32_t peerRank, int32_t peerDevice, const void* peerPointers16, int32_t nSend,
                                     const int32_t* sendMeshVerts, const int32_t* remoteSlots);

/* Slab partitioning of one tet mesh (csrc/partition.cpp; host code, no device needed): owner of every vertex, and for one
 * rank its local mesh (tets with an owned vertex: owned vertices + a one-ring ghost layer, global order kept), the per-peer
 * send / recv vertex lists (ascending global id on both sides) and the restriction of the global colouring (colour order
 * kept, the vertices a rank sends first inside every colour). The reference has no counterpart (one device). */
typedef struct gaia_partition gaia_partition;
int gaia_b200_partition_slab_owner(int32_t nV, const float* xyz, int32_t world, int32_t axis /* -1: longest */, int32_t* owner);
int gaia_b200_partition_create(int32_t nV, int64_t nT, const int32_t* tets, const int32_t* owner, int32_t rank, int32_t world, gaia_partition** out);
void gaia_b200_partition_destroy(gaia_partition* p);
int gaia_b200_partition_sizes(const gaia_partition* p, int32_t* nLocalV, int64_t* nLocalT, int32_t* nPeers);
int gaia_b200_partition_local_verts(const gaia_partition* p, int32_t* globalIds, uint8_t* owned);
int gaia_b200_partition_local_tets(const gaia_partition* p, int32_t* tetsLocalIds, int64_t* globalTetIds);
int gaia_b200_partition_peer(const gaia_partition* p, int32_t i, int32_t* peerRank, int32_t* nSend, int32_t* nRecv);
int gaia_b200_partition_peer_lists(const gaia_partition* p, int32_t i, int32_t* sendLocal, int32_t* recvLocal);
int gaia_b200_partition_restrict_ids(const gaia_partition* p, int32_t n, const int32_t* globalIds, int32_t* localOut);
int gaia_b200_partition_restrict_colors(const gaia_partition* p, int32_t nColors, const int32_t* offsets, const int32_t* verts,
                                        int32_t* outOffsets, int32_t* outVerts);
/* Surface of a tet mesh without a device (the orderings writeAllToBinary / writeAllToPLY use): call with NULL outputs for the
 * counts. surfaceVIds ascending; surfaceFaces: 3 tet-mesh vertex ids per face, in the reference's face order. */
int gaia_b200_tet_surface(int32_t nV, int32_t nT, const int32_t* tets, int32_t* nSurfV, int32_t* surfaceVIds, int32_t* nSurfF, int32_t* surfaceFaces);

/* statistics */
int gaia_b200_get_stats(gaia_handle* h, gaia_stats* out);
int gaia_b200_reset_stats(gaia_handle* h);
/* average device time (ms) of the colour-solve kernels since the last reset, measured with CUDA
 * events on the launching stream when profiling is enabled */
int gaia_b200_set_profiling(gaia_handle* h, int enabled);   /* 0 off; 1 phase timers + per-launch events; 2 phase timers only (the colour loop stays a CUDA graph) */
void* gaia_b200_cuda_stream(gaia_handle* h);   /* cudaStream_t, for callers that time with their own events */
/* Kept for round-1 callers: the colour sweep always runs 4 lanes per vertex (the lane count fixes the record order and the
 * association of the per-vertex sums, so it is a property of the layout built at finalize, not a run-time switch).
 * Accepts 0, 4 or 8 and has no effect. */
int gaia_b200_set_solver_lanes(gaia_handle* h, int lanes);
/* Contact-free stretches of the colour loop (all colours of the iterations up to the next intermediate detection) run as ONE
 * cooperative launch with a grid-wide barrier between colours (csrc/sweep.cu, MULTI) - the device-side form of the loop
 * M/VBD/VBDPhysics.cpp:1659-1702. Results are bit-identical to one launch per colour. enabled: 0 = one launch per colour
 * everywhere; 1 (default) = multi-colour launches where the host issues the loop (intermediate detection cuts it into
 * stretches), while loops that replay from a CUDA graph keep one launch per colour (measured faster on large colours);
 * 2 = multi-colour launches everywhere (no graph replay). Colours with active contacts always use one launch per colour
 * (their contact pre-pass is a separate launch). */
int gaia_b200_set_multi_colour_launch(gaia_handle* h, int enabled);
/* Backtracking line search on the vertex merit energy after the 3x3 solve: the reference's APPLY_LOCAL_LINE_SEARCH
 * (M/VBD/VBDPhysics.cpp:25 - compiled out there -, call site :2467-2474; backTracingLineSearchVBD, M/VBD/VBD_NeoHookean.h:115-153;
 * evaluateVertexMeritEnergy, M/VBD/VBD_NeoHookean.cpp:90-124). Merit energy = inertia + Neo-Hookean energy of the incident
 * tets (no boundary / contact / damping terms, as in the reference); first trial alpha = stepSize, c = 0, alpha *= tau after
 * a rejected trial (tauGD when the 3x3 solve failed and the direction is the force), at most maxIters trials, the vertex stays
 * at the LAST trial. tau / tauGD / maxIters = PhysicsParams tau (0.5), tau_GD (0.1), backtracingLineSearchMaxIters (16).
 * A feature switch (csrc/sweep.cu: k_sweep_linesearch, one thread per vertex over plain per-tet tables uploaded on first use,
 * ~100 B per tet), not a fast path; with it the colour loop uses neither the multi-colour launch nor the fused halo exchange.
 * After finalize. line_search_shrinks: rejected trials so far (diagnostic). */
int gaia_b200_set_line_search(gaia_handle* h, int32_t enabled, float tau, float tauGD, int32_t maxIters);
/* The bowl ground of a scene created with gaia_physics_params::useBowlGround = 1: BasePhysicsParams::bowlCenter (0, 0, 0),
 * bowlRadius (1), bowlCap (false) (M/Parameters/PhysicsParameters.h:18, 40-45). The inside of the sphere is the ground: a vertex
 * further than `radius` from `center` - below the centre only, unless `cap` - is pushed back with boundaryCollisionStiffness and
 * rubbed with the boundary friction, exactly as VBDPhysics::accumlateBoundaryForceAndHessian writes it (M/VBD/VBDPhysics.cpp:
 * 2824-2864; the world box is then not applied, :2865). The term is built into the contact variant of the colour sweep, to which
 * such scenes are routed; not combined with the fused halo exchange, the multi-colour launch or the line search. Any time after create. */
int gaia_b200_set_bowl_ground(gaia_handle* h, const float center[3], float radius, int32_t cap);
int64_t gaia_b200_line_search_shrinks(gaia_handle* h);
/* Lazy refit of the DCD trees (csrc/collision.cu): boxes are inflated by a margin and refitted only when some vertex has moved
 * further than that from where the boxes were computed. `edges` = the margin in typical tet edges (cube root of 6 x the mean
 * rest volume); 0 = refit at every detection; negative (the default) = adaptive: starts at 0.2, grows by 1.3x per refit while
 * refits come more often than every 6 detections, stays below 0.48 (at half an edge the candidate neighbourhood of a vertex
 * doubles), and goes back to 0.2 for good when growing did not space the refits out (fast-moving bodies) or when more than
 * 2 % of the candidate lists overflow. A tuning knob without a reference counterpart (Embree refits at
 * every detection, M/CollisionDetector/DiscreteCollisionDetector.cpp): contact sets do not depend on it, queries test the
 * primitives themselves. Before finalize. */
int gaia_b200_set_lazy_refit_margin(gaia_handle* h, float edges);
/* Bytes of the solver's HBM layout, for the bench's "inputs vs L2" note and for capacity planning:
 * out[0] per-tet record stream, out[1] per-neighbour entry stream, out[2] per-vertex state read or written by one iteration
 * (headers, positions, inertia, previous positions), out[3] tile table; out[4] = shared-memory position reads per
 * iteration (3 per strip restart, 1 per strip step), out[5] = incidences (vertex, tet) per iteration. */
int gaia_b200_get_layout_bytes(gaia_handle* h, int64_t out[6]);

/* Vertex graph colouring (host code; no device needed) -------------------------
 * The reference colours its meshes with an offline tool (S/GraphColoring/GraphColoring.cpp: `-v -a greedy -b -r ratio`;
 * OrderedGreedy, M/GraphColoring/gready.cpp:16-117; balanceColoredCategories, M/GraphColoring/ColoringAlgorithms.cpp:82-198)
 * whose output is the `*.vertexColoring.json` that add_tet_mesh / add_tri_mesh take. These calls restate both of the tool's
 * algorithms - `-a greedy` (OrderedGreedy) and `-a mcs` (Mcs::color, M/GraphColoring/mcs.cpp:14-168, the one the colourings
 * shipped under Data/mesh_models were made with: 8 colours) - with the same node order, colour choice and balancing moves,
 * in O(E log V) instead of the tool's O(V^2) scans. Pinned: colours[] are identical to the tool's own code (compiled into
 * oracle/_ref) on random graphs and on the vertex graphs of reference meshes, before and after balancing
 * (tests/test_coloring.py, fixtures tests/golden/ref_coloring.npz). One documented difference: the tool's balancing loop is
 * unbounded and can shuttle a node between two colours for ever; here it stops after 4 V + 16 moves (a colouring is valid
 * after every move).
 * `balance` is a flag word: GAIA_COLOR_BALANCE (1) balances the categories towards goalMaxMinRatio, GAIA_COLOR_MCS (2)
 * orders the nodes by maximum cardinality search instead of smallest remaining degree.
 * colors: numNodes colour ids (may be NULL); categoryOffsets (numColours + 1, sized numNodes + 1 by the caller) and
 * categoryNodes (numNodes): the colour categories in the tool's order (both NULL to skip). Return the number of colours.
 * color_graph takes the undirected edge list as given; the mesh variants build the vertex graph of TetMeshVertexGraph
 * (vertices sharing a tet) / TriMeshVertexGraph (vertices sharing a triangle, plus the two wing vertices of every interior
 * edge when edgeEnergy != 0: they share a bending hinge), with unique edges. */
int gaia_b200_color_graph(int32_t numNodes, int64_t nEdges, const int32_t* edges, int32_t balance, float goalMaxMinRatio,
                          int32_t* colors, int32_t* categoryOffsets, int32_t* categoryNodes);
int gaia_b200_color_tet_mesh(int32_t nV, int32_t nT, const int32_t* tets, int32_t balance, float goalMaxMinRatio,
                             int32_t* colors, int32_t* categoryOffsets, int32_t* categoryNodes);
int gaia_b200_color_tri_mesh(int32_t nV, int32_t nF, const int32_t* faces, int32_t edgeEnergy, int32_t balance, float goalMaxMinRatio,
                             int32_t* colors, int32_t* categoryOffsets, int32_t* categoryNodes);
/* The same contract computed ON THE DEVICE `device` (csrc/coloring_gpu.cu): deterministic speculative greedy (largest degree
 * first) + batched largest-to-smallest balancing; for graphs too large for an offline pass and for re-colouring generated or
 * partitioned meshes. A valid colouring with few, balanced colours - NOT the tool's colours (use the host calls above where
 * the tool's exact output matters). flags: GAIA_COLOR_BALANCE. At most 64 colours (GAIA_ERR_UNSUPPORTED beyond). */
int gaia_b200_color_graph_device(int device, int32_t numNodes, int64_t nEdges, const int32_t* edges, int32_t flags, float goalMaxMinRatio,
                                 int32_t* colors, int32_t* categoryOffsets, int32_t* categoryNodes);
/* Allocation-free variant of the call above. It has NO counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it). It runs the same deterministic speculative greedy (no balancing pass) and calls no cudaMalloc / cudaFree / allocating
 * library routine: its scratch is a file-scope device array in csrc/coloring_gpu.cu sized for GAIA_NOALLOC_COLOR_MAX_NODES
 * nodes (12 MB + 512 B, held by every GPU that loads the library until the process exits); more nodes -> GAIA_ERR_OVERFLOW.
 * Inputs and output are the CALLER's device memory on `device`: the undirected graph in CSR form (dOffsets [numNodes + 1],
 * dAdjacency [dOffsets[numNodes]], every edge listed from both ends, no self loops) and dColors [numNodes]. Returns the
 * number of colours (same colours as gaia_b200_color_graph_device with flags = 0) or a negative gaia_status; errMsg
 * (optional, errMsgCap bytes) receives the reason. Work is issued on the legacy default stream and complete on return.
 * All calls share the one scratch array: calls that overlap on the same GPU race - serialise them. */
#define GAIA_NOALLOC_COLOR_MAX_NODES (1 << 20)
int gaia_b200_color_graph_device_noalloc(int device, int32_t numNodes, const int32_t* dOffsets, const int32_t* dAdjacency,
                                         int32_t* dColors, char* errMsg, int32_t errMsgCap);

#ifdef __cplusplus
}
#endif
#endif /* GAIA_B200_H */
