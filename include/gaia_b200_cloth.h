/*
 * gaia_b200_cloth.h — C ABI of the B200-native VBD *cloth* step.
 *
 * Replaces, for triangle-mesh scenes, the call sites of
 *   VBDClothSimulationFramework::runStep_CDEveryIter          M/VBDCloth/VBDClothPhysics.cpp:389-518
 *   ClothContactDetector (VF / EE radius queries)              M/CollisionDetector/ClothContactDetector.{h,cpp}
 *   VBDTriMeshStVk (StVK membrane + quadratic bending)         M/VBDCloth/VBDTriMeshStVK.cpp:50-98, 721-982
 * Same conventions as gaia_b200.h (status codes, one host thread per handle, xyz-interleaved float32 in the ORIGINAL
 * vertex order of each mesh, no CPU fallback).
 *
 * Meshes: the first `numSimMeshes` meshes added with add_tri_mesh(simulated=1) are cloth (baseTriMeshesForSimulation);
 * meshes added with simulated=0 are colliders (colliderTriMeshes: every vertex fixed, positions may be replaced
 * between frames with set_positions) and must come after all simulated meshes, like triMeshesAll
 * (M/Framework/BaseClothSimPhsicsFramework.cpp:57-63).
 */
#ifndef GAIA_B200_CLOTH_H
#define GAIA_B200_CLOTH_H

#include "gaia_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* PhysicsParams + ContactDetectorParams of the cloth Parameters.json: BasePhysicsParams
 * (M/Parameters/PhysicsParameters.h:9-105), VBDClothPhysicsParameters (M/VBDCloth/VBDClothPhysicsParameters.h:5-80),
 * ClothContactDetectorParameters (M/CollisionDetector/ClothContactDetector.h:13-30). Same names, same defaults. */
typedef struct gaia_cloth_physics_params {
    float gravity[3];                       /* (0,-10,0) */
    float timeStep;                         /* 0.0166666 */
    int32_t numSubsteps;                    /* 20 */
    int32_t iterations;                     /* 1 */
    float worldBoundsMin[3];                /* -1e7 */
    float worldBoundsMax[3];                /* +1e7 */
    float boundaryFrictionDynamic;          /* 0.1 */
    /* VBDClothPhysicsParameters */
    float stepSize;                         /* 1.0 */
    int32_t associateGravityWithInertia;    /* 1 */
    int32_t applyAcceleration;              /* 0 */
    float acceleratorRho;                   /* 0.95 */
    int32_t handleCollision;                /* 1 */
    float contactStiffness;                 /* 1e4 */
    float contactRadius;                    /* 1.0 */
    float thickness;                        /* 0.0 */
    float conservativeStepRelaxation;       /* 0.45 */
    float boundaryContactStiffness;         /* 1e4 */
    float boundaryFrictionEpsValue;         /* 1e-2 */
    int32_t contactEnergyTypes;             /* 1: quadratic + logarithmic two stages (0: quadratic, 2: IPC-like) */
    int32_t contactBVHReconstructionSteps;  /* 32 */
    float degenerateTriangleThres;          /* 1e-6 */
    /* ClothContactDetectorParameters */
    float maxQueryDis;                      /* 0.4 */
} gaia_cloth_physics_params;

/* One entry of the cloth Models.json: ObjectParams (M/Materials/Materials.h:12-53), TriMeshParams
 * (M/TriMesh/TriMesh.cpp:449-463), VBDObjectParamsTriMesh (M/VBDCloth/VBDBaseTriMesh.h:8-20),
 * VBDObjectParamsTriMeshStVk (M/VBDCloth/VBDTriMeshStVK.h:12-22). Rest shape: use3DRestpose = true
 * (TriMeshFEM::computeRestposeTrianglesFrom3D, M/TriMesh/TriMesh.cpp:213-244). */
typedef struct gaia_cloth_object_params {
    float scale[3];                 /* (1,1,1) */
    float rotation[3];              /* axis*angle, (0,0,0) */
    float translation[3];           /* (0,0,0) */
    float density;                  /* 1 */
    int32_t hasNoGravZone;          /* 0 */
    float noGravZoneThreshold;      /* 0 */
    float maxVelocityMagnitude;     /* -1 */
    float frictionDynamic;          /* 0.1 */
    float frictionEpsV;             /* 0.1 */
    float bendingStiffness;         /* 0 */
    float exponentialVelDamping;    /* 1.0 (used as is: the cloth framework does not raise it to dt) */
    float constantVelDamping;       /* 0.0 */
    int32_t initializationType;     /* 0 inertia, 1 adaptive gravity, 2 previous position, 3 inertia truncated to the
                                       conservative bound, 4 keep (M/VBDCloth/VBDBaseTriMesh.cpp:77-163) */
    float miu;                      /* 1e4 */
    float lambda;                   /* 1e4 */
    float dampingStVK;              /* 0 */
    float dampingBending;           /* 0 */
} gaia_cloth_object_params;

/* Detected contacts, for bit-exact parity of the contact sets (VFContactPointInfo / EEContactPointInfo,
 * M/CollisionDetector/ClothContactDetector.h:32-103). VF list sorted by (vertexMesh, vertex, faceMesh, face);
 * EE list by (mesh1, edge1, mesh2, edge2). Edge ids are the reference's (creation order of MeshFrame edges,
 * M/TriMesh/TriMesh.cpp:499-560). */
typedef struct gaia_cloth_vf_contact {
    int32_t vertexMesh, vertex, faceMesh, face;
    int32_t primitiveType;   /* ClosestPointOnPrimitiveType: 0 vertex, 1 edge, 2 face */
    int32_t primitiveId;     /* vertex / edge / face id the closest point lies on */
} gaia_cloth_vf_contact;
typedef struct gaia_cloth_ee_contact {
    int32_t mesh1, edge1, mesh2, edge2;
} gaia_cloth_ee_contact;

typedef struct gaia_cloth_stats {
    double timeCsmpUpdatingCollisionInfoDCD;   /* detection phases (BVH + queries + bounds), ms, profiling mode */
    double timeCsmpInitialStep;
    double timeCsmpMaterialSolve;
    double solveKernelMs;           /* colour-solve kernels alone */
    int64_t vertexSolves;
    int64_t colourSweeps;
    int64_t substeps;
    int64_t frames;
    int64_t kernelLaunches;
    int64_t algorithmicBytes;       /* SURVEY §8(d) cloth formula, summed over colour sweeps */
    int64_t numberCollisionDetection;   /* RunningTimeStatistics::numberCollisionDetection */
    int64_t numContactsVF;          /* last detection */
    int64_t numContactsEE;
} gaia_cloth_stats;

typedef struct gaia_cloth_handle gaia_cloth_handle;

void gaia_b200_cloth_default_physics_params(gaia_cloth_physics_params* p);
void gaia_b200_cloth_default_object_params(gaia_cloth_object_params* p);
const char* gaia_b200_cloth_last_error(const gaia_cloth_handle* h);

/* create: BaseClothPhsicsFramework::parseRunningParameters + timeStepEqualize */
int gaia_b200_cloth_create(const gaia_cloth_physics_params* phys, int device, gaia_cloth_handle** out);
void gaia_b200_cloth_destroy(gaia_cloth_handle* h);

/* add_tri_mesh: TriMeshFEM::initialize (loadObj is the caller's: xyz = 3*nV floats, faces = 3*nF vertex ids) +
 * TriMeshTopology::initialize (M/TriMesh/TriMesh.cpp:15-111, 481-753) + VBDTriMeshStVk::initialize. Colouring as in
 * gaia_b200_add_tet_mesh (ignored for colliders). Returns the mesh id. */
int gaia_b200_cloth_add_tri_mesh(gaia_cloth_handle* h, int32_t simulated, int32_t nV, int32_t nF,
                                 const float* xyz, const int32_t* faces, const gaia_cloth_object_params* obj,
                                 int32_t nColors, const int32_t* color_offsets, const int32_t* color_verts,
                                 int32_t nFixed, const int32_t* fixedPoints);
/* Deformers of the cloth Parameters.json (loadClothDeformers, M/VBDCloth/VBDClothDeformer.h:202-242), applied at the first
 * iteration of every substep (VBDClothSimulationFramework::applyDeformers, M/VBDCloth/VBDClothPhysics.cpp:408-412, 2464-2470).
 * Call after add_tri_mesh of `mesh` and before finalize; the selected vertices become fixed points (setFixedPointCloth).
 * Rotator (:43-132): p = root + R(curTime * angularVelocity) * RootP with root = ((P - center).axis) axis and
 * RootP = P - root exactly as the reference computes them (so the turn is about the axis through the origin and `center`
 * only enters through its component along the axis); rotationEndTime <= 0: never stops. Translator (:134-200): p += speed * dt while deformationStartTime <= curTime < deformationEndTime (end < 0:
 * never stops). Return the deformer id. */
int gaia_b200_cloth_add_rotator(gaia_cloth_handle* h, int mesh, const int32_t* verts, int32_t n, const float center[3],
                                const float axis[3], double angularVelocity, double rotationEndTime);
int gaia_b200_cloth_add_translator(gaia_cloth_handle* h, int mesh, const int32_t* verts, int32_t n, const float speed[3],
                                   float deformationStartTime, float deformationEndTime);
/* Key-frame sequence of a collider mesh (ColliderParams.ColliderMeshes[i] with colliderType "TriMeshSequence":
 * ColliderTrimeshSequenceParams / ColliderTrimeshSequence, M/SpatialQuery/ColiiderTriMeshSequence.h:4-151). `mesh` was added
 * with simulated = 0; xyz holds nKeys meshes of its topology (nKeys * 3*nV floats, transformed like the mesh); keyFrames are
 * increasing frame ids (NULL: 0, 1, 2, ...). Every iteration the collider is moved to prev * (1 - t) + next * t with the
 * reference's t (per frame; per substep when interpolateSubstep; per iteration for the first interpolateIter iterations),
 * and a moved collider triggers collision detection (VBDClothPhysics.cpp:413-418). Before finalize. */
int gaia_b200_cloth_set_collider_sequence(gaia_cloth_handle* h, int mesh, int32_t nKeys, const int32_t* keyFrames, const float* xyz,
                                          int32_t interpolateSubstep, int32_t interpolateIter);
/* Sewing of a simulated mesh (ObjectParams::sewingPath = {"vertices": [[v0, v1, v2], ...], "ratios": [...]} and
 * ::sewingStiffness, M/TriMesh/TriMesh.cpp:95-110, 458; VBDBaseTriMesh::accumlateSewingForceAndHessian,
 * M/VBDCloth/VBDBaseTriMesh.cpp:320-344, called from the vertex solve at VBDClothPhysics.cpp:1112-1114): seam i pulls vertex
 * v0 onto the point (1 - ratio_i) x_v1 + ratio_i x_v2 of the SAME mesh; every vertex of the seam gets the force
 * -m k (x_v0 - (1 - ratio) x_v1 - ratio x_v2) and the Hessian m^2 k I with m = 1, -(1 - ratio), -ratio for v0, v1, v2.
 * Replaces the mesh's seams; before finalize. The colouring passed to add_tri_mesh must keep the three vertices of a seam in
 * different colours (as for the reference, whose colour files are an input). */
int gaia_b200_cloth_set_sewing(gaia_cloth_handle* h, int mesh, int32_t nSeams, const int32_t* verts3, const float* ratios, float sewingStiffness);
/* Contact pairs the seams of a mesh exclude (ObjectParams::filteredPairsPath = {"vf": [[vertex, face], ...], "ee": [[edge, edge],
 * ...]}, TriMeshFEM::generate_sewing_filters, M/TriMesh/TriMesh.h:596-622): a listed vertex never pairs with the listed face,
 * a listed edge never with the listed edge (either way round), of the SAME mesh - neither as a contact nor in the minimum
 * distances behind the conservative bounds (the ...WithSewing callbacks return before measuring, M/CollisionDetector/
 * ClothContactDetector.cpp:290-320, 815-852). Edge ids are the reference's (gaia_b200_cloth_get_edges). As in the reference the
 * lists only take effect for a mesh that also has seams (:969-974, 1011-1016). Replaces the mesh's lists; before finalize. */
int gaia_b200_cloth_set_filtered_pairs(gaia_cloth_handle* h, int mesh, int32_t nVF, const int32_t* vfPairs, int32_t nEE, const int32_t* eePairs);
/* One vertex-face attachment (PhysicsParams.VertexFaceAttachmentConstraintParams: vertexConstraintVertexInfos,
 * vertexConstraintTargetFacesInfos, vertexConstraintBarys, vertexConstraintStiffness; M/VBDCloth/VBDTriMeshConstraints.h:6-52,
 * set-up VBDClothPhysics.cpp:152-220, force ::accumulateConstraintForceAndHessian :1464-1498): vertex `vertex` of the simulated
 * mesh `meshV` is tied with `stiffness` (the reference's default: 1e4) to the point bary . (three vertices faceVerts of mesh
 * `meshF`), or, with meshF = -1, to the fixed position baryOrPos. When meshF is simulated too, each face vertex whose bary is
 * not 0 receives the reaction, computed exactly as the reference does: from x(meshV, THAT face vertex's id) - attached point
 * (the reference reads the V-side mesh with the id of the vertex being solved, :1472), so such ids must exist in meshV.
 * Before finalize; the vertices an attachment couples must have different colours. */
int gaia_b200_cloth_add_attachment(gaia_cloth_handle* h, int meshV, int32_t vertex, int meshF, const int32_t faceVerts[3],
                                   const float baryOrPos[3], float stiffness);
/* finalize: VBDClothSimulationFramework::initialize (parallel groups :72-138, contact detector) */
int gaia_b200_cloth_finalize(gaia_cloth_handle* h);

/* run_frames: BasePhysicFramework::simulate loop body with runStep = runStep_CDEveryIter */
int gaia_b200_cloth_run_frames(gaia_cloth_handle* h, int32_t nFrames);
/* parity entry: one substep's first `nIters` iterations (detection, initial guess, colour sweeps), no velocity update */
int gaia_b200_cloth_run_iterations(gaia_cloth_handle* h, int32_t nIters);
int gaia_b200_cloth_synchronize(gaia_cloth_handle* h);
/* frameId / curTime of the framework (curTime += dt every substep; the deformers' clock); set_time: ClothPhysicsState::
 * initializePhysics when a run is recovered (M/Framework/BaseClothSimPhsicsFramework.h:94-96) */
int gaia_b200_cloth_frame_id(const gaia_cloth_handle* h);
double gaia_b200_cloth_cur_time(const gaia_cloth_handle* h);
int gaia_b200_cloth_set_time(gaia_cloth_handle* h, int32_t frameId, double curTime);

int gaia_b200_cloth_num_meshes(const gaia_cloth_handle* h);
int gaia_b200_cloth_mesh_sizes(const gaia_cloth_handle* h, int mesh, int32_t* nV, int32_t* nF, int32_t* nE);
int gaia_b200_cloth_get_positions(gaia_cloth_handle* h, int mesh, float* xyz);
int gaia_b200_cloth_get_velocities(gaia_cloth_handle* h, int mesh, float* xyz);
int gaia_b200_cloth_set_state(gaia_cloth_handle* h, int mesh, const float* pos, const float* vel);   /* vel may be NULL */
int64_t gaia_b200_cloth_total_vertices(const gaia_cloth_handle* h);
int gaia_b200_cloth_get_all_positions(gaia_cloth_handle* h, float* xyz);
int gaia_b200_cloth_set_all_state(gaia_cloth_handle* h, const float* pos, const float* vel);

/* integer parity surfaces */
int gaia_b200_cloth_get_edges(const gaia_cloth_handle* h, int mesh, int32_t* out);   /* per edge: eV1,eV2,eV12Next,eV21Next,fId1,fId2 */
int gaia_b200_cloth_num_parallel_groups(const gaia_cloth_handle* h);
int gaia_b200_cloth_get_parallel_group(const gaia_cloth_handle* h, int group, int32_t* out, int32_t* nPairs);
int gaia_b200_cloth_get_vertex_masses(const gaia_cloth_handle* h, int mesh, float* out);
/* applyCollisionDetection on the current state without stepping; fills the lists below and the conservative bounds */
int gaia_b200_cloth_detect_contacts(gaia_cloth_handle* h);
int gaia_b200_cloth_get_vf_contacts(gaia_cloth_handle* h, gaia_cloth_vf_contact* out, int64_t capacity, int64_t* n);
int gaia_b200_cloth_get_ee_contacts(gaia_cloth_handle* h, gaia_cloth_ee_contact* out, int64_t capacity, int64_t* n);
int gaia_b200_cloth_get_conservative_bounds(gaia_cloth_handle* h, int mesh, float* out);   /* vertexConvervativeBounds */

int gaia_b200_cloth_get_stats(gaia_cloth_handle* h, gaia_cloth_stats* out);
int gaia_b200_cloth_reset_stats(gaia_cloth_handle* h);
int gaia_b200_cloth_set_profiling(gaia_cloth_handle* h, int enabled);
void* gaia_b200_cloth_cuda_stream(gaia_cloth_handle* h);

#ifdef __cplusplus
}
#endif
#endif /* GAIA_B200_CLOTH_H */
