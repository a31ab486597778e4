// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// Scene/mesh state of the CPU restatement of the reference VBD step.
#pragma once
#include "oracle_common.h"
#include "../include/gaia_b200.h"

namespace orc {

// CollidingPointInfo + CollisionDetectionResult + VBDCollisionInfo
// (M/CollisionDetector/CollisionDetertionParameters.h:166-222, M/VBD/VBD_CollisionInfo.h:12-22),
// restricted to maxNumberOfCollisions == 1 (the reference default, PREALLOCATED_NUM_COLLISIONS 1).
struct ColResult {
    int numIntersections = 0;
    int idVQuery = -1, idTMQuery = -1;
    bool fromCCD = false;
    // CollidingPointInfo
    V3 closestSurfacePt, closestSurfacePtBarycentrics, closestPointNormal;
    int intersectedElement = -1, intersectedMeshId = -1, closestSurfaceFaceId = -1;
    int closestPointType = 7;  // NotFound
    bool shortestPathFound = false;
    // VBDCollisionInfo
    float penetrationDepth = 0.f;
    V3 collisionForce;
    M3 collisionHessian;
    void clear() { numIntersections = 0; fromCCD = false; shortestPathFound = false; closestSurfaceFaceId = -1;
                   intersectedElement = -1; intersectedMeshId = -1; closestPointType = 7; }
};

// CollisionRelation (M/VBD/ActiveCollisionList.h:9-18)
struct ColRelation { int meshId, surfaceVertexId, collisionId, collisionType, collisionVertexOrder; };

struct Mesh {
    int id = 0;
    int nV = 0, nT = 0;
    gaia_object_params obj;
    // state: xyz interleaved (Eigen 3xN column-major)
    std::vector<V3> pos, prev, vel, velPrev, inertia, accel;
    std::vector<int> tets;            // 4*nT
    std::vector<float> DmInv;         // 9*nT, each 3x3 column-major (M/TetMesh/TetMeshFEM.h:410-413)
    std::vector<float> restVol, mass;
    std::vector<uint8_t> fixedMask;
    std::vector<int> fixedPoints;
    // CSR vertex -> tets (M/TetMesh/TetMeshFEM.cpp:275-325): infos[2v]=start, [2v+1]=count
    std::vector<int> vNbrInfos, vNbrTets, vNbrOrder;
    std::vector<std::vector<int>> colors;   // verticesColoringCategories, size-descending
    std::vector<int> vertexParallelGroup;   // group of each vertex
    // surface topology (M/TetMesh/TetMeshFEM.cpp:136-221)
    std::vector<int> surfaceVIds, tetVertToSurf;
    std::vector<int> surfFaces;             // 3*nF tet-mesh vertex ids
    std::vector<int> surfFaceTet, surfFaceIdAtTet, surfFace3Nbr; // nF, nF, 3*nF
    std::vector<int> tetNbrTets;            // 4*nT
    std::vector<int> tetXorSums;            // nT
    std::vector<std::vector<int>> surfVertNbrSurfVerts; // per surface vertex: neighbour tet-mesh vertex ids
    std::vector<std::vector<int>> surfVertNbrSurfFaces;
    // flags
    bool activeForCollision = true, activeForMaterialSolve = true;
    bool hasVelocitiesPrev = false, hasApproxAcceleration = false;
    std::vector<uint8_t> activeCollisionMask, penetratedMask;
    // collision results per surface vertex, relations per vertex
    std::vector<ColResult> colResults;
    std::vector<std::vector<ColRelation>> colRelations;
    int nSurfV() const { return (int)surfaceVIds.size(); }
    int nSurfF() const { return (int)surfFaces.size() / 3; }
};

// DeformerRotator / DeformerTranslater (M/VBD/VBDDeformer.h:45-177)
struct Deformer {
    int type = 0, mesh = 0;
    std::vector<int> verts;
    V3 origin, axis, speed;
    double angularVelocity = 0, endTime = -1;
    std::vector<V3> roots, rootPs;   // translator: original positions in roots
};

struct Scene {
    gaia_physics_params phys;
    gaia_collision_params col;
    float dt = 0.f, dtSqrReciprocal = 0.f;
    std::vector<Mesh> meshes;
    std::vector<std::vector<int>> vertexParallelGroups;  // [mesh,v,mesh,v,...]
    // ActiveCollisionList (M/VBD/ActiveCollisionList.h:20-52): per group [mesh,v,...]
    std::vector<std::vector<int>> activeCollisionsEachGroup;
    bool finalized = false;
    bool keepColourOrder = false;   // partition of a multi-GPU scene: colours stay in the given (global) order
    // APPLY_LOCAL_LINE_SEARCH (M/VBD/VBDPhysics.cpp:25, compiled out in the reference): backtracking on the vertex merit energy
    struct { bool enabled = false; float tau = 0.5f, tauGD = 0.1f; int maxIters = 16; } lineSearch;   // VBDPhysicsParameters.h:24-29
    int64_t lineSearchShrinks = 0;  // trial steps rejected so far (diagnostic)
    struct { V3 center; float radius = 1.f; bool cap = false; } bowl;   // BasePhysicsParams::bowlCenter / bowlRadius / bowlCap (M/Parameters/PhysicsParameters.h:18, 40-45), used when phys.useBowlGround
    int frameId = 0, substep = 0, iIter = 0;
    float curTime = 0.f;
    std::vector<Deformer> deformers;
    Pool* pool = nullptr;
    // optional schedule trace of run_step (tests; single-threaded scenes only): events of 4 ints {code, a, b, c} -
    // 1 dcd, 2 initial step (mesh), 3 deformers, 4 ccd, 5 prepare, 6 update collision info (mesh, vertex), 7 vertex solve (mesh, vertex,
    // friction), 8 intermediate detection, 9 velocity update, 11 active-list entry after a 5 / 8 (group, mesh, vertex)
    std::vector<int32_t>* trace = nullptr;
    void ev(int code, int a = 0, int b = 0, int c = 0) { if (trace) { trace->push_back(code); trace->push_back(a); trace->push_back(b); trace->push_back(c); } }
    gaia_stats stats;
    std::string err;
    // collision acceleration (oracle_collision.cpp)
    void* bvhTet = nullptr;   // all-tets scene
    void* bvhSurf = nullptr;  // per-mesh surface scenes
    ~Scene();
};

// oracle_mesh.cpp
void mesh_initialize(Mesh& m, int nV, int nT, const float* xyz, const int32_t* tets, const gaia_object_params& obj,
                     int nColors, const int32_t* colorOffsets, const int32_t* colorVerts, int nFixed, const int32_t* fixed,
                     bool keepColourOrder = false);
void scene_finalize(Scene& s);
// oracle_step.cpp
void run_step(Scene& s);                  // VBDPhysics::runStep_hybridCollisionHandling, one frame
void run_iterations(Scene& s, int nIters); // initial step + nIters colour sweeps
// phase-level stepping (multi-process tests drive the colour loop themselves and exchange ghosts between phases)
void phase_begin_substep(Scene& s);       // curTime += dt, dcd, initial step, deformers
void phase_prepare(Scene& s);             // ccd + prepareCollisionData
void phase_solve_group(Scene& s, int g, bool applyFriction);
void vertex_merit_energy_parts(const Scene& s, const Mesh& m, int iV, const V3& x, float& inertia, float& elastic);   // evaluateVertexMeritEnergy
bool tetrahedral_traverse_to(const Mesh& m, V3 rayOrigin, V3 rayDir, float maxDis, int startTet, int startFace, int targetTet, float eps,
                             int* numTraversed);      // TetMeshFEM::tetrahedralTraverseTo (oracle_collision.cpp)
bool probe_feasible_region(const Mesh& m, V3 p, int faceId, int type, float eps);   // DiscreteCollisionDetector::checkFeasibleRegion
void phase_intermediate_cd(Scene& s);    // dcd (intermediate) + ccd + prepareCollisionData, every intermediateCollisionIterations
void phase_end_substep(Scene& s);         // velocities
void enable_models(Scene& s);
void neohookean_force_hessian(const Scene& s, const Mesh& m, int iV, V3& force, M3& h);
V3 neohookean_force(const Mesh& m, int iV);
void inertia_force_hessian(float mass, V3 inertia, V3 x, float dtSqrReciprocal, V3& force, M3& h);
float accelerator_omega(int order, float rho, float prevOmega);
V3 accelerate_position(V3 y, V3 yPrevPrev, float omega);
int vertex_step_direction(float stepSizeNewton, float stepSizeGD, V3 force, const M3& h, V3& dir, float& stepSize, bool& solverOk);      // accumlateMaterialForce: elastic force alone
bool solve3x3_psd_stable(const float* m, const float* b, float* out);
void vertex_friction(float mu, float lambda, const float T[6], const float u[2], float epsU, V3& force, M3& hessian);
// oracle_collision.cpp
void collision_initialize(Scene& s);
void collision_free(Scene& s);
void dcd(Scene& s, bool intermediate);
void ccd(Scene& s, bool intermediate = false);
void prepare_collision_data(Scene& s);
void update_collision_info(Scene& s, ColResult& r);

}  // namespace orc
