// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// State of the CPU restatement of the reference VBD *cloth* step (M/VBDCloth, M/TriMesh,
// M/CollisionDetector/ClothContactDetector.*). Parity status: the per-vertex arithmetic (membrane + bending force / Hessian,
// bending forms, VF / EE contact force / Hessian with friction, world box, contact energy derivatives, segment distances) is
// PINNED to the reference's own lines compiled into oracle/_ref (tests/test_oracle_vs_ref.py; ref_shim/refcloth*.cpp);
// mesh set-up, contact detection and the step loop are UNPINNED - their reference code needs Embree / MeshFrame2 (absent here)
// and ships no golden vectors - and are checked against closed-form properties and brute force in tests/test_oracle_cloth.py.
#pragma once
#include "oracle_common.h"
#include "../include/gaia_b200_cloth.h"

namespace orc {

// EdgeInfo (M/TriMesh/TriMesh.h:51-60)
struct ClothEdge { int eV1, eV2, eV12Next, eV21Next, fId1, fId2; };

// ClosestPointOnTriangleType (M/CollisionDetector/CollisionDetertionParameters.h): AtA..NotFound as in the tet oracle
enum { cAtA = 0, cAtB, cAtC, cAtAB, cAtBC, cAtAC, cAtInterior, cNotFound };

// VFContactPointInfo (M/CollisionDetector/ClothContactDetector.h:32-49)
struct VFContact {
    int vMesh, v, fMesh, f;
    int closestType, primType, primId;
    float d;
    V3 cp, bary;
};
// EEContactPointInfo (:86-103)
struct EEContact {
    int m1, e1, m2, e2;
    float mu1, mu2;
    V3 c1, c2;
    float d;
};
// FVContactInfo (:79-84)
struct FVLink { int vMesh, v, contactId; };

// VertexFaceAttachmentConstraint (M/VBDCloth/VBDTriMeshConstraints.h:31-45)
struct Attachment {
    float stiffness = 1e4f;
    int vertexOrder = -1;                   // 0-2: face side, 3: vertex side
    int vertexId = -1, meshV = -1, meshF = -1;
    int face[3] = {-1, -1, -1};
    V3 attachedPos;                         // barycentrics (meshF >= 0) or a fixed position (meshF = -1)
};

struct TriMesh {
    int id = 0, nV = 0, nF = 0, nE = 0;
    bool simulated = true;
    gaia_cloth_object_params obj;
    std::vector<V3> pos, prev, vel, velPrev, inertia, accel, next, posAtCD, prevIter;
    std::vector<int> faces;                 // 3*nF (facePos)
    std::vector<float> DmInv;               // 4*nF, each 2x2 column-major
    std::vector<float> restArea, mass;
    std::vector<uint8_t> fixedMask;
    std::vector<int> fixedPoints;
    // TriMeshTopology (M/TriMesh/TriMesh.cpp:481-753)
    std::vector<ClothEdge> edges;
    std::vector<int> face3Edges, face3NbrFaces;          // 3*nF each
    std::vector<int> vNbrVOff, vNbrV;                    // CSR: neighbour vertices
    std::vector<int> vNbrFOff, vNbrF, vNbrFOrder;        // CSR: incident faces + order of the vertex in the face
    std::vector<int> vNbrEOff, vNbrE, vNbrEOrder;        // CSR: incident edges + order (0: eV1, 1: eV2)
    std::vector<int> vBendOff, vBend, vBendOrder;        // CSR: relevant bendings (edge id, order 0..3)
    std::vector<std::array<float, 16>> Q;                // edgeLaplacianQuadraticForms, row-major 4x4
    std::vector<std::vector<int>> colors;
    std::vector<int> globalColor;
    // contact state
    std::vector<float> bounds;              // vertexConvervativeBounds
    std::vector<uint8_t> activeCollisionMask;
    std::vector<std::vector<VFContact>> vf; // per vertex
    std::vector<float> vfMinDis;            // per vertex
    std::vector<std::vector<EEContact>> ee; // per edge (simulated meshes only)
    std::vector<float> eeMinDis;
    std::vector<float> faceMinDis;          // per face
    std::vector<std::vector<FVLink>> faceLinks;
    bool hasVelocitiesPrev = false, hasApproxAcceleration = false;
    // sewing (M/TriMesh/TriMesh.h:226-228; sewingVMap built as in generate_sewing_filters, :596-633) and attachment constraints
    std::vector<std::array<int, 3>> sewingVertices;
    std::vector<float> sewingRatio;
    float sewingStiffness = 0.f;
    std::vector<std::vector<std::pair<int, int>>> sewingVMap;      // per vertex: (seam, order)
    std::vector<std::vector<Attachment>> attachments;              // per vertex (VertexConstraints::attachmentConstraints)
    std::vector<std::vector<int>> sewingVF, sewingEE;              // filteredPairsPath: per vertex the faces, per edge the edges (sorted) not to be paired with
};

// ClothDeformerRotator / ClothDeformerTranslater (M/VBDCloth/VBDClothDeformer.h:43-200)
struct ClothDeformer {
    int type = 0;                           // 0 rotator, 1 translator
    int mesh = 0;
    std::vector<int> verts;
    V3 o, axis, speed;
    double angularVelocity = 0.0, rotationEnd = -1.0;
    float deformationStartTime = 0.f, deformationEndTime = -1.f;
    std::vector<V3> roots, rootPs;
};

// ColliderTrimeshSequence (M/SpatialQuery/ColiiderTriMeshSequence.h:4-151)
struct ColliderSequence {
    int mesh = 0;
    std::vector<int> keyFrames;
    std::vector<std::vector<V3>> keyPos;
    bool interpolateSubstep = true;
    int interpolateIter = 1;
    bool firstTime = true;
};

struct ClothScene {
    gaia_cloth_physics_params phys;
    std::vector<ClothDeformer> deformers;
    std::vector<ColliderSequence> colliderSeqs;
    float dt = 0.f, dtSqrReciprocal = 0.f;
    std::vector<TriMesh> meshes;
    int numSim = 0;
    bool finalized = false;
    bool bruteForce = false;                // contact candidates by brute force instead of the grid
    std::vector<std::vector<int>> groups;   // [mesh,v,...]
    bool collisionDetectionRequired = false;
    int frameId = 0, substep = 0, iIter = 0;
    float curTime = 0.f;
    gaia_cloth_stats stats;
    Pool* pool = nullptr;
    ~ClothScene() { delete pool; }
};

void cloth_mesh_initialize(TriMesh& m, bool simulated, int nV, int nF, const float* xyz, const int32_t* faces,
                           const gaia_cloth_object_params& obj, int nColors, const int32_t* colorOffsets, const int32_t* colorVerts,
                           int nFixed, const int32_t* fixed);
void cloth_finalize(ClothScene& s);
void cloth_add_rotator(ClothScene& s, int mesh, const int32_t* verts, int n, const float center[3], const float axis[3],
                       double angularVelocity, double rotationEnd);
void cloth_add_translator(ClothScene& s, int mesh, const int32_t* verts, int n, const float speed[3], float tStart, float tEnd);
void cloth_set_collider_sequence(ClothScene& s, int mesh, int nKeys, const int32_t* keyFrames, const float* xyz, bool interpolateSubstep, int interpolateIter);
void cloth_set_sewing(ClothScene& s, int mesh, int nSeams, const int32_t* verts3, const float* ratios, float stiffness);
int cloth_add_attachment(ClothScene& s, int meshV, int vertex, int meshF, const int32_t faceVerts[3], const float baryOrPos[3], float stiffness);
void cloth_set_filtered_pairs(ClothScene& s, int mesh, int nVF, const int32_t* vfPairs, int nEE, const int32_t* eePairs);
void cloth_sewing_force_hessian(const TriMesh& m, int iV, V3& force, M3& hessian);
void cloth_constraint_force_hessian(const ClothScene& s, int iMesh, int vId, V3& force, M3& hessian);
void cloth_transform_points(const gaia_cloth_object_params& obj, std::vector<V3>& pos);
void cloth_detect(ClothScene& s);
float cloth_accelerator_omega(float rho, int order, float prevOmega);
void cloth_conservative_bounds(const ClothScene& s, TriMesh& m);
void cloth_run_substeps(ClothScene& s, int nSubsteps);
void cloth_run_iterations(ClothScene& s, int nIters);
// probes for the unit tests
void cloth_material_force_hessian(const ClothScene& s, const TriMesh& m, int v, V3& force, M3& hessian);
void cloth_vertex_force_hessian(ClothScene& s, int mesh, int v, V3& force, M3& hessian);
float cloth_elastic_energy(const ClothScene& s, const TriMesh& m);
void cloth_contact_repulsion(const ClothScene& s, float dis, float& dEdD, float& d2E);
float cloth_probe_vf(const ClothScene& s, int vMesh, int v, int fMesh, int f, int vertexOrder, V3& force, M3& hessian);
float cloth_probe_ee(const ClothScene& s, int m1, int e1, int m2, int e2, int vertexOrder, V3& force, M3& hessian);
bool cloth_probe_feasible_region(const TriMesh& m, V3 p, int faceId, int type, float eps);
void cloth_probe_initial_step(ClothScene& s);
bool cloth_probe_update_colliders(ClothScene& s, int frameId, int substep, int iIter);
void cloth_probe_update_velocity(ClothScene& s);
bool cloth_probe_truncate(const ClothScene& s, int mesh, int v, V3& newPos);
void cloth_probe_boundary(const ClothScene& s, int mesh, int v, V3& force, M3& hessian);
V3 closest_point_triangle(V3 p, V3 a, V3 b, V3 c, V3& bary, int& type);   // oracle_collision.cpp
void closest_points_segments(V3 p0, V3 p1, V3 q0, V3 q1, V3& ps, V3& qt, float& s, float& t);   // oracle_cloth_contact.cpp
bool solve3x3_psd_stable(const float* m, const float* b, float* out);      // oracle_step.cpp

}  // namespace orc
