// Host-side scene of the cloth path (include/gaia_b200_cloth.h): triangle meshes, their topology in the reference's
// orderings, and the colour-sorted packing the kernels of cloth.cu read.
//
// Layout in HBM (cloth): all vertices of all meshes live in ONE array in "sorted" order
//   [ parallel group 0 | group 1 | ... | collider vertices ],
// a group being one merged colour (VBDClothSimulationFramework::initializeParallelGroups, M/VBDCloth/VBDClothPhysics.cpp:72-138),
// so that a colour sweep is a contiguous range. The per-(vertex, incident face) and per-(vertex, relevant hinge) records
// are stored PER COLOUR in ELL form: record k of the i-th vertex of the colour sits at base + k * |colour| + i, which makes
// the one-thread-per-vertex sweep read them with fully coalesced 16-byte loads and needs no reduction across threads.
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "../../include/gaia_b200_cloth.h"

namespace gbc {

struct ClothMesh {
    int id = 0, nV = 0, nF = 0, nE = 0;
    bool simulated = true;
    gaia_cloth_object_params obj;
    std::vector<float> pos0;                 // 3*nV, transformed rest positions (original order)
    std::vector<int32_t> faces;              // 3*nF
    std::vector<float> DmInv;                // 4*nF (2x2 column-major)
    std::vector<float> restArea, mass;
    std::vector<uint8_t> fixedMask;
    std::vector<int32_t> edges;              // 6*nE: eV1, eV2, eV12Next, eV21Next, fId1, fId2 (EdgeInfo, M/TriMesh/TriMesh.h:51-60)
    std::vector<int32_t> face3Edges;         // 3*nF
    std::vector<int32_t> vNbrVOff, vNbrV;    // CSR neighbour vertices
    std::vector<int32_t> vNbrFOff, vNbrF, vNbrFOrder;
    std::vector<int32_t> vNbrEOff, vNbrE, vNbrEOrder;
    std::vector<int32_t> vBendOff, vBend, vBendOrder;
    std::vector<float> Q;                    // 16*nE: edgeLaplacianQuadraticForms, row-major 4x4
    std::vector<std::vector<int32_t>> colors;
    std::vector<int32_t> sortedId;           // original vertex -> sorted id
    // sewing (M/TriMesh/TriMesh.h:226-227 + ObjectParams::sewingStiffness): seam i = (v0, v1, v2) with ratio i
    std::vector<int32_t> sewVerts;           // 3 * nSeams
    std::vector<float> sewRatios;
    float sewStiffness = 0.f;
    std::vector<int32_t> filtVF, filtEE;     // filteredPairsPath: (vertex, face) and (edge, edge) pairs, 2 ints each
    int64_t vBase = 0, fBase = 0, eBase = 0; // offsets in "meshes concatenated" order
};

// ClothDeformerRotator / ClothDeformerTranslater (M/VBDCloth/VBDClothDeformer.h:43-200): selected vertices become fixed and
// are moved by the host schedule at the first iteration of every substep
struct ClothDeformer {
    int type = 0;                      // 0 rotator, 1 translator
    int mesh = 0;
    std::vector<int32_t> verts;
    float o[3] = {0, 0, 0}, axis[3] = {0, 0, 0}, speed[3] = {0, 0, 0};
    double angularVelocity = 0.0, rotationEnd = -1.0;
    float deformationStartTime = 0.f, deformationEndTime = -1.f;
    std::vector<float> roots, rootPs;  // rotator: 3 per selected vertex, from the positions at creation
};

// ColliderTrimeshSequence (M/SpatialQuery/ColiiderTriMeshSequence.h:4-151): key-frame meshes of one collider, interpolated
// per frame / substep / iteration
struct ColliderSequence {
    int mesh = 0;
    std::vector<int32_t> keyFrames;
    std::vector<float> keyPos;         // nKeys * 3*nV, transformed like the mesh
    bool interpolateSubstep = true;
    int interpolateIter = 1;
    bool firstTime = true;
};

// one entry of VertexFaceAttachmentConstraintParams (M/VBDCloth/VBDTriMeshConstraints.h:6-29)
struct ClothAttachment {
    int meshV = 0, vertex = 0, meshF = -1;
    int32_t face[3] = {-1, -1, -1};
    float pos[3] = {0, 0, 0};          // barycentrics (meshF >= 0) or the fixed position (meshF = -1)
    float stiffness = 1e4f;
};

struct ClothHost {
    gaia_cloth_physics_params phys;
    std::vector<ClothAttachment> attachments;
    std::vector<ClothDeformer> deformers;
    std::vector<ColliderSequence> colliderSeqs;
    float dt = 0.f, dtSqrReciprocal = 0.f;
    std::vector<ClothMesh> meshes;
    int numSim = 0;
    std::vector<std::vector<int32_t>> groups;     // [mesh, v, ...]
    std::vector<int32_t> groupStart;              // G + 1 sorted-id offsets; groupStart[G] = number of simulated vertices
    std::vector<int32_t> sortedToOrig;            // sorted id -> global original index (vBase + v)
    int64_t N = 0, nSimV = 0, nF = 0, nE = 0, nSimE = 0, nSimF = 0;
    // per group: ELL widths and SURVEY §8(d) algorithmic bytes of one sweep
    std::vector<int32_t> ellF, ellH;
    std::vector<int64_t> ellFBase, ellHBase;      // record offsets
    std::vector<int64_t> groupSolved, groupAlgBytes;
    int64_t nRecF = 0, nRecH = 0;
};

void cloth_transform_points(const gaia_cloth_object_params& obj, int32_t nV, std::vector<float>& pos);
void cloth_mesh_init(ClothMesh& m, bool simulated, int32_t nV, int32_t nF, const float* xyz, const int32_t* faces,
                     const gaia_cloth_object_params& obj, int32_t nColors, const int32_t* colorOffsets, const int32_t* colorVerts,
                     int32_t nFixed, const int32_t* fixed);
bool cloth_host_finalize(ClothHost& s, std::string& err);

}  // namespace gbc
