// Internal scene representation of libgaia_b200.so (host side + device mirrors).
// Layout in HBM: all meshes are merged into ONE colour-sorted vertex array; a vertex keeps a
// "sorted id" s = position in [group 0 | group 1 | ...], groups being the reference's merged parallel groups
// (M/VBD/VBDPhysics.cpp:236-274). Every per-vertex array is SoA over s; per-incidence (vertex,tet) records are
// stored contiguously per vertex in two float4 streams so that a colour sweep reads them with coalesced 16-byte loads.
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/gaia_b200.h"

namespace gb {

struct HostMesh {
    int id = 0, nV = 0, nT = 0;
    gaia_object_params obj;            // damping already made step-invariant
    std::vector<float> pos0;           // 3*nV transformed rest positions (original order)
    std::vector<int32_t> tets;         // 4*nT
    std::vector<float> restVol;        // nT  (det(Dm)/6, float as the reference computes it)
    std::vector<float> mass;           // nV
    std::vector<uint8_t> fixedMask;
    std::vector<uint8_t> ghostMask;    // owned by another partition (multi-GPU): never solved here, overwritten by the halo
    std::vector<int32_t> fixedPoints;
    std::vector<std::vector<int32_t>> colors;   // size-descending (stable)
    std::vector<int32_t> groupOfVertex;
    // surface topology (reference orderings, see host_setup.cpp)
    std::vector<int32_t> surfaceVIds, tetVertToSurf, surfFaces, surfFaceTet, surfFaceIdAtTet, surfFace3Nbr, tetNbrTets;
    std::vector<int32_t> surfNbrOff, surfNbrVerts;   // CSR: neighbour surface vertices (tet-mesh ids) of each surface vertex
    // placement in the global arrays
    int64_t vertBase = 0;              // offset of this mesh in "original order, meshes concatenated"
    int64_t tetBase = 0, surfVBase = 0, surfFBase = 0;
    std::vector<int32_t> sortedId;     // nV: original vertex -> sorted id
    bool activeForCollision = true, activeForMaterialSolve = true;
    int64_t nSolved = 0;               // non-fixed vertices
    int64_t nIncSolved = 0;            // incidences of non-fixed vertices
    std::vector<int64_t> nbrDistinctPerGroup;   // |N(c)| restricted to this mesh, per global group
    int nSurfV() const { return (int)surfaceVIds.size(); }
    int nSurfF() const { return (int)surfFaces.size() / 3; }
};

// per-mesh constants read by the kernels (small table in global memory, L1/L2 resident)
struct MeshParamsDev {
    float miu, lmbd, a;                // a = 1 + miu/lmbd
    float dampH, dampD;                // dampingHydrostatic, dampingDeviatoric
    float frictionDynamic, frictionEpsV;
    float expVelDamp, constVelDamp, maxVel, noGravThreshold;
    int32_t hasNoGravZone, initType;
    float initRatio_g;
    int32_t activeSolve, activeCollide, hasVelPrev, hasApproxAccel;
    int32_t vertBase, tetBase, surfVBase, surfFBase;   // bases into the global (original-order) arrays
    int32_t pad0, pad1;
};

struct StepParamsDev {
    float dt, dtSqrReciprocal;
    float gravity[3];
    float boundsMin[3], boundsMax[3];
    float kBoundary, muBoundary, epsVBoundary;
    float stepSize, stepSizeGD;
    float kCollision;
    int32_t usePlaneGround, associateGravityWithInertia;
    int32_t useBowlGround, bowlCap;      // BasePhysicsParams::useBowlGround / bowlCap / bowlCenter / bowlRadius (M/Parameters/PhysicsParameters.h:18, 40-45)
    float bowlCenter[3], bowlRadius;
};

struct Scene;   // defined in scene_impl.h (needs CUDA types)

}  // namespace gb
