// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own tetrahedral traversal, compiled from where it lies under /root/reference:
//   TetMeshFEM::tetrahedralTraverseTo (the static-stack variant DCD calls first)     Simulator/Modules/TetMesh/TetMeshFEM.cpp:503-742
//   its stack sizes and face tables                                                  Simulator/Modules/TetMesh/TetMeshFEM.cpp:14-25
//   TetMeshFEM::exitFaceSelection, ::projectTo2DCoordinates, ::copyRepermutedVerts   Simulator/Modules/TetMesh/TetMeshFEM.h:420-489, 525-561
//   TetMeshFEM::originBarycentric2DTriangle                                          Simulator/Modules/TetMesh/TetMeshFEM.h:625-633
//   CPArrayStatic / CircularArray (MeshFrame2, vendored by the reference)            3rdParty/MeshFrame2/MeshFrame/MeshFrame/Memory/Array.h
//   CuMatrix::buildOrthonormalBasis, vec2CrossProduct (vendored)                     3rdParty/CuMatrix/CuMatrix/Geometry/Geometry.h, MatrixOps/CuMatrix.h
// oracle/Makefile extracts the line ranges into the build directory (deleted after linking); the two vendored headers are
// included from where they lie. TetMeshFEM itself needs Eigen + MeshFrame2's mesh classes, so this file supplies the stand-in
// <Eigen/Dense> (ref_shim/stub/Eigen/Dense) and a stand-in TetMeshFEM with the members those lines read, backed by flat arrays:
// vertex positions, tet vertex ids, neighbour tets across each face, xor of each tet's vertex ids (the oracle's set-up,
// M/TetMesh/TetMeshFEM.cpp:136-221 restated). Everything lives in namespace GAIA_TT.
#include <cstddef>
#include <cstdint>
#include <cmath>
#include <cassert>
#include <atomic>
#include <iostream>
#include <string>
#include <vector>
#include <Eigen/Dense>
#include <MeshFrame/Memory/Array.h>
#include <CuMatrix/Geometry/Geometry.h>

namespace GAIA_TT {
typedef float FloatingType;
typedef int32_t IdType;
typedef Eigen::Matrix<FloatingType, 2, 1> Vec2;
typedef Eigen::Matrix<FloatingType, 3, 1> Vec3;

// Simulator/Modules/TetMesh/TetMeshFEM.h:36-53
enum class TraverseStopReason { querySuccess = 0, overflow = 1, passedMaximumDis = 2, reachedBoundary = 3, emptyStack = 4 };
struct TraverseStatistics { int numTetsTraversed = 0; TraverseStopReason stopReason; };

// 4 x N / 1 x N integer tables of the topology: what tetVIds()(i, t), .col(t).data() and tetsXorSums()(t) need
struct IntMat4 {
    int32_t* p = nullptr;
    struct Col { int32_t* q; int32_t* data() { return q; } };
    Col col(int j) { return Col{p + 4 * (size_t)j}; }
    int32_t operator()(int i, int j) const { return p[4 * (size_t)j + i]; }
};
struct IntVec { int32_t* p = nullptr; int32_t operator()(int j) const { return p[j]; } };
struct RefTetTopology { IntMat4 tetVIds, tetsNeighborTets; IntVec tetsXorSums; };

struct TetMeshFEM {
    const float* pos = nullptr;
    RefTetTopology topo;
    RefTetTopology* pTopology = &topo;
    Vec3 vertex(int i) const { return Vec3(pos[3 * (size_t)i], pos[3 * (size_t)i + 1], pos[3 * (size_t)i + 2]); }
    IntMat4& tetVIds() { return topo.tetVIds; }
    IntMat4& tetsNeighborTets() { return topo.tetsNeighborTets; }
    IntVec& tetsXorSums() { return topo.tetsXorSums; }
    static const int32_t passThroughCheckSteps;
    static const int32_t tet4Faces[4][3];
    static const int32_t posibleExit3Faces[3][3];
    // defined by the extracted lines
    bool tetrahedralTraverseTo(const Vec3& rayOrigin, const Vec3& rayDir, const FloatingType maxTraversalDis, int32_t startTetId, int32_t startFaceId,
                               int32_t targetTetId, FloatingType rayTriIntersectionEpsilon, TraverseStatistics& statistics);
    int exitFaceSelection(Eigen::Matrix<FloatingType, 2, 4>& ptsProj2D, Eigen::Vector4i& possibleExitFace, FloatingType rayTriIntersectionEpsilon);
    void projectTo2DCoordinates(Eigen::Matrix<FloatingType, 2, 4>& ptsProj2D, int32_t tetId, int32_t incomingFaceId,
                                const Eigen::Matrix<FloatingType, 2, 3>& axesT, const Vec3& origin);
    void copyRepermutedVerts(Eigen::Matrix<FloatingType, 3, 4>& ptsPermuted3D, int32_t tetId, int32_t incomingFaceId);
    template <typename Derived2x1>
    void originBarycentric2DTriangle(const Eigen::MatrixBase<Derived2x1>& a, const Eigen::MatrixBase<Derived2x1>& b, const Eigen::MatrixBase<Derived2x1>& c,
                                     Vec3& barys);
};

#include "ref_tetmesh_h_420_489.inc"
#include "ref_tetmesh_h_525_561.inc"
#include "ref_tetmesh_h_625_633.inc"
#include "ref_tetmesh_cpp_14_25.inc"
#include "ref_tetmesh_cpp_503_742.inc"
}  // namespace GAIA_TT

extern "C" int ref_tet_traverse(int nV, const float* pos, int nT, int32_t* tets, int32_t* nbrTets, int32_t* xorSums, const float* origin3, const float* dir3,
                                float maxDis, int startTet, int startFace, int targetTet, float eps, int32_t* numTraversed, int32_t* stopReason) {
    GAIA_TT::TetMeshFEM m;
    m.pos = pos; m.topo.tetVIds.p = tets; m.topo.tetsNeighborTets.p = nbrTets; m.topo.tetsXorSums.p = xorSums;
    GAIA_TT::TraverseStatistics st;
    const bool ok = m.tetrahedralTraverseTo(GAIA_TT::Vec3(origin3[0], origin3[1], origin3[2]), GAIA_TT::Vec3(dir3[0], dir3[1], dir3[2]), maxDis, startTet,
                                            startFace, targetTet, eps, st);
    if (numTraversed) *numTraversed = st.numTetsTraversed;
    if (stopReason) *stopReason = (int)st.stopReason;
    return ok ? 1 : 0;
}
