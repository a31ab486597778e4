// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own feasible-region test of the discrete collision detector, compiled from where it lies
// under /root/reference (Simulator/Modules/CollisionDetector):
//   DiscreteCollisionDetector::checkFeasibleRegion                       DiscreteCollisionDetector.cpp:664-700
//   checkEdgeFeasibleRegion, checkVertexFeasibleRegion                   DiscreteCollisionDetector.cpp:709-815
//   loadVertexPos, faceNormal                                            DiscreteCollisionDetector.h:102-121
// with the Embree math headers the reference vendors under Modules/common. oracle/Makefile extracts those line ranges into
// the build directory (deleted after linking). The detector class binds Embree scenes and TetMeshFEM needs Eigen + MeshFrame2,
// so this file supplies stand-ins with the members those lines read, backed by flat arrays (positions padded by one float:
// Vec3fa::loadu reads four): surface faces, neighbour face across each edge, tet vertex -> surface vertex, and the
// neighbouring surface vertices of each surface vertex (the oracle's set-up, M/TetMesh/TetMeshFEM.cpp:136-221 restated).
#include "common/math/vec2.h"
#include "common/math/vec3.h"
#include "common/math/vec3fa.h"
#include <cassert>
#include <cstdint>
#include <cstdio>
#include <vector>

namespace GAIA {
using embree::Vec3fa;
// Simulator/Modules/CollisionDetector/CollisionDetertionParameters.h:141-151
enum class ClosestPointOnTriangleType { AtA, AtB, AtC, AtAB, AtBC, AtAC, AtInterior, NotFound };

struct FloatCols {
    const float* p = nullptr;
    struct Col { const float* q; const float* data() { return q; } };
    Col col(int i) { return Col{p + 3 * (size_t)i}; }
};
struct IntMat3 {
    const int32_t* p = nullptr;
    struct Col { const int32_t* q; const int32_t* data() { return q; } };
    Col col(int j) { return Col{p + 3 * (size_t)j}; }
    int32_t operator()(int i, int j) const { return p[3 * (size_t)j + i]; }
};
struct IntVec { const int32_t* p = nullptr; int32_t operator()(int j) const { return p[j]; } };
struct TetMeshFEM {
    FloatCols mVertPos;
    IntMat3 faces, faceNbr;
    IntVec tetToSurf;
    std::vector<std::vector<int32_t>> nbrVerts;
    IntMat3& surfaceFacesTetMeshVIds() { return faces; }
    IntMat3& surfaceFaces3NeighborFaces() { return faceNbr; }
    IntVec& tetVertIndicesToSurfaceVertIndices() { return tetToSurf; }
    std::vector<std::vector<int32_t>>& surfaceVertexNeighborSurfaceVertices() { return nbrVerts; }
};
// DiscreteCollisionDetector.h:92-100
embree::Vec3fa loadVertexPos(TetMeshFEM* pTM, int32_t vId);
embree::Vec3fa faceNormal(TetMeshFEM* pTM, int32_t faceId);
bool checkEdgeFeasibleRegion(embree::Vec3fa& p, TetMeshFEM* pTM, int32_t faceId, int32_t edgeId, int32_t edgeVId1, int32_t edgeVId2, float feasibleRegionEpsilon);
bool checkVertexFeasibleRegion(embree::Vec3fa& p, TetMeshFEM* pTM, int32_t vId, float feasibleRegionEpsilon);
struct DiscreteCollisionDetector {
    bool checkFeasibleRegion(embree::Vec3fa& p, TetMeshFEM* pTM, int32_t faceId, ClosestPointOnTriangleType pointType, float feasibleRegionEpsilon);
};
}  // namespace GAIA

using namespace GAIA;
using embree::Vec3fa;
#define ABSOLUTE_RELAXIATION 0.f      // DiscreteCollisionDetector.cpp:8
#include "ref_dcd_h_102_121.inc"
#include "ref_dcd_cpp_664_700.inc"
#include "ref_dcd_cpp_709_815.inc"

extern "C" int ref_check_feasible_region(int nV, const float* pos, int nF, const int32_t* faces, const int32_t* face3Nbr, const int32_t* tetVertToSurf,
                                         int nSV, const int32_t* nbrOff, const int32_t* nbrVerts, const float* p3, int faceId, int type, float eps) {
    std::vector<float> padded(3 * (size_t)nV + 1, 0.f);
    for (size_t i = 0; i < 3 * (size_t)nV; i++) padded[i] = pos[i];
    TetMeshFEM m;
    m.mVertPos.p = padded.data(); m.faces.p = faces; m.faceNbr.p = face3Nbr; m.tetToSurf.p = tetVertToSurf;
    m.nbrVerts.resize(nSV);
    for (int sv = 0; sv < nSV; sv++) m.nbrVerts[sv].assign(nbrVerts + nbrOff[sv], nbrVerts + nbrOff[sv + 1]);
    DiscreteCollisionDetector d;
    embree::Vec3fa p(p3[0], p3[1], p3[2]);
    return d.checkFeasibleRegion(p, &m, faceId, (ClosestPointOnTriangleType)type, eps) ? 1 : 0;
}
