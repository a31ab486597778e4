// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own feasible-region test of the cloth contact detector, compiled from where it lies
// under /root/reference (Simulator/Modules/CollisionDetector/TriMeshCollisionGeometry.cpp):
//   checkFeasibleRegion (TriMeshFEM)                                     :7-42
//   checkEdgeFeasibleRegion, checkVertexFeasibleRegion                   :91-186
// with the Embree math headers the reference vendors under Modules/common. oracle/Makefile extracts those line ranges into
// the build directory (deleted after linking). TriMeshFEM needs Eigen + MeshFrame2, so this file supplies a stand-in with the
// members those lines read, backed by flat arrays (positions padded by one float: Vec3fa::loadu reads four): faces, the edge
// behind each face edge, edge infos and the neighbour vertices of each vertex (the cloth oracle's set-up, restated).
#include "common/math/vec2.h"
#include "common/math/vec3.h"
#include "common/math/vec3fa.h"
#include <cstdint>
#include <vector>

#define CMP_EPSILON2 (1e-10f)         // Modules/Types/Types.h:21

namespace GAIA_CF {
typedef float FloatingType;
typedef const float CFloatingType;
typedef int32_t IdType;
typedef const int32_t CIdType;
enum class ClosestPointOnTriangleType { AtA, AtB, AtC, AtAB, AtBC, AtAC, AtInterior, NotFound };
struct EdgeInfo { int eV1, eV2, eV12Next, eV21Next, fId1, fId2; };
struct IntMat3 {
    const int32_t* p = nullptr;
    struct Col { const int32_t* q; const int32_t* data() const { return q; } };
    Col col(int j) const { return Col{p + 3 * (size_t)j}; }
    int32_t operator()(int i, int j) const { return p[3 * (size_t)j + i]; }
};
struct RefTopology { IntMat3 faces3NeighborEdges; std::vector<EdgeInfo> edgeInfos; };
struct TriMeshFEM {
    const float* pos = nullptr;
    IntMat3 facePos;
    RefTopology topo;
    const RefTopology* pTopology = &topo;
    const int32_t* nbrOff = nullptr; const int32_t* nbrV = nullptr;
    struct VRef { const float* q; const float* data() const { return q; } };
    VRef vertex(size_t i) const { return VRef{pos + 3 * i}; }
    IdType facePosVId(int f, int k) const { return facePos(k, f); }
    size_t numNeiVertices(IdType v) const { return (size_t)(nbrOff[v + 1] - nbrOff[v]); }
    IdType getVertexIthNeiVertex(IdType v, IdType i) const { return nbrV[nbrOff[v] + i]; }
};
// TriMeshCollisionGeometry.h:10-13, 19-27
inline embree::Vec3fa loadVertexPos(const TriMeshFEM* pMesh, int32_t vId) { return embree::Vec3fa::loadu(pMesh->vertex(vId).data()); }
bool checkFeasibleRegion(const embree::Vec3fa& p, const TriMeshFEM* pMesh, int32_t faceId, ClosestPointOnTriangleType pointType, float feasibleRegionEpsilon);
bool checkEdgeFeasibleRegion(const embree::Vec3fa& p, const TriMeshFEM* pMesh, int32_t faceId, int32_t edgeOrderInFace, float feasibleRegionEpsilon);
bool checkVertexFeasibleRegion(const embree::Vec3fa& p, const TriMeshFEM* pMesh, int32_t vId, float feasibleRegionEpsilon);
}  // namespace GAIA_CF

#define GAIA GAIA_CF
using namespace GAIA_CF;
#define ABSOLUTE_RELAXIATION (1e-6f)  // TriMeshCollisionGeometry.cpp:4
#include "ref_trigeom_cpp_7_42.inc"
#include "ref_trigeom_cpp_91_186.inc"
#undef GAIA

extern "C" int ref_cloth_check_feasible_region(int nV, const float* pos, int nF, const int32_t* faces, const int32_t* face3Edges, int nE, const int32_t* edge6,
                                               const int32_t* nbrOff, const int32_t* nbrV, const float* p3, int faceId, int type, float eps) {
    std::vector<float> padded(3 * (size_t)nV + 1, 0.f);
    for (size_t i = 0; i < 3 * (size_t)nV; i++) padded[i] = pos[i];
    GAIA_CF::TriMeshFEM m;
    m.pos = padded.data(); m.facePos.p = faces; m.topo.faces3NeighborEdges.p = face3Edges; m.nbrOff = nbrOff; m.nbrV = nbrV;
    m.topo.edgeInfos.resize(nE);
    for (int e = 0; e < nE; e++) {
        GAIA_CF::EdgeInfo& E = m.topo.edgeInfos[e];
        E.eV1 = edge6[6 * e]; E.eV2 = edge6[6 * e + 1]; E.eV12Next = edge6[6 * e + 2]; E.eV21Next = edge6[6 * e + 3]; E.fId1 = edge6[6 * e + 4]; E.fId2 = edge6[6 * e + 5];
    }
    return GAIA_CF::checkFeasibleRegion(embree::Vec3fa(p3[0], p3[1], p3[2]), &m, faceId, (GAIA_CF::ClosestPointOnTriangleType)type, eps) ? 1 : 0;
}
