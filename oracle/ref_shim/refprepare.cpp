// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own construction of the active collision lists and the face-side relations (SURVEY §8 row
// a15), compiled from where it lies under /root/reference (Simulator/Modules/VBD):
//   CollisionRelation, CollisionRelationList, ActiveCollisionList                       ActiveCollisionList.h:9-53
//   ActiveCollisionList::initializeParallelGroup, ::clear, ::addToActiveCollisionList   ActiveCollisionList.cpp:6-110
//   VBDPhysics::prepareCollisionDataCPU                                                 VBDPhysics.cpp:1709-1810
// with MeshFrame2's CPArray (3rdParty/MeshFrame2/MeshFrame/MeshFrame/Memory/Array.h, vendored by the reference). oracle/Makefile
// extracts the line ranges into the build directory (deleted after linking). Stand-ins with the members those lines read: meshes
// over flat integer arrays (group of every vertex, surface vertex ids, surface faces), detection results (query mesh / vertex,
// from CCD, and per intersection: shortest path found, intersected mesh, closest surface face). Everything in namespace GAIA_PC.
#include <atomic>
#include <cassert>
#include <cstdint>
#include <iostream>
#include <memory>
#include <string>
#include <utility>
#include <vector>
#include <MeshFrame/Memory/Array.h>

#define GAIA GAIA_PC
#define COLLISION_RELATION_PREALLOCATE 4      // ActiveCollisionList.h:6

namespace GAIA {
typedef float FloatingType;
typedef const float CFloatingType;
typedef int32_t IdType;
template <class F> static void cpu_parallel_for(int b, int e, F&& fn) { for (int i = b; i < e; i++) fn(i); }

struct VecDynamicI {      // Eigen::VectorXi: what the lines use of it
    std::vector<int> d;
    size_t size() const { return d.size(); }
    void resize(size_t n) { d.resize(n); }
    int& operator()(size_t i) { return d[i]; }
};
struct CollidingPointInfo { bool shortestPathFound = false; int intersectedMeshId = -1, closestSurfaceFaceId = -1; };
struct CollisionDetectionResult {
    int idTMQuery = -1, idVQuery = -1;
    bool fromCCD = false;
    std::vector<CollidingPointInfo> collidingPts;
    size_t numIntersections() const { return collidingPts.size(); }
};
typedef CollisionDetectionResult VBDCollisionDetectionResult;
struct RefIntMat3 { const int32_t* p = nullptr; int32_t operator()(size_t i, size_t f) const { return p[3 * f + i]; } };
struct RefIntVec { const int32_t* p = nullptr; size_t n = 0; size_t size() const { return n; } int32_t operator()(size_t i) const { return p[i]; } };
struct RefMask { std::vector<uint8_t> v; void setZero() { for (auto& x : v) x = 0; } uint8_t& operator()(size_t i) { return v[i]; } };
struct VBDBaseTetMesh {
    int nV = 0;
    std::vector<int> vertexParallelGroups;
    RefIntMat3 faces; RefIntVec svids; RefMask activeCollisionMask;
    size_t numVertices() const { return (size_t)nV; }
    const RefIntMat3& surfaceFacesTetMeshVIds() const { return faces; }
    const RefIntVec& surfaceVIds() const { return svids; }
};

#include "ref_activelist_h_9_53.inc"

struct VBDPhysics {
    std::vector<std::shared_ptr<VBDBaseTetMesh>> tMeshes;
    std::vector<std::vector<VBDCollisionDetectionResult>> collisionResultsAll;
    ActiveCollisionList activeColllisionList;
    CollisionRelationList& getCollisionRelationList(int meshId, int vId) { return activeColllisionList.vertexCollisionRelations[meshId][vId]; }   // VBDPhysics.h
    void prepareCollisionDataCPU();
};
}  // namespace GAIA

using namespace GAIA;
#include "ref_activelist_cpp_6_110.inc"
#include "ref_vbdphysics_cpp_1709_1810.inc"
#undef GAIA

// Meshes concatenated: vOff / svOff / fOff [nMesh + 1] into vertexGroup (group of every vertex), surfaceVIds, surfaceFaces (3 per
// face). Detection results per surface vertex (svOff indexing): has (0 / 1 intersection), fromCCD, shortestPathFound,
// intersectedMesh, closestFace. groupSizes [nGroups] = number of (mesh, vertex) pairs of each parallel group (the lists are
// pre-allocated from it: 2 ints per pair times preAllocRatio, as initializeParallelGroup does).
// Outputs: activeOff [nGroups + 1] + activePairs (mesh, vertex per entry, in the reference's order); relOff [total vertices + 1] +
// relRows (meshId, surfaceVertexId, collisionId, collisionType, collisionVertexOrder per relation, in order); mask [total vertices].
// Returns the number of relation rows (-1 if a capacity is too small).
extern "C" int64_t ref_prepare_collision_data(int nMesh, const int32_t* vOff, const int32_t* svOff, const int32_t* fOff, const int32_t* vertexGroup,
                                              const int32_t* surfaceVIds, const int32_t* surfaceFaces, const uint8_t* has, const uint8_t* fromCCD,
                                              const uint8_t* pathFound, const int32_t* hitMesh, const int32_t* hitFace, int nGroups, const int32_t* groupSizes,
                                              float preAllocRatio, int32_t* activeOff, int32_t* activePairs, int64_t activeCap, int32_t* relOff, int32_t* relRows,
                                              int64_t relCap, uint8_t* mask) {
    GAIA_PC::VBDPhysics ph;
    std::vector<std::vector<GAIA_PC::IdType>> groups(nGroups);
    for (int g = 0; g < nGroups; g++) groups[g].assign(2 * (size_t)groupSizes[g], 0);
    ph.collisionResultsAll.resize(nMesh);
    for (int m = 0; m < nMesh; m++) {
        auto p = std::make_shared<GAIA_PC::VBDBaseTetMesh>();
        p->nV = vOff[m + 1] - vOff[m];
        p->vertexParallelGroups.assign(vertexGroup + vOff[m], vertexGroup + vOff[m + 1]);
        p->faces.p = surfaceFaces + 3 * (size_t)fOff[m];
        p->svids.p = surfaceVIds + svOff[m]; p->svids.n = (size_t)(svOff[m + 1] - svOff[m]);
        p->activeCollisionMask.v.assign(p->nV, 1);      // stale values: prepareCollisionDataCPU has to clear them
        ph.tMeshes.push_back(p);
        ph.collisionResultsAll[m].resize(p->svids.n);
        for (size_t sv = 0; sv < p->svids.n; sv++) {
            GAIA_PC::VBDCollisionDetectionResult& r = ph.collisionResultsAll[m][sv];
            const size_t q = (size_t)svOff[m] + sv;
            r.idTMQuery = m; r.idVQuery = p->svids.p[sv]; r.fromCCD = fromCCD[q] != 0;
            if (has[q]) { r.collidingPts.resize(1); r.collidingPts[0].shortestPathFound = pathFound[q] != 0; r.collidingPts[0].intersectedMeshId = hitMesh[q]; r.collidingPts[0].closestSurfaceFaceId = hitFace[q]; }
        }
    }
    ph.activeColllisionList.initialize(ph.tMeshes, groups, preAllocRatio);
    ph.prepareCollisionDataCPU();
    int64_t at = 0;
    for (int g = 0; g < nGroups; g++) {
        activeOff[g] = (int32_t)at;
        const size_t n = ph.activeColllisionList.numActiveCollisionsEachParallelGroup[g];
        for (size_t i = 0; i < n; i++) {
            if (at + 1 > activeCap) return -1;
            activePairs[2 * at] = ph.activeColllisionList.activeCollisionsEachParallelGroup[g](2 * i);
            activePairs[2 * at + 1] = ph.activeColllisionList.activeCollisionsEachParallelGroup[g](2 * i + 1);
            at++;
        }
    }
    activeOff[nGroups] = (int32_t)at;
    int64_t nr = 0;
    for (int m = 0; m < nMesh; m++)
        for (int v = 0; v < vOff[m + 1] - vOff[m]; v++) {
            relOff[vOff[m] + v] = (int32_t)nr;
            auto& rl = ph.activeColllisionList.vertexCollisionRelations[m][v];
            for (size_t i = 0; i < rl.size(); i++) {
                if (nr + 1 > relCap) return -1;
                relRows[5 * nr] = rl[i].meshId; relRows[5 * nr + 1] = rl[i].surfaceVertexId; relRows[5 * nr + 2] = rl[i].collisionId;
                relRows[5 * nr + 3] = rl[i].collisionType; relRows[5 * nr + 4] = rl[i].collisionVertexOrder;
                nr++;
            }
            mask[vOff[m] + v] = ph.tMeshes[m]->activeCollisionMask.v[v];
        }
    relOff[vOff[nMesh]] = (int32_t)nr;
    return nr;
}
