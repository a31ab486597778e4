// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own merge of the meshes' colour classes into the global parallel groups,
// compiled from where it lies under /root/reference: findSmallestParallelGroup and the merge loop of VBDPhysics::initialize
// (Simulator/Modules/VBD/VBDPhysics.cpp:99-122, 236-274). oracle/Makefile extracts the lines into the build directory (deleted
// after linking). The colour classes come in already sorted by size (TetMeshFEM.cpp:131-132 sorts them with an unstable std::sort;
// the rule stated for this repo is the stable order). Namespace GAIA_PG.
#include <cassert>
#include <cstdint>
#include <memory>
#include <vector>

namespace GAIA_PG {
typedef int32_t IdType;
#include "ref_vbdphysics_cpp_99_122.inc"
struct VBDBaseTetMesh {
    typedef std::shared_ptr<VBDBaseTetMesh> SharedPtr;
    std::vector<std::vector<IdType>> cats; size_t nV = 0;
    std::vector<std::vector<IdType>>& verticesColoringCategories() { return cats; }
    size_t numVertices() const { return nV; }
};
static void merge(std::vector<VBDBaseTetMesh::SharedPtr>& tMeshes, std::vector<std::vector<IdType>>& vertexParallelGroups, size_t& numAllVertices) {
    std::vector<VBDBaseTetMesh::SharedPtr>& basetetMeshes = tMeshes;
#include "ref_vbdphysics_cpp_236_274.inc"
}
}  // namespace GAIA_PG

// meshes: nColors [nMesh], classOff (offsets of every class of every mesh, in order, + 1) into classVerts; out: groupOff [maxColors + 1],
// groupPairs (mesh, vertex per entry). Returns the number of groups.
extern "C" int ref_merge_groups(int nMesh, const int32_t* nVerts, const int32_t* nColors, const int32_t* classOff, const int32_t* classVerts, int32_t* groupOff,
                                int32_t* groupPairs) {
    std::vector<GAIA_PG::VBDBaseTetMesh::SharedPtr> meshes;
    int k = 0;
    for (int m = 0; m < nMesh; m++) {
        auto p = std::make_shared<GAIA_PG::VBDBaseTetMesh>();
        p->nV = (size_t)nVerts[m];
        for (int c = 0; c < nColors[m]; c++, k++) p->cats.emplace_back(classVerts + classOff[k], classVerts + classOff[k + 1]);
        meshes.push_back(p);
    }
    std::vector<std::vector<GAIA_PG::IdType>> groups; size_t nAll = 0;
    GAIA_PG::merge(meshes, groups, nAll);
    int at = 0;
    for (size_t g = 0; g < groups.size(); g++) { groupOff[g] = at; for (int x : groups[g]) groupPairs[at++] = x; }
    groupOff[groups.size()] = at;
    return (int)groups.size();
}
