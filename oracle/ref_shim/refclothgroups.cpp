// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own merge of the cloths' colour classes into parallel groups, compiled from where it lies
// under /root/reference: VBDClothSimulationFramework::initializeParallelGroups and ::findSmallestParallelGroup (Simulator/Modules/
// VBDCloth/VBDClothPhysics.cpp:72-138; unlike the tet framework the group is NOT taken off the mesh's list, so two classes of one
// mesh can share a group). oracle/Makefile extracts the lines into the build directory (deleted after linking). Namespace GAIA_CG.
#include <cassert>
#include <cstdint>
#include <memory>
#include <vector>

#define GAIA GAIA_CG
namespace GAIA {
typedef int32_t IdType;
struct RefMesh {
    std::vector<std::vector<IdType>> cats; std::vector<IdType> globalColors;
    std::vector<std::vector<IdType>>& verticesColoringCategories() { return cats; }
};
struct VBDClothSimulationFramework {
    std::vector<std::shared_ptr<RefMesh>> baseTriMeshesForSimulation;
    std::vector<std::vector<IdType>> vertexParallelGroups;
    void initializeParallelGroups();
    IdType findSmallestParallelGroup(std::vector<IdType>& availableColors, const std::vector<std::vector<IdType>>& vertexParallelGroups, bool removeSmallestAvailableColor);
};
}  // namespace GAIA
using namespace GAIA;
#include "ref_clothphysics_cpp_72_138.inc"
#undef GAIA

// as ref_merge_groups; also globalColor [total vertices] (vOff from nVerts)
extern "C" int ref_cloth_merge_groups(int nMesh, const int32_t* nVerts, const int32_t* nColors, const int32_t* classOff, const int32_t* classVerts, int32_t* groupOff,
                                      int32_t* groupPairs, int32_t* globalColor) {
    GAIA_CG::VBDClothSimulationFramework fw;
    int k = 0;
    for (int m = 0; m < nMesh; m++) {
        auto p = std::make_shared<GAIA_CG::RefMesh>();
        p->globalColors.assign((size_t)nVerts[m], -1);
        for (int c = 0; c < nColors[m]; c++, k++) p->cats.emplace_back(classVerts + classOff[k], classVerts + classOff[k + 1]);
        fw.baseTriMeshesForSimulation.push_back(p);
    }
    fw.initializeParallelGroups();
    int at = 0;
    for (size_t g = 0; g < fw.vertexParallelGroups.size(); g++) { groupOff[g] = at; for (int x : fw.vertexParallelGroups[g]) groupPairs[at++] = x; }
    groupOff[fw.vertexParallelGroups.size()] = at;
    int v0 = 0;
    for (int m = 0; m < nMesh; m++) { for (int v = 0; v < nVerts[m]; v++) globalColor[v0 + v] = fw.baseTriMeshesForSimulation[m]->globalColors[v]; v0 += nVerts[m]; }
    return (int)fw.vertexParallelGroups.size();
}
