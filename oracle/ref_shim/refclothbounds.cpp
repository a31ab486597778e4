// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own computation of the cloth conservative bounds and the active-collision mask: the last
// loop of VBDClothSimulationFramework::applyCollisionDetection, Simulator/Modules/VBDCloth/VBDClothPhysics.cpp:775-819, compiled
// from where it lies under /root/reference (oracle/Makefile extracts the lines into the build directory, deleted after linking).
// It folds what the VF / EE queries left behind - per vertex, per face, per edge: the smallest distance met and whether a
// contact was recorded - over each vertex's incident faces and edges. The queries themselves are Embree callbacks (unbuildable
// here); this file supplies stand-ins with the members the lines read, over flat arrays. Everything lives in namespace GAIA_CB.
#include <algorithm>
#include <cstdint>
#include <memory>
#include <vector>

#define GAIA GAIA_CB

namespace GAIA {
typedef float FloatingType;
typedef const float CFloatingType;
typedef int32_t IdType;

struct RefCountOnly { size_t n = 0; size_t size() const { return n; } };
struct RefQueryResult { FloatingType minDisToPrimitives = 0.f; RefCountOnly contactPts; };
typedef RefQueryResult ClothVFContactQueryResult;
typedef RefQueryResult ClothEEContactQueryResult;

struct RefMesh {
    int nV = 0;
    const int32_t *fOff = nullptr, *fAdj = nullptr, *eOff = nullptr, *eAdj = nullptr;      // vertex -> incident faces / edges (CSR)
    std::vector<float> vertexConvervativeBounds;
    std::vector<uint8_t> mask;
    int positionsAtPrevCollisionDetection = 0;
    int positions() const { return 1; }
    size_t numVertices() const { return (size_t)nV; }
    size_t numNeiFaces(int v) const { return (size_t)(fOff[v + 1] - fOff[v]); }
    IdType getVertexIthNeiFace(int v, size_t i) const { return fAdj[fOff[v] + i]; }
    size_t numNeiEdges(int v) const { return (size_t)(eOff[v + 1] - eOff[v]); }
    IdType getVertexIthNeiEdge(int v, size_t i) const { return eAdj[eOff[v] + i]; }
    uint8_t& activeCollisionMask(int v) { return mask[v]; }
};
struct RefDetector {
    std::vector<float> faceMin;
    std::vector<RefCountOnly> faceInfo;
    FloatingType getFaceMinDis(size_t, IdType f) const { return faceMin[f]; }
    RefCountOnly& getFaceContactInfo(size_t, IdType f) { return faceInfo[f]; }
};
struct RefDetectorParams { FloatingType maxQueryDis = 0.f; };
struct RefPhysicsParams { FloatingType thickness = 0.f, conservativeStepRelaxation = 0.f; };
template <class F> static void cpu_parallel_for(int b, int e, F&& fn) { for (int i = b; i < e; i++) fn(i); }

struct VBDClothSimulationFramework {
    std::vector<std::shared_ptr<RefMesh>> baseTriMeshesForSimulation;
    RefDetector det; RefDetector* pClothContactDetector = &det;
    RefDetectorParams dpar; RefDetectorParams* pClothContactDetectorParameters = &dpar;
    RefPhysicsParams phys;
    std::vector<RefQueryResult> vf, ee;
    RefPhysicsParams& physicsParams() { return phys; }
    RefMesh& getSimulatedMesh(size_t i) { return *baseTriMeshesForSimulation[i]; }
    ClothVFContactQueryResult& getVFContactResult(size_t, int v) { return vf[v]; }
    ClothEEContactQueryResult& getEEContactResult(size_t, int e) { return ee[e]; }
    void boundsLoop() {
#include "ref_clothphysics_cpp_775_819.inc"
    }
};
}  // namespace GAIA
#undef GAIA

// one simulated mesh: vertex -> faces / edges incidence (CSR), what the queries left per vertex / face / edge (smallest distance,
// number of contacts recorded), the mask before; out: conservative bounds [nV], mask after [nV]
extern "C" void ref_cloth_bounds(int nV, int nF, int nE, const int32_t* fOff, const int32_t* fAdj, const int32_t* eOff, const int32_t* eAdj, float maxQueryDis,
                                 float thickness, float relaxation, const float* vfMinDis, const int32_t* vfCount, const float* faceMinDis, const int32_t* faceCount,
                                 const float* eeMinDis, const int32_t* eeCount, const uint8_t* maskBefore, float* bounds, uint8_t* maskAfter) {
    GAIA_CB::VBDClothSimulationFramework fw;
    auto m = std::make_shared<GAIA_CB::RefMesh>();
    m->nV = nV; m->fOff = fOff; m->fAdj = fAdj; m->eOff = eOff; m->eAdj = eAdj;
    m->vertexConvervativeBounds.assign(nV, -1.f); m->mask.assign(maskBefore, maskBefore + nV);
    fw.baseTriMeshesForSimulation.push_back(m);
    fw.dpar.maxQueryDis = maxQueryDis; fw.phys.thickness = thickness; fw.phys.conservativeStepRelaxation = relaxation;
    fw.vf.resize(nV); fw.ee.resize(nE); fw.det.faceMin.assign(faceMinDis, faceMinDis + nF); fw.det.faceInfo.resize(nF);
    for (int v = 0; v < nV; v++) { fw.vf[v].minDisToPrimitives = vfMinDis[v]; fw.vf[v].contactPts.n = (size_t)vfCount[v]; }
    for (int f = 0; f < nF; f++) fw.det.faceInfo[f].n = (size_t)faceCount[f];
    for (int e = 0; e < nE; e++) { fw.ee[e].minDisToPrimitives = eeMinDis[e]; fw.ee[e].contactPts.n = (size_t)eeCount[e]; }
    fw.boundsLoop();
    for (int v = 0; v < nV; v++) { bounds[v] = m->vertexConvervativeBounds[v]; maskAfter[v] = m->mask[v]; }
}
