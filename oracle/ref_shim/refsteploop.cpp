// ORACLE — TEST INFRASTRUCTURE ONLY.
// The SCHEDULE of the reference's tet step, from its own code: VBDPhysics::runStep_hybridCollisionHandling, Simulator/Modules/VBD/
// VBDPhysics.cpp:1634-1707, compiled from where it lies under /root/reference (oracle/Makefile extracts the lines into the build
// directory, deleted after linking) against a stand-in VBDPhysics whose member functions only RECORD that they were called:
// which operations run, in which order, for which (mesh, vertex), with which friction flag, in every substep and iteration.
// The detection passes are the reference's own as well - VBDPhysics::dcd, ::intermediateDCD, ::ccd, ::intermediateCCD,
// ::intermediateCollisionDetection, VBDPhysics.cpp:1958-2172 - run against stand-in detectors that record every query (which
// surface vertices are queried by which pass, when the penetrated mask is cleared and read, when the trees are rebuilt); the
// OUTCOME of a DCD query (intersections found or not) is data taken from the caller's recording, like the active lists.
// Events are 4 ints {code, a, b, c} with the codes of oracle/vbd_oracle.h (Scene::trace): 1 dcd, 2 initial step (mesh),
// 3 deformers, 4 ccd, 5 prepare, 6 update collision info (mesh, vertex), 7 vertex solve (mesh, vertex, friction),
// 8 intermediate detection, 9 velocity update, 11 active-list entry after a 5 (group, mesh, vertex), 12 DCD trees updated (rebuild
// tets, rebuild surface), 13 DCD query (mesh, vertex, intersections), 14 closest-point query (mesh, vertex), 15 CCD tree updated
// (rebuild), 16 CCD query (mesh, vertex), 17 result cleared (mesh, surface vertex), 18 penetrated mask cleared (mesh).
// The active collision lists are data, not schedule: whenever the stand-in "prepares" (5, 8) it takes the lists the caller recorded
// at that point (the 11-events of the oracle's own trace) and repeats them, so that the two traces can be compared whole.
// Everything lives in namespace GAIA_SL.
#include <cstdint>
#include <functional>
#include <iostream>
#include <memory>
#include <vector>

#define GAIA GAIA_SL
#define DEBUG_LVL_DEBUG 2
#define DEBUG_LVL_DEBUG_VEBOSE 3
#define TICK(x)
#define TOCK_STRUCT(a, b)
#define GRAVITY_AXIS 1      // Modules/Types/Types.h:9

namespace GAIA {
typedef float FloatingType;
typedef int32_t IdType;
template <class F> static void debugOperation(int, F&&) {}
template <class F> static void cpu_parallel_for(int b, int e, F&& fn) { for (int i = b; i < e; i++) fn(i); }

struct Trace {
    std::vector<int32_t> ev;
    void add(int code, int a = 0, int b = 0, int c = 0) { ev.push_back(code); ev.push_back(a); ev.push_back(b); ev.push_back(c); }
};
struct RefMask { std::vector<uint8_t> v; void setZero(); uint8_t& operator()(size_t i) { return v[i]; } int meshId = 0; Trace* tr = nullptr; };
struct RefIntVec { const int32_t* p = nullptr; size_t n = 0; size_t size() const { return n; } int32_t operator()(size_t i) const { return p[i]; } };
struct RefHeight { float y; float operator[](int) const { return y; } };
struct RefMesh {
    int id = 0; Trace* tr = nullptr;
    bool activeForMaterialSolve = true, activeForCollision = true;
    std::vector<uint8_t> fixedMask;
    RefMask penetratedMask;
    RefIntVec svids;
    void evaluateExternalForce() {}
    void applyInitialStep() { tr->add(2, id); }
    const RefIntVec& surfaceVIds() const { return svids; }
    RefHeight vertex(int) const { return RefHeight{-1e30f}; }      // always below collisionOffHeight (its default is 1e7)
};
typedef RefMesh VBDTetMeshNeoHookean;
typedef RefMesh VBDBaseTetMesh;
typedef RefMesh TetMeshFEM;
struct VBDCollisionDetectionResult {
    int mesh = 0, sv = 0, v = 0, n = 0; Trace* tr = nullptr;
    int numIntersections() const { return n; }
    void clear() { tr->add(17, mesh, sv); n = 0; }
};
struct ClosestPointQueryResult {};
struct VBDPhysics;
struct RefDCD {
    VBDPhysics* ph = nullptr;
    void vertexCollisionDetection(int32_t v, int iMesh, VBDCollisionDetectionResult* r);
    void closestPointQuery(VBDCollisionDetectionResult* r, ClosestPointQueryResult*);
};
struct RefCCD { VBDPhysics* ph = nullptr; void vertexContinuousCollisionDetection(int32_t v, int iMesh, VBDCollisionDetectionResult* r); };
struct RefPhysicsParams {
    int numSubsteps = 1, iterations = 1, frictionStartIter = 0, intermediateCollisionIterations = 0; FloatingType dt = 0.f;
    int dcdTetMeshSceneBVHRebuildSteps = 16, dcdSurfaceSceneBVHRebuildSteps = 16, ccdBVHRebuildSteps = 16;
    FloatingType collisionOffHeight = 1e7f; bool useGPU = false;
};
struct RefCollisionParams { bool allowDCD = true, allowCCD = true; };
struct RefActiveList {
    std::vector<size_t> numActiveCollisionsEachParallelGroup;
    std::vector<std::vector<IdType>> activeCollisionsEachParallelGroup;
};
struct RefColResult { IdType mesh, v; };

struct VBDPhysics {
    Trace tr;
    RefPhysicsParams p;
    int substep = 0, iIter = 0;
    double curTime = 0.0;
    std::vector<std::shared_ptr<RefMesh>> tMeshes;
    std::vector<std::vector<IdType>> vertexParallelGroups;
    RefActiveList activeColllisionList;
    RefColResult cur;
    // the caller's recording of the lists: 11-events in the order they were taken
    const int32_t* rec = nullptr; int64_t nRec = 0, at = 0;
    RefPhysicsParams& physicsParams() { return p; }
    int timeStatistics() { return 0; }
    size_t numTetMeshes() const { return tMeshes.size(); }
    void outputPosVel() {} void clearForces() {} void outputForces() {}
    void loadLists(int code) {       // skip to the next `code` event of the recording and take the 11-events that follow it
        for (auto& l : activeColllisionList.activeCollisionsEachParallelGroup) l.clear();
        while (at < nRec && rec[4 * at] != code) at++;
        at++;
        while (at < nRec && rec[4 * at] == 11) {
            activeColllisionList.activeCollisionsEachParallelGroup[rec[4 * at + 1]].push_back(rec[4 * at + 2]);
            activeColllisionList.activeCollisionsEachParallelGroup[rec[4 * at + 1]].push_back(rec[4 * at + 3]);
            tr.add(11, rec[4 * at + 1], rec[4 * at + 2], rec[4 * at + 3]);
            at++;
        }
        for (size_t g = 0; g < vertexParallelGroups.size(); g++)
            activeColllisionList.numActiveCollisionsEachParallelGroup[g] = activeColllisionList.activeCollisionsEachParallelGroup[g].size() / 2;
    }
    RefCollisionParams cp;
    RefCollisionParams& collisionParams() { return cp; }
    int frameId = 0;
    std::vector<std::vector<VBDCollisionDetectionResult>> collisionResultsAll;
    RefDCD dcdStub; RefDCD* pDCD = &dcdStub;
    RefCCD ccdStub; RefCCD* pCCD = &ccdStub;
    int64_t atQuery = 0;      // cursor over the recording for the outcomes of the DCD queries
    int dcdOutcome(int mesh, int v) {      // the next recorded DCD query must be this one; its outcome is data
        while (atQuery < nRec && rec[4 * atQuery] != 13) atQuery++;
        if (atQuery >= nRec || rec[4 * atQuery + 1] != mesh || rec[4 * atQuery + 2] != v) return 0;
        return rec[4 * atQuery++ + 3];
    }
    void updateDCDBVH(bool a, bool b) { tr.add(12, a, b); }
    void updateCCDBVH(bool a) { tr.add(15, a); }
    void syncAllToCPUVertPosOnly(bool) {} void prepareCollisionDataGPU() {} void syncAllToGPU(bool) {}
    // the step loop calls these names; the reference's own bodies are compiled under the names with _ref (see the #defines below)
    void dcd_ref(); void intermediateDCD(); void ccd_ref(); void intermediateCCD(); void intermediateCollisionDetection_ref();
    void dcd() { tr.add(1); dcd_ref(); }
    void applyDeformers() { tr.add(3); }
    void ccd() { tr.add(4); ccd_ref(); }
    void prepareCollisionDataCPU() { tr.add(5); loadLists(5); }
    void intermediateCollisionDetection() { tr.add(8); intermediateCollisionDetection_ref(); }
    void updateVelocities() { tr.add(9); }
    RefColResult& getCollisionDetectionResultFromTetMeshId(IdType iMesh, int vId) { cur.mesh = iMesh; cur.v = vId; return cur; }
    void updateCollisionInfo(RefColResult& r) { tr.add(6, r.mesh, r.v); }
    void VBDStepWithCollision(VBDTetMeshNeoHookean*, IdType iMesh, int vId, bool apply_friction) { tr.add(7, iMesh, vId, apply_friction ? 1 : 0); }
    void runStep_hybridCollisionHandling();
};
}  // namespace GAIA

void GAIA::RefMask::setZero() { tr->add(18, meshId); for (auto& x : v) x = 0; }
void GAIA::RefDCD::vertexCollisionDetection(int32_t v, int iMesh, VBDCollisionDetectionResult* r) { r->n = ph->dcdOutcome(iMesh, v); ph->tr.add(13, iMesh, v, r->n); }
void GAIA::RefDCD::closestPointQuery(VBDCollisionDetectionResult* r, ClosestPointQueryResult*) { ph->tr.add(14, r->mesh, r->v); }
void GAIA::RefCCD::vertexContinuousCollisionDetection(int32_t v, int iMesh, VBDCollisionDetectionResult*) { ph->tr.add(16, iMesh, v); }

using namespace GAIA;
#include "ref_vbdphysics_cpp_1634_1707.inc"
// the detection passes: dcd / ccd / intermediateCollisionDetection are defined under *_ref (the stand-in's members of those names
// record the call and then run these)
#define dcd dcd_ref
#define ccd ccd_ref
#define intermediateCollisionDetection intermediateCollisionDetection_ref
#include "ref_vbdphysics_cpp_1958_2172.inc"
#undef dcd
#undef ccd
#undef intermediateCollisionDetection
#undef GAIA

// params10: numSubsteps, iterations, frictionStartIter, intermediateCollisionIterations, frameId, allowDCD, allowCCD, dcdTetMeshSceneBVHRebuildSteps,
// dcdSurfaceSceneBVHRebuildSteps, ccdBVHRebuildSteps; svOff [nMesh + 1] into surfaceVIds; penetrated [total vertices]: the mask before the
// frame; meshes: active flag and fixed mask each
// (fixedOff [nMesh + 1] into fixedMask); groups: groupOff [nGroups + 1] into groupPairs (mesh, vertex pairs); rec: the recording
// (events, 4 ints each). Returns the number of ints written to out (or needed, if cap is too small).
extern "C" int64_t ref_step_trace(const int32_t* params4, int nMesh, const int32_t* active, const int32_t* fixedOff, const uint8_t* fixedMask, int nGroups,
                                  const int32_t* groupOff, const int32_t* groupPairs, const int32_t* rec, int64_t nRecEvents, int32_t* out, int64_t cap,
                                  const int32_t* svOff, const int32_t* surfaceVIds, uint8_t* penetrated) {
    GAIA_SL::VBDPhysics ph;
    ph.p.numSubsteps = params4[0]; ph.p.iterations = params4[1]; ph.p.frictionStartIter = params4[2]; ph.p.intermediateCollisionIterations = params4[3];
    ph.frameId = params4[4]; ph.cp.allowDCD = params4[5] != 0; ph.cp.allowCCD = params4[6] != 0;
    ph.p.dcdTetMeshSceneBVHRebuildSteps = params4[7]; ph.p.dcdSurfaceSceneBVHRebuildSteps = params4[8]; ph.p.ccdBVHRebuildSteps = params4[9];
    ph.dcdStub.ph = &ph; ph.ccdStub.ph = &ph;
    ph.collisionResultsAll.resize(nMesh);
    for (int i = 0; i < nMesh; i++) {
        auto m = std::make_shared<GAIA_SL::RefMesh>();
        m->id = i; m->tr = &ph.tr; m->activeForMaterialSolve = active[i] != 0; m->activeForCollision = active[i] != 0;
        m->fixedMask.assign(fixedMask + fixedOff[i], fixedMask + fixedOff[i + 1]);
        m->penetratedMask.v.assign(penetrated + fixedOff[i], penetrated + fixedOff[i + 1]); m->penetratedMask.meshId = i; m->penetratedMask.tr = &ph.tr;
        m->svids.p = surfaceVIds + svOff[i]; m->svids.n = (size_t)(svOff[i + 1] - svOff[i]);
        ph.collisionResultsAll[i].resize(m->svids.n);
        for (size_t sv = 0; sv < m->svids.n; sv++) { auto& r = ph.collisionResultsAll[i][sv]; r.mesh = i; r.sv = (int)sv; r.v = m->svids.p[sv]; r.tr = &ph.tr; }
        ph.tMeshes.push_back(m);
    }
    ph.vertexParallelGroups.resize(nGroups);
    for (int g = 0; g < nGroups; g++) ph.vertexParallelGroups[g].assign(groupPairs + groupOff[g], groupPairs + groupOff[g + 1]);
    ph.activeColllisionList.numActiveCollisionsEachParallelGroup.assign(nGroups, 0);
    ph.activeColllisionList.activeCollisionsEachParallelGroup.resize(nGroups);
    ph.rec = rec; ph.nRec = nRecEvents;
    ph.runStep_hybridCollisionHandling();
    for (int i = 0; i < nMesh; i++) for (size_t k = 0; k < ph.tMeshes[i]->penetratedMask.v.size(); k++) penetrated[fixedOff[i] + k] = ph.tMeshes[i]->penetratedMask.v[k];   // carried to the next frame
    const int64_t n = (int64_t)ph.tr.ev.size();
    if (out && cap >= n) for (int64_t i = 0; i < n; i++) out[i] = ph.tr.ev[i];
    return n;
}
