// ORACLE — TEST INFRASTRUCTURE ONLY.
// The SCHEDULE of the reference's tet step, from its own code: VBDPhysics::runStep_hybridCollisionHandling, Simulator/Modules/VBD/
// VBDPhysics.cpp:1634-1707, compiled from where it lies under /root/reference (oracle/Makefile extracts the lines into the build
// directory, deleted after linking) against a stand-in VBDPhysics whose member functions only RECORD that they were called:
// which operations run, in which order, for which (mesh, vertex), with which friction flag, in every substep and iteration.
// Events are 4 ints {code, a, b, c} with the codes of oracle/vbd_oracle.h (Scene::trace): 1 dcd, 2 initial step (mesh),
// 3 deformers, 4 ccd, 5 prepare, 6 update collision info (mesh, vertex), 7 vertex solve (mesh, vertex, friction),
// 8 intermediate detection, 9 velocity update, 11 active-list entry after a 5 / 8 (group, mesh, vertex).
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

namespace GAIA {
typedef float FloatingType;
typedef int32_t IdType;
template <class F> static void debugOperation(int, F&&) {}
template <class F> static void cpu_parallel_for(int b, int e, F&& fn) { for (int i = b; i < e; i++) fn(i); }

struct Trace {
    std::vector<int32_t> ev;
    void add(int code, int a = 0, int b = 0, int c = 0) { ev.push_back(code); ev.push_back(a); ev.push_back(b); ev.push_back(c); }
};
struct RefMesh {
    int id = 0; Trace* tr = nullptr;
    bool activeForMaterialSolve = true;
    std::vector<uint8_t> fixedMask;
    void evaluateExternalForce() {}
    void applyInitialStep() { tr->add(2, id); }
};
typedef RefMesh VBDTetMeshNeoHookean;
struct RefPhysicsParams { int numSubsteps = 1, iterations = 1, frictionStartIter = 0, intermediateCollisionIterations = 0; FloatingType dt = 0.f; };
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
    void dcd() { tr.add(1); }
    void applyDeformers() { tr.add(3); }
    void ccd() { tr.add(4); }
    void prepareCollisionDataCPU() { tr.add(5); loadLists(5); }
    void intermediateCollisionDetection() { tr.add(8); loadLists(8); }
    void updateVelocities() { tr.add(9); }
    RefColResult& getCollisionDetectionResultFromTetMeshId(IdType iMesh, int vId) { cur.mesh = iMesh; cur.v = vId; return cur; }
    void updateCollisionInfo(RefColResult& r) { tr.add(6, r.mesh, r.v); }
    void VBDStepWithCollision(VBDTetMeshNeoHookean*, IdType iMesh, int vId, bool apply_friction) { tr.add(7, iMesh, vId, apply_friction ? 1 : 0); }
    void runStep_hybridCollisionHandling();
};
}  // namespace GAIA

using namespace GAIA;
#include "ref_vbdphysics_cpp_1634_1707.inc"
#undef GAIA

// params5: numSubsteps, iterations, frictionStartIter, intermediateCollisionIterations; meshes: active flag and fixed mask each
// (fixedOff [nMesh + 1] into fixedMask); groups: groupOff [nGroups + 1] into groupPairs (mesh, vertex pairs); rec: the recording
// (events, 4 ints each). Returns the number of ints written to out (or needed, if cap is too small).
extern "C" int64_t ref_step_trace(const int32_t* params4, int nMesh, const int32_t* active, const int32_t* fixedOff, const uint8_t* fixedMask, int nGroups,
                                  const int32_t* groupOff, const int32_t* groupPairs, const int32_t* rec, int64_t nRecEvents, int32_t* out, int64_t cap) {
    GAIA_SL::VBDPhysics ph;
    ph.p.numSubsteps = params4[0]; ph.p.iterations = params4[1]; ph.p.frictionStartIter = params4[2]; ph.p.intermediateCollisionIterations = params4[3];
    for (int i = 0; i < nMesh; i++) {
        auto m = std::make_shared<GAIA_SL::RefMesh>();
        m->id = i; m->tr = &ph.tr; m->activeForMaterialSolve = active[i] != 0;
        m->fixedMask.assign(fixedMask + fixedOff[i], fixedMask + fixedOff[i + 1]);
        ph.tMeshes.push_back(m);
    }
    ph.vertexParallelGroups.resize(nGroups);
    for (int g = 0; g < nGroups; g++) ph.vertexParallelGroups[g].assign(groupPairs + groupOff[g], groupPairs + groupOff[g + 1]);
    ph.activeColllisionList.numActiveCollisionsEachParallelGroup.assign(nGroups, 0);
    ph.activeColllisionList.activeCollisionsEachParallelGroup.resize(nGroups);
    ph.rec = rec; ph.nRec = nRecEvents;
    ph.runStep_hybridCollisionHandling();
    const int64_t n = (int64_t)ph.tr.ev.size();
    if (out && cap >= n) for (int64_t i = 0; i < n; i++) out[i] = ph.tr.ev[i];
    return n;
}
