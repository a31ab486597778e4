// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own key-frame collider, ColliderTrimeshSequence::update and ::set_state (Simulator/Modules/
// SpatialQuery/ColiiderTriMeshSequence.h:59-124), compiled from where they lie under /root/reference (oracle/Makefile extracts the
// lines into the build directory, deleted after linking): which calls move the collider (the first one, then per frame / per
// substep / per iteration as interpolateSubstep and interpolateIter say, only between the first and the last key frame) and the
// interpolation weights. The rest of the class loads .obj files through the mesh class; here the key meshes are flat arrays with
// the two whole-array operations the lines use. Everything lives in namespace GAIA_CO.
#include <algorithm>
#include <cassert>
#include <cstdint>
#include <iterator>
#include <vector>

namespace GAIA_CO {
typedef float FloatingType;
typedef int32_t IdType;
struct Arr { std::vector<float> d; };
inline Arr operator*(const Arr& a, float s) { Arr r = a; for (auto& x : r.d) x = x * s; return r; }
inline Arr operator+(const Arr& a, const Arr& b) { Arr r = a; for (size_t i = 0; i < r.d.size(); i++) r.d[i] = a.d[i] + b.d[i]; return r; }
struct RefKeyMesh { Arr p; const Arr& positions() const { return p; } };
struct RefSeqParams { std::vector<int> keyFrames; bool interpolateSubstep = true; int interpolateIter = 1; };
struct ColliderTrimeshSequence {
    RefSeqParams par;
    Arr cur;
    bool updated = false;
    RefSeqParams& colliderParameters() { return par; }
    Arr& positions() { return cur; }
#include "ref_collider_seq_h_59_124.inc"
    std::vector<RefKeyMesh> meshes{};
    bool first_time = true;
};
}  // namespace GAIA_CO

// calls [nCalls][3] = (frameId, substep, iteration) in the order the step makes them; out [nCalls][3 nV] = the collider after each call,
// updatedOut [nCalls] = its `updated` flag
extern "C" void ref_collider_sequence(int nV, int nKeys, const int32_t* keyFrames, const float* keyPos, int interpolateSubstep, int interpolateIter, int nCalls,
                                      const int32_t* calls, int numSubsteps, int numIters, float* out, int32_t* updatedOut) {
    GAIA_CO::ColliderTrimeshSequence c;
    c.par.keyFrames.assign(keyFrames, keyFrames + nKeys); c.par.interpolateSubstep = interpolateSubstep != 0; c.par.interpolateIter = interpolateIter;
    c.meshes.resize(nKeys);
    for (int k = 0; k < nKeys; k++) c.meshes[k].p.d.assign(keyPos + 3 * (size_t)nV * k, keyPos + 3 * (size_t)nV * (k + 1));
    c.cur.d.assign(3 * (size_t)nV, 0.f);
    for (int i = 0; i < nCalls; i++) {
        c.update(calls[3 * i], calls[3 * i + 1], calls[3 * i + 2], (size_t)numSubsteps, (size_t)numIters);
        for (size_t k = 0; k < 3 * (size_t)nV; k++) out[3 * (size_t)nV * i + k] = c.cur.d[k];
        updatedOut[i] = c.updated ? 1 : 0;
    }
}
