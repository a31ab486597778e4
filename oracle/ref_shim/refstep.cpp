// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own initial step and velocity update of the tet VBD path, compiled from where they lie
// under /root/reference (Simulator/Modules/VBD):
//   VBDBaseTetMesh::forwardStep, ::forwardStepSimplecticEuler             VBD_BaseMaterial.cpp:7-132, 134-147
//   VBDBaseTetMesh::computeInertia, ::applyInitialStep, ::handleVelocityConstraint      VBD_BaseMaterial.h:201-221, 244-259, 300-306
//   VBDPhysics::updateVelocities, ::updateVelocity, ::dampVelocities     VBDPhysics.cpp:2486-2506, 2514-2518, 2520-2561
// oracle/Makefile extracts those line ranges into the build directory (deleted after linking). The classes need Eigen,
// MeshFrame2 and the whole framework, so this file supplies stand-ins with the members those lines read: 3 x N arrays with
// the whole-array operations the lines use (sum, difference, scalar multiple, colwise +=, column access - all elementwise),
// parameter structs, and a cpu_parallel_for that runs serially. Everything lives in namespace GAIA_ST.
#include <Eigen/Dense>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define GRAVITY_AXIS 1      // Modules/Types/Types.h:9
#define GAIA GAIA_ST

namespace GAIA {
typedef float FloatingType;
typedef const float CFloatingType;
typedef int32_t IdType;
typedef Eigen::Matrix<FloatingType, 3, 1> Vec3;

// a 3 x N float array (Eigen::Matrix<float, 3, Dynamic> in the reference): elementwise whole-array operations only
struct Mat3N {
    std::vector<float> d;      // column-major
    int n = 0;
    void resize(int cols) { n = cols; d.assign(3 * (size_t)cols, 0.f); }
    float& operator()(int r, int c) { return d[3 * (size_t)c + r]; }
    float operator()(int r, int c) const { return d[3 * (size_t)c + r]; }
    struct Col : Vec3 {
        Mat3N* o; int j;
        Col(Mat3N* o_, int j_) : o(o_), j(j_) { for (int k = 0; k < 3; k++) this->d[k] = (*o_)(k, j_); }
        Col& operator=(const Vec3& v) { for (int k = 0; k < 3; k++) { (*o)(k, j) = v.d[k]; this->d[k] = v.d[k]; } return *this; }
        Col& operator=(const Col& v) { return *this = static_cast<const Vec3&>(v); }
        Col& operator+=(const Vec3& v) { for (int k = 0; k < 3; k++) { (*o)(k, j) += v.d[k]; this->d[k] = (*o)(k, j); } return *this; }
        Col& operator*=(float s) { for (int k = 0; k < 3; k++) { (*o)(k, j) *= s; this->d[k] = (*o)(k, j); } return *this; }
        void setZero() { for (int k = 0; k < 3; k++) { (*o)(k, j) = 0.f; this->d[k] = 0.f; } }
    };
    Col col(int j) { return Col(this, j); }
    struct Colwise { Mat3N* o; void operator+=(const Vec3& v) { for (int j = 0; j < o->n; j++) for (int k = 0; k < 3; k++) (*o)(k, j) += v.d[k]; } };
    Colwise colwise() { return Colwise{this}; }
};
inline Mat3N operator+(const Mat3N& a, const Mat3N& b) { Mat3N m = a; for (size_t i = 0; i < m.d.size(); i++) m.d[i] = a.d[i] + b.d[i]; return m; }
inline Mat3N operator-(const Mat3N& a, const Mat3N& b) { Mat3N m = a; for (size_t i = 0; i < m.d.size(); i++) m.d[i] = a.d[i] - b.d[i]; return m; }
inline Mat3N operator*(float s, const Mat3N& a) { Mat3N m = a; for (size_t i = 0; i < m.d.size(); i++) m.d[i] = s * a.d[i]; return m; }
inline Mat3N operator/(const Mat3N& a, float s) { Mat3N m = a; for (size_t i = 0; i < m.d.size(); i++) m.d[i] = a.d[i] / s; return m; }

struct RefPhysicsParams { float dt = 0.f; Vec3 gravity; bool associateGravityWithInertia = true; };
struct RefObjectParams {
    int initializationType = 0; float initRatio_g = 0.f; std::vector<int> fixedPoints;
    float maxVelocityMagnitude = -1.f, noGravZoneThreshold = 0.f, exponentialVelDamping = 1.f, constantVelDamping = 0.f; bool hasNoGravZone = false;
};

template <class F> static void cpu_parallel_for(int b, int e, F&& fn) { for (int i = b; i < e; i++) fn(i); }

struct VBDBaseTetMesh {
    Mat3N mVertPos, mVertPrevPos, mVelocity, mVelocitiesPrev, inertia, acceletration, vertexExternalForces, vertexInternalForces;
    const float* internalForcesNow = nullptr;       // the caller's elastic forces for this substep (evaluateInternalForce needs the mesh class)
    std::vector<float> invMass;
    RefPhysicsParams* pPhysicsParams = nullptr;
    RefObjectParams* pObjParamsVBD = nullptr;
    RefObjectParams* pObjectParams = nullptr;
    bool hasVelocitiesPrev = false, hasApproxAcceleration = false, activeForMaterialSolve = true;
    int numVertices() const { return mVertPos.n; }
    Mat3N& positions() { return mVertPos; }
    Mat3N& positionsPrev() { return mVertPrevPos; }
    Mat3N& velocities() { return mVelocity; }
    Mat3N& velocitiesPrev() { return mVelocitiesPrev; }
    Mat3N::Col vertex(size_t i) { return mVertPos.col((int)i); }
    Mat3N::Col vertexPrevPos(size_t i) { return mVertPrevPos.col((int)i); }
    Mat3N::Col velocity(size_t i) { return mVelocity.col((int)i); }
    float vertexInvMass(int i) const { return invMass[i]; }
    void evaluateExternalForce() {}                 // VBD_BaseMaterial.h:308-311: the external forces stay zero on this path
    void evaluateInternalForce() { for (size_t i = 0; i < vertexInternalForces.d.size(); i++) vertexInternalForces.d[i] = internalForcesNow ? internalForcesNow[i] : 0.f; }
    void forwardStepSimplecticEuler();              // initializationType 5: defined by the extracted lines
    // defined by the extracted lines
    void forwardStep();
    void computeInertia();
    void applyInitialStep();
    void handleVelocityConstraint();
};
typedef VBDBaseTetMesh TetMeshFEM;

struct VBDPhysics {
    std::vector<std::shared_ptr<VBDBaseTetMesh>> tMeshes;
    std::vector<RefObjectParams*> objParams;
    RefPhysicsParams phys;
    RefPhysicsParams& physicsParams() { return phys; }
    size_t numTetMeshes() const { return tMeshes.size(); }
    RefObjectParams& getObjectParam(size_t i) { return *objParams[i]; }
    void updateVelocities();
    void updateVelocity(TetMeshFEM* pMesh, IdType vertexId);
    void dampVelocities();
};

#include "ref_basemat_h_201_221.inc"
#include "ref_basemat_h_244_259.inc"
#include "ref_basemat_h_300_306.inc"
}  // namespace GAIA

using namespace GAIA;
#include "ref_basemat_cpp_7_132.inc"
#include "ref_basemat_cpp_134_147.inc"
#include "ref_vbdphysics_cpp_2486_2506.inc"
#include "ref_vbdphysics_cpp_2514_2518.inc"
#include "ref_vbdphysics_cpp_2520_2561.inc"
#undef GAIA

// One mesh through nSteps substeps of: applyInitialStep -> (the solver is replaced by the caller's positions) -> updateVelocities.
// params12: dt, gravity xyz, associateGravityWithInertia, initializationType, initRatio_g, maxVelocityMagnitude, hasNoGravZone,
// noGravZoneThreshold, exponentialVelDamping, constantVelDamping. solved: [nSteps][3 nV] positions "after the iterations".
// Outputs per step: positions after the initial step, inertia, velocities after the update; each [nSteps][3 nV].
// internalForces: [nSteps][3 nV] elastic forces at the positions each substep starts from (read by type 5 only), or NULL;
// invMass: [nV] vertexInvMass (type 5 only), or NULL for 1.
extern "C" void ref_step_sequence(int nV, const float* pos0, const float* vel0, const float* params12, int nFixed, const int32_t* fixed, int nSteps,
                                  const float* solved, float* outInitPos, float* outInertia, float* outVel, const float* internalForces, const float* invMass) {
    GAIA_ST::VBDPhysics ph;
    GAIA_ST::RefObjectParams obj;
    ph.phys.dt = params12[0]; ph.phys.gravity = GAIA_ST::Vec3(params12[1], params12[2], params12[3]); ph.phys.associateGravityWithInertia = params12[4] != 0.f;
    obj.initializationType = (int)params12[5]; obj.initRatio_g = params12[6]; obj.maxVelocityMagnitude = params12[7]; obj.hasNoGravZone = params12[8] != 0.f;
    obj.noGravZoneThreshold = params12[9]; obj.exponentialVelDamping = params12[10]; obj.constantVelDamping = params12[11];
    obj.fixedPoints.assign(fixed, fixed + nFixed);
    auto m = std::make_shared<GAIA_ST::VBDBaseTetMesh>();
    m->pPhysicsParams = &ph.phys; m->pObjParamsVBD = &obj; m->pObjectParams = &obj;
    for (GAIA_ST::Mat3N* a : {&m->mVertPos, &m->mVertPrevPos, &m->mVelocity, &m->mVelocitiesPrev, &m->inertia, &m->acceletration, &m->vertexExternalForces,
                              &m->vertexInternalForces}) a->resize(nV);
    m->invMass.assign(nV, 1.f);
    if (invMass) m->invMass.assign(invMass, invMass + nV);
    std::memcpy(m->mVertPos.d.data(), pos0, 3 * (size_t)nV * sizeof(float));
    std::memcpy(m->mVelocity.d.data(), vel0, 3 * (size_t)nV * sizeof(float));
    ph.tMeshes.push_back(m); ph.objParams.push_back(&obj);
    const size_t S = 3 * (size_t)nV;
    for (int s = 0; s < nSteps; s++) {
        m->internalForcesNow = internalForces ? internalForces + s * S : nullptr;
        m->applyInitialStep();
        std::memcpy(outInitPos + s * S, m->mVertPos.d.data(), S * sizeof(float));
        std::memcpy(outInertia + s * S, m->inertia.d.data(), S * sizeof(float));
        std::memcpy(m->mVertPos.d.data(), solved + s * S, S * sizeof(float));
        ph.updateVelocities();
        std::memcpy(outVel + s * S, m->mVelocity.d.data(), S * sizeof(float));
    }
}
