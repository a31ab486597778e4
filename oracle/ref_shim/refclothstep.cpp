// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own cloth initial step, velocity update and conservative-bound truncation, compiled from
// where they lie under /root/reference (Simulator/Modules/VBDCloth):
//   VBDBaseTriMesh::forwardStep_intertia, ::applyInitialStep, ::handleVeclocityConstraint     VBDBaseTriMesh.cpp:58-161, 163-178, 354-360
//   VBDClothSimulationFramework::updateVelocity                                               VBDClothPhysics.cpp:2318-2359
//   VBDClothSimulationFramework::applyConservativeBoundTruncation                             VBDClothPhysics.cpp:1223-1246
// oracle/Makefile extracts those line ranges into the build directory (deleted after linking). Stand-ins as in refstep.cpp:
// 3 x N arrays with the elementwise whole-array operations the lines use, parameter structs, a serial cpu_parallel_for.
// Everything lives in namespace GAIA_CS.
#include <Eigen/Dense>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define GRAVITY_AXIS 1      // Modules/Types/Types.h:9
#define GAIA GAIA_CS

namespace GAIA {
typedef float FloatingType;
typedef const float CFloatingType;
typedef int32_t IdType;
typedef Eigen::Matrix<FloatingType, 3, 1> Vec3;

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
    Vec3 col(int j) const { return Vec3((*this)(0, j), (*this)(1, j), (*this)(2, j)); }
    struct Colwise { Mat3N* o; void operator+=(const Vec3& v) { for (int j = 0; j < o->n; j++) for (int k = 0; k < 3; k++) (*o)(k, j) += v.d[k]; } };
    Colwise colwise() { return Colwise{this}; }
};
inline Mat3N operator+(const Mat3N& a, const Mat3N& b) { Mat3N m = a; for (size_t i = 0; i < m.d.size(); i++) m.d[i] = a.d[i] + b.d[i]; return m; }
inline Mat3N operator-(const Mat3N& a, const Mat3N& b) { Mat3N m = a; for (size_t i = 0; i < m.d.size(); i++) m.d[i] = a.d[i] - b.d[i]; return m; }
inline Mat3N operator*(float s, const Mat3N& a) { Mat3N m = a; for (size_t i = 0; i < m.d.size(); i++) m.d[i] = s * a.d[i]; return m; }
inline Mat3N operator/(const Mat3N& a, float s) { Mat3N m = a; for (size_t i = 0; i < m.d.size(); i++) m.d[i] = a.d[i] / s; return m; }

struct RefPhysicsParams { float dt = 0.f; Vec3 gravity; bool associateGravityWithInertia = true, handleCollision = true; };
struct RefObjectParams {
    int initializationType = 0; std::vector<int> fixedPoints;
    float maxVelocityMagnitude = -1.f, noGravZoneThreshold = 0.f, exponentialVelDamping = 1.f, constantVelDamping = 0.f; bool hasNoGravZone = false;
};
template <class F> static void cpu_parallel_for(int b, int e, F&& fn) { for (int i = b; i < e; i++) fn(i); }

struct VBDBaseTriMesh {
    Mat3N pos, positionsPrev, velocities, velocitiesPrev, inertia, acceletration, vertexExternalForces, positionsAtPrevCollisionDetection;
    std::vector<float> vertexConvervativeBounds;
    RefPhysicsParams* pPhysicsParams = nullptr;
    RefObjectParams* pObjectParams = nullptr;
    bool hasVelocitiesPrev = false, hasApproxAcceleration = false;
    size_t numVertices() const { return (size_t)pos.n; }
    Mat3N& positions() { return pos; }
    Mat3N::Col vertex(size_t i) { return pos.col((int)i); }
    float vertexInvMass(size_t) const { return 1.f; }
    const RefObjectParams& objectParamsBase() const { return *pObjectParams; }
    void evaluateExternalForce() {}                 // VBDBaseTriMesh.cpp:205-208: zero on this path
    // defined by the extracted lines
    void forwardStep_intertia();
    void applyInitialStep();
    void handleVeclocityConstraint();
};

struct VBDClothSimulationFramework {
    std::vector<std::shared_ptr<VBDBaseTriMesh>> baseTriMeshesForSimulation;
    std::vector<RefObjectParams*> objParams;
    RefPhysicsParams phys;
    RefPhysicsParams& physicsParams() { return phys; }
    VBDBaseTriMesh& getSimulatedMesh(size_t i) { return *baseTriMeshesForSimulation[i]; }
    RefObjectParams& getObjectParam(size_t i) { return *objParams[i]; }
    void updateVelocity();
    bool applyConservativeBoundTruncation(IdType iMesh, IdType vId, Vec3& newPos);
};
}  // namespace GAIA

using namespace GAIA;
#include "ref_basetrimesh_cpp_58_161.inc"
#include "ref_basetrimesh_cpp_163_178.inc"
#include "ref_basetrimesh_cpp_354_360.inc"
#include "ref_clothphysics_cpp_2318_2359.inc"
#include "ref_clothphysics_cpp_1223_1246.inc"
#undef GAIA

// One cloth through nSteps substeps of: applyInitialStep -> (solver replaced by the caller's positions) -> updateVelocity.
// params11: dt, gravity xyz, associateGravityWithInertia, handleCollision, initializationType, maxVelocityMagnitude, hasNoGravZone,
// noGravZoneThreshold, exponentialVelDamping; params11[10] is followed by constantVelDamping in params12[11].
// bounds [nV] (conservative bounds, constant over the sequence). Outputs per step, each [nSteps][3 nV]: positions after the initial
// step, inertia, velocities after the update.
extern "C" void ref_cloth_step_sequence(int nV, const float* pos0, const float* vel0, const float* params12, int nFixed, const int32_t* fixed, const float* bounds,
                                        int nSteps, const float* solved, float* outInitPos, float* outInertia, float* outVel) {
    GAIA_CS::VBDClothSimulationFramework fw;
    GAIA_CS::RefObjectParams obj;
    fw.phys.dt = params12[0]; fw.phys.gravity = GAIA_CS::Vec3(params12[1], params12[2], params12[3]); fw.phys.associateGravityWithInertia = params12[4] != 0.f;
    fw.phys.handleCollision = params12[5] != 0.f;
    obj.initializationType = (int)params12[6]; obj.maxVelocityMagnitude = params12[7]; obj.hasNoGravZone = params12[8] != 0.f; obj.noGravZoneThreshold = params12[9];
    obj.exponentialVelDamping = params12[10]; obj.constantVelDamping = params12[11];
    obj.fixedPoints.assign(fixed, fixed + nFixed);
    auto m = std::make_shared<GAIA_CS::VBDBaseTriMesh>();
    m->pPhysicsParams = &fw.phys; m->pObjectParams = &obj;
    for (GAIA_CS::Mat3N* a : {&m->pos, &m->positionsPrev, &m->velocities, &m->velocitiesPrev, &m->inertia, &m->acceletration, &m->vertexExternalForces,
                              &m->positionsAtPrevCollisionDetection}) a->resize(nV);
    m->vertexConvervativeBounds.assign(bounds, bounds + nV);
    std::memcpy(m->pos.d.data(), pos0, 3 * (size_t)nV * sizeof(float));
    std::memcpy(m->velocities.d.data(), vel0, 3 * (size_t)nV * sizeof(float));
    fw.baseTriMeshesForSimulation.push_back(m); fw.objParams.push_back(&obj);
    const size_t S = 3 * (size_t)nV;
    for (int s = 0; s < nSteps; s++) {
        m->applyInitialStep();
        std::memcpy(outInitPos + s * S, m->pos.d.data(), S * sizeof(float));
        std::memcpy(outInertia + s * S, m->inertia.d.data(), S * sizeof(float));
        std::memcpy(m->pos.d.data(), solved + s * S, S * sizeof(float));
        fw.updateVelocity();
        std::memcpy(outVel + s * S, m->velocities.d.data(), S * sizeof(float));
    }
}

// applyConservativeBoundTruncation for n candidate positions of vertex v_i: newPos [n][3] in / out, truncated [n] out
extern "C" void ref_cloth_truncate(int nV, const float* posAtCD, const float* bounds, int n, const int32_t* verts, float* newPos, int32_t* truncated) {
    GAIA_CS::VBDClothSimulationFramework fw;
    auto m = std::make_shared<GAIA_CS::VBDBaseTriMesh>();
    m->positionsAtPrevCollisionDetection.resize(nV);
    std::memcpy(m->positionsAtPrevCollisionDetection.d.data(), posAtCD, 3 * (size_t)nV * sizeof(float));
    m->vertexConvervativeBounds.assign(bounds, bounds + nV);
    fw.baseTriMeshesForSimulation.push_back(m);
    for (int i = 0; i < n; i++) {
        GAIA_CS::Vec3 p(newPos[3 * i], newPos[3 * i + 1], newPos[3 * i + 2]);
        truncated[i] = fw.applyConservativeBoundTruncation(0, verts[i], p) ? 1 : 0;
        newPos[3 * i] = p.d[0]; newPos[3 * i + 1] = p.d[1]; newPos[3 * i + 2] = p.d[2];
    }
}
