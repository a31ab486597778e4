// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own CPU world-box term of the tet VBD path, compiled from where it lies under
// /root/reference (Simulator/Modules/VBD/VBDPhysics.cpp):
//   VBDPhysics::accumlateBoundaryForceAndHessian (CPU semantics: penalty + friction per violated axis)      :2821-2954
//   VBDPhysics::computeVertexFriction                                                                       :2563-2592
// (refshim.cu pins the GPU variant, which applies friction once; this is the variant the product implements.) oracle/Makefile
// extracts those line ranges into the build directory (deleted after linking); stand-ins: <Eigen/Dense> from ref_shim/stub and
// the members the lines read. Everything lives in namespace GAIA_BC.
#include <Eigen/Dense>
#include <cstdint>
#include <vector>

#define GAIA GAIA_BC
#define DEBUG_LVL_DEBUG_VEBOSE 3

namespace GAIA {
typedef float FloatingType;
typedef const float CFloatingType;
typedef int32_t IdType;
typedef Eigen::Matrix<FloatingType, 2, 1> Vec2;
typedef Eigen::Matrix<FloatingType, 3, 1> Vec3;
typedef Eigen::Matrix<FloatingType, 2, 2> Mat2;
typedef Eigen::Matrix<FloatingType, 3, 3> Mat3;
typedef Eigen::Matrix<FloatingType, 3, 2> Mat3x2;

struct TetMeshFEM {
    Vec3 x, xPrev;
    Vec3 vertex(IdType) const { return x; }
    Vec3 vertexPrevPos(IdType) const { return xPrev; }
};
struct RefPhysicsParams {
    float boundaryCollisionStiffness = 0.f, boundaryFrictionDynamic = 0.f, boundaryFrictionEpsV = 0.f, dt = 0.f, bowlRadius = 1.f;
    bool useBowlGround = false, bowlCap = false, usePlaneGround = true;
    Vec3 bowlCenter;
    Eigen::Matrix<FloatingType, 3, 2> worldBounds;
};
// debug recording arrays the lines touch inside debugOperation lambdas (never run here)
struct DebugCols {
    struct Col { void operator+=(const Vec3&) {} };
    Col col(IdType) { return Col(); }
    float& operator()(size_t, IdType) { static float sink; return sink; }
};
template <class F> static void debugOperation(int, F&&) {}

struct VBDPhysics {
    RefPhysicsParams p;
    RefPhysicsParams& physicsParams() { return p; }
    std::vector<DebugCols> boundaryCollisionForce, boundaryFrictionForce;
    void accumlateBoundaryForceAndHessian(TetMeshFEM* pMesh, IdType meshId, IdType vertexId, Vec3& force, Mat3& hessian, bool apply_friction);
    void computeVertexFriction(CFloatingType mu, CFloatingType lambda, const Mat3x2& T, const Vec2& u, CFloatingType epsU, Vec3& force, Mat3& hessian);
};
}  // namespace GAIA

using namespace GAIA;
#include "ref_vbdphysics_cpp_2563_2592.inc"
#include "ref_vbdphysics_cpp_2821_2954.inc"
#undef GAIA

extern "C" void ref_boundary_cpu(const float* p3, const float* pPrev3, const float* boundsMin3, const float* boundsMax3, float kB, float muB, float epsVB,
                                 float dt, int applyFriction, float* f3, float* h9) {
    GAIA_BC::VBDPhysics ph;
    ph.p.boundaryCollisionStiffness = kB; ph.p.boundaryFrictionDynamic = muB; ph.p.boundaryFrictionEpsV = epsVB; ph.p.dt = dt;
    for (int d = 0; d < 3; d++) { ph.p.worldBounds(d, 0) = boundsMin3[d]; ph.p.worldBounds(d, 1) = boundsMax3[d]; }
    ph.boundaryCollisionForce.resize(1); ph.boundaryFrictionForce.resize(1);
    GAIA_BC::TetMeshFEM m;
    m.x = GAIA_BC::Vec3(p3[0], p3[1], p3[2]); m.xPrev = GAIA_BC::Vec3(pPrev3[0], pPrev3[1], pPrev3[2]);
    GAIA_BC::Vec3 force; GAIA_BC::Mat3 H;
    ph.accumlateBoundaryForceAndHessian(&m, 0, 0, force, H, applyFriction != 0);
    for (int k = 0; k < 3; k++) f3[k] = force.d[k];
    for (int k = 0; k < 9; k++) h9[k] = H.d[k];
}

// the same function with useBowlGround: the inside of a sphere (center, radius; cap: also above the centre) as the ground
extern "C" void ref_boundary_bowl_cpu(const float* p3, const float* pPrev3, const float* center3, float radius, int cap, float kB, float muB, float epsVB,
                                      float dt, int applyFriction, float* f3, float* h9) {
    GAIA_BC::VBDPhysics ph;
    ph.p.boundaryCollisionStiffness = kB; ph.p.boundaryFrictionDynamic = muB; ph.p.boundaryFrictionEpsV = epsVB; ph.p.dt = dt;
    ph.p.useBowlGround = true; ph.p.bowlCap = cap != 0; ph.p.bowlRadius = radius; ph.p.bowlCenter = GAIA_BC::Vec3(center3[0], center3[1], center3[2]);
    ph.boundaryCollisionForce.resize(1); ph.boundaryFrictionForce.resize(1);
    GAIA_BC::TetMeshFEM m;
    m.x = GAIA_BC::Vec3(p3[0], p3[1], p3[2]); m.xPrev = GAIA_BC::Vec3(pPrev3[0], pPrev3[1], pPrev3[2]);
    GAIA_BC::Vec3 force; GAIA_BC::Mat3 H;
    ph.accumlateBoundaryForceAndHessian(&m, 0, 0, force, H, applyFriction != 0);
    for (int k = 0; k < 3; k++) f3[k] = force.d[k];
    for (int k = 0; k < 9; k++) h9[k] = H.d[k];
}
