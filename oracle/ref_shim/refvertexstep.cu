// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own per-vertex step, VBDPhysics::VBDStepWithCollision (Simulator/Modules/VBD/VBDPhysics.cpp:
// 2402-2484), compiled from where it lies under /root/reference, with the reference's own CuMatrix::solve3x3_psd_stable
// (3rdParty/CuMatrix/CuMatrix/MatrixOps/CuMatrix.h). What it pins is what happens AFTER the terms are summed: the threshold on the
// force (CMP_EPSILON2), the 3x3 solve in float or double (useDouble3x3), the fall-back to a gradient step of stepSizeGD when the
// solve fails, the update x += stepSize * direction. The terms themselves (material, inertia, world box, contacts) are pinned one by
// one elsewhere; here the stand-in's material routine adds the caller's TOTAL force / Hessian and the other three add nothing, and
// the order in which the function calls them is recorded. Also here: VBDBaseTetMesh::accumlateInertiaForceAndHessian
// (VBD_BaseMaterial.h:359-366), compiled against a second stand-in mesh over the caller's masses / inertia / positions. oracle/Makefile extracts the lines into the build directory (deleted
// after linking). Stand-in <Eigen/Dense> from ref_shim/stub. Everything lives in namespace GAIA_VS.
#include <cstdint>
#include <cstdlib>
#include <cstdio>
#include <cmath>
#include <iostream>
#include <vector>
#include <Eigen/Dense>
#include <CuMatrix/MatrixOps/CuMatrix.h>

#define GAIA GAIA_VS
#define DEBUG_LVL_DEBUG_VEBOSE 3
#define CMP_EPSILON 0.00001f                       // Modules/Types/Types.h:20-21
#define CMP_EPSILON2 (CMP_EPSILON * CMP_EPSILON)

namespace GAIA {
typedef float FloatingType;
typedef int32_t IdType;
typedef Eigen::Matrix<FloatingType, 3, 1> Vec3;
typedef Eigen::Matrix<FloatingType, 3, 3> Mat3;
template <class F> static void debugOperation(int, F&&) {}
inline std::ostream& operator<<(std::ostream& o, const Vec3&) { return o; }
inline std::ostream& operator<<(std::ostream& o, const Mat3&) { return o; }
inline std::ostream& operator<<(std::ostream& o, const Eigen::Matrix<FloatingType, 1, 3>&) { return o; }

struct DebugCols { struct Col : Vec3 { void operator+=(const Vec3&) {} }; Col col(IdType) { return Col(); } };
struct ExtForces { Vec3 col(IdType) const { return Vec3(); } };      // vertexExternalForces: zero on this path
struct RefMesh {
    Vec3 x, fTotal; Mat3 hTotal;
    std::vector<int>* order = nullptr;
    ExtForces vertexExternalForces;
    void accumlateMaterialForceAndHessian2(IdType, Vec3& force, Mat3& h) { order->push_back(0); force += fTotal; h += hTotal; }
    void accumlateInertiaForceAndHessian(IdType, Vec3&, Mat3&) { order->push_back(1); }
    Vec3& vertex(IdType) { return x; }
};
typedef RefMesh TetMeshFEM;
typedef RefMesh VBDTetMeshNeoHookean;
struct RefPhysicsParams { FloatingType stepSize = 1.f, stepSizeGD = 1e-3f, tau = 0.5f, tau_GD = 0.1f; bool useDouble3x3 = false; };
struct VBDPhysics {
    RefPhysicsParams p;
    std::vector<int> order;
    std::vector<DebugCols> MaterialForce, InertiaForce;
    RefPhysicsParams& physicsParams() { return p; }
    void accumlateBoundaryForceAndHessian(TetMeshFEM*, IdType, IdType, Vec3&, Mat3&, bool) { order.push_back(2); }
    void accumlateCollisionForceAndHessian(TetMeshFEM*, IdType, IdType, Vec3&, Mat3&, bool) { order.push_back(3); }
    void VBDStepWithCollision(TetMeshFEM* pMesh_, IdType meshId, IdType vertexId, bool apply_friction);
};
}  // namespace GAIA

using namespace GAIA;
#include "ref_vbdphysics_cpp_2402_2484.inc"

namespace GAIA {
typedef const float CFloatingType;
struct RefInertiaParams { FloatingType dtSqrReciprocal = 0.f; };
struct Cols3 { const float* p = nullptr; Vec3 col(int i) const { return Vec3(p[3 * i], p[3 * i + 1], p[3 * i + 2]); } };
struct VBDBaseTetMesh {
    const float* mass = nullptr; Cols3 inertia, pos;
    RefInertiaParams par; RefInertiaParams* pPhysicsParams = &par;
    FloatingType vertexMass(int i) const { return mass[i]; }
    Vec3 vertex(int i) const { return pos.col(i); }
    void accumlateInertiaForceAndHessian(int iV, Vec3& force, Mat3& hessian);
};
#include "ref_basemat_h_359_366.inc"
}  // namespace GAIA
#undef GAIA

extern "C" void ref_inertia_terms(int n, const float* mass, const float* inertia, const float* x, float dtSqrReciprocal, float* f, float* h) {
    GAIA_VS::VBDBaseTetMesh m;
    m.mass = mass; m.inertia.p = inertia; m.pos.p = x; m.par.dtSqrReciprocal = dtSqrReciprocal;
    for (int i = 0; i < n; i++) {
        GAIA_VS::Vec3 force(f[3 * i], f[3 * i + 1], f[3 * i + 2]); GAIA_VS::Mat3 H;
        for (int k = 0; k < 9; k++) H.d[k] = h[9 * i + k];
        m.accumlateInertiaForceAndHessian(i, force, H);
        for (int k = 0; k < 3; k++) f[3 * i + k] = force.d[k];
        for (int k = 0; k < 9; k++) h[9 * i + k] = H.d[k];
    }
}

// n cases: total force f [n][3], total Hessian h [n][9] (column-major), position x [n][3] in / out; order4 [4]: the order in which
// the four accumulate routines were called in the last case (0 material, 1 inertia, 2 world box, 3 contacts)
extern "C" void ref_vertex_step_tail(int n, const float* f, const float* h, float* x, float stepSize, float stepSizeGD, int useDouble, int32_t* order4) {
    for (int i = 0; i < n; i++) {
        GAIA_VS::VBDPhysics ph;
        ph.p.stepSize = stepSize; ph.p.stepSizeGD = stepSizeGD; ph.p.useDouble3x3 = useDouble != 0;
        ph.MaterialForce.resize(1); ph.InertiaForce.resize(1);
        GAIA_VS::RefMesh m;
        m.order = &ph.order;
        m.x = GAIA_VS::Vec3(x[3 * i], x[3 * i + 1], x[3 * i + 2]);
        m.fTotal = GAIA_VS::Vec3(f[3 * i], f[3 * i + 1], f[3 * i + 2]);
        for (int k = 0; k < 9; k++) m.hTotal.d[k] = h[9 * i + k];
        ph.VBDStepWithCollision(&m, 0, 0, true);
        x[3 * i] = m.x.d[0]; x[3 * i + 1] = m.x.d[1]; x[3 * i + 2] = m.x.d[2];
        if (i == n - 1) for (int k = 0; k < 4; k++) order4[k] = ph.order[k];
    }
}

// the position update of the accelerator, the body of VBDSolveParallelGroup_applyAcceleration_kernel for one vertex
// (Simulator/Modules/VBD/VBDPhysicsCompute.cu:932-935: y = (y_k+1 - y_k-1) * omega + y_k-1, with CuMatrix's vec3 helpers)
extern "C" void ref_accelerate(int n, float* vAll, const float* yPrevPrevAll, float acceleratorOmega) {
    typedef float FloatingTypeGPU;
    for (int i = 0; i < n; i++) {
        float* v = vAll + 3 * i;
        const float* y_prevprev = yPrevPrevAll + 3 * i;
#include "ref_compute_cu_932_935.inc"
    }
}
