// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own CCD solver: Simulator/Modules/CollisionDetector/CCDSolver.h with the vendored
// Simulator/3rdParty/CyCodeBase, compiled from where they lie under /root/reference (include paths from oracle/Makefile;
// `intrin.h` -> <x86intrin.h> forwarding header generated at build time). Output: oracle/_ref/.
#include "CCDSolver.h"

extern "C" {
// cy::IntersectContinuousTriPoint<float> (CCDSolver.h:170-202)
int ref_ccd_tri_point(const float* x0, const float* x1, const float* p0, const float* p1, float* t, float* bary3) {
    cy::Vec3<float> x[2][3], p[2];
    for (int k = 0; k < 3; k++) { x[0][k] = cy::Vec3<float>(x0[3 * k], x0[3 * k + 1], x0[3 * k + 2]); x[1][k] = cy::Vec3<float>(x1[3 * k], x1[3 * k + 1], x1[3 * k + 2]); }
    p[0] = cy::Vec3<float>(p0[0], p0[1], p0[2]); p[1] = cy::Vec3<float>(p1[0], p1[1], p1[2]);
    float tt = -1; cy::Vec3<float> b(0, 0, 0);
    bool hit = cy::IntersectContinuousTriPoint<float>(tt, x, p, b);
    *t = tt; bary3[0] = b.x; bary3[1] = b.y; bary3[2] = b.z;
    return hit ? 1 : 0;
}
// cy::CubicRoots<float> (cyPolynomial.h:799-902)
int ref_cubic_roots(const float* coef4, float x0, float x1, float* roots3) { return cy::CubicRoots<float>(roots3, coef4, x0, x1); }
}
