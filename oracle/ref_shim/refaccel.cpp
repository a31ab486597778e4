// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own Chebyshev weights, compiled from where they lie under /root/reference:
//   VBDPhysics::getAcceleratorOmega (the tet step)                Simulator/Modules/VBD/VBDPhysics.cpp:3558-3574
//   ChebysevAccelerator (the cloth step)                          Simulator/Modules/SolverUtils/ChebysevAccelerator.h:6-38
// oracle/Makefile extracts the line ranges into the build directory (deleted after linking). Namespace GAIA_AC.
#include <cassert>
#include <memory>
#define GAIA GAIA_AC
#define SQR(x) ((x) * (x))        // Modules/Types/Types.h:17
namespace GAIA {
typedef float FloatingType;
typedef const float CFloatingType;
struct VBDPhysics { FloatingType getAcceleratorOmega(int order, CFloatingType pho, CFloatingType prevOmega); };
#include "ref_chebyshev_h_6_38.inc"
}  // namespace GAIA
using namespace GAIA;
#include "ref_vbdphysics_cpp_3558_3574.inc"
#undef GAIA

// the sequence omega_1 .. omega_n for one rho: which = 0 tet step, 1 cloth step
extern "C" void ref_accelerator_omegas(int which, float rho, int n, float* out) {
    float omega = 1.f;
    GAIA_AC::VBDPhysics ph;
    GAIA_AC::ChebysevAccelerator acc(rho);
    for (int order = 1; order <= n; order++) {
        omega = which == 0 ? ph.getAcceleratorOmega(order, rho, omega) : acc.getAcceleratorOmega(order, omega);
        out[order - 1] = omega;
    }
}
