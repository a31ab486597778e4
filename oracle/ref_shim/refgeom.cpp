// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own collision predicates: pointInTet, closestPointTriangle and
// get_closest_points_between_segments, Simulator/Modules/CollisionDetector/CollisionGeometry.h:56-267. That header also pulls
// in TriMesh.h / TetMeshFEM.h (Eigen, absent here), so oracle/Makefile extracts exactly lines 56-267 from the file where it
// lies under /root/reference into the build directory (ref_collision_geometry_56_267.inc, deleted after linking) and this
// file supplies the four names those lines use: the Embree math headers the reference vendors under Modules/common,
// FloatingType, CMP_EPSILON (Modules/Types/Types.h:20) and ClosestPointOnTriangleType
// (Modules/CollisionDetector/CollisionDetertionParameters.h:141-151). Nothing of the reference is copied into the repo.
#include "common/math/vec2.h"
#include "common/math/vec3.h"
#include "common/math/vec3fa.h"
#include <algorithm>
#include <cmath>
#include <cstdio>

typedef float FloatingType;
#define CMP_EPSILON 0.00001f
#define d_of(m, n, o, p) ((m.x - n.x) * (o.x - p.x) + (m.y - n.y) * (o.y - p.y) + (m.z - n.z) * (o.z - p.z))

namespace GAIA {
enum class ClosestPointOnTriangleType { AtA, AtB, AtC, AtAB, AtBC, AtAC, AtInterior, NotFound };
#include "ref_collision_geometry_56_267.inc"
}  // namespace GAIA

static inline embree::Vec3fa V(const float* p) { return embree::Vec3fa(p[0], p[1], p[2]); }

extern "C" {
int ref_point_in_tet(const float* p, const float* a, const float* b, const float* c, const float* d) {
    return GAIA::pointInTet(V(p), V(a), V(b), V(c), V(d)) ? 1 : 0;
}
int ref_closest_point_triangle(const float* p, const float* a, const float* b, const float* c, float* cp3, float* bary3) {
    embree::Vec3fa bary;
    GAIA::ClosestPointOnTriangleType type = GAIA::ClosestPointOnTriangleType::NotFound;
    const embree::Vec3fa cp = GAIA::closestPointTriangle(V(p), V(a), V(b), V(c), bary, type);
    cp3[0] = cp.x; cp3[1] = cp.y; cp3[2] = cp.z; bary3[0] = bary.x; bary3[1] = bary.y; bary3[2] = bary.z;
    return (int)type;
}
void ref_closest_points_segments(const float* p0, const float* p1, const float* q0, const float* q1, float* out8) {
    embree::Vec3fa ps, qt;
    float s = 0.f, t = 0.f;
    GAIA::get_closest_points_between_segments(V(p0), V(p1), V(q0), V(q1), ps, qt, s, t);
    out8[0] = ps.x; out8[1] = ps.y; out8[2] = ps.z; out8[3] = qt.x; out8[4] = qt.y; out8[5] = qt.z; out8[6] = s; out8[7] = t;
}
}

// ---- the cloth contact energy derivative, VBDClothSimulationFramework::computeContactRepulsiveForce
// (Simulator/Modules/VBDCloth/VBDClothPhysics.cpp:1533-1579): the member function's own lines, extracted by oracle/Makefile,
// compiled against a stand-in class that only offers the three physicsParams() fields the function reads.
typedef const float CFloatingType;
#define SQR(x) ((x) * (x))
namespace GAIA {
struct RefClothContactParams { float contactRadius, contactStiffness; int contactEnergyTypes; };
struct VBDClothSimulationFramework {
    RefClothContactParams p;
    RefClothContactParams& physicsParams() { return p; }
    void computeContactRepulsiveForce(CFloatingType dis_, CFloatingType thickness, FloatingType& dEdD, FloatingType& d2E_dDdD);
};
}  // namespace GAIA
#include "ref_contact_repulsive_1533_1579.inc"

extern "C" void ref_cloth_contact_repulsion(float dis, float thickness, float contactRadius, float contactStiffness, int energyType, float* dEdD, float* d2E) {
    GAIA::VBDClothSimulationFramework f;
    f.p.contactRadius = contactRadius; f.p.contactStiffness = contactStiffness; f.p.contactEnergyTypes = energyType;
    *dEdD = 0.f; *d2E = 0.f;
    f.computeContactRepulsiveForce(dis, thickness, *dEdD, *d2E);
}
