// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own CPU contact term of the tet VBD path, compiled from where it lies under /root/reference:
//   VBDPhysics::updateCollisionInfo (closest point, normal, penetration, friction, penalty force / Hessian)      Modules/VBD/VBDPhysics.cpp:2205-2311
//   VBDPhysics::accumlateCollisionForceAndHessianPerCollision (barycentric weights per vertex order)            Modules/VBD/VBDPhysics.cpp:3005-3047
//   VBDPhysics::computeVertexFriction                                                                           Modules/VBD/VBDPhysics.cpp:2563-2592
//   TetMeshFEM::computeFaceOrientedVolume, ::computeFaceNormal                                                   Modules/TetMesh/TetMeshFEM.h:562-575
//   loadVertexPos                                                                    Modules/CollisionDetector/DiscreteCollisionDetector.h:102-105
//   closestPointTriangle (with the other predicates of that range)                   Modules/CollisionDetector/CollisionGeometry.h:56-267
// with the Embree math headers the reference vendors under Modules/common. (refshim.cu pins the GPU variant of the same term.)
// oracle/Makefile extracts those line ranges into the build directory (deleted after linking); stand-ins: <Eigen/Dense> from
// ref_shim/stub and the members the lines read. Everything lives in namespace GAIA_CC.
#include "common/math/vec2.h"
#include "common/math/vec3.h"
#include "common/math/vec3fa.h"
#include <Eigen/Dense>
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <memory>
#include <vector>

#define GAIA GAIA_CC
#define DEBUG_LVL_DEBUG_VEBOSE 3
#define CMP_EPSILON 0.00001f
#define d_of(m, n, o, p) ((m.x - n.x) * (o.x - p.x) + (m.y - n.y) * (o.y - p.y) + (m.z - n.z) * (o.z - p.z))

namespace GAIA {
typedef float FloatingType;
typedef const float CFloatingType;
typedef int32_t IdType;
typedef Eigen::Matrix<FloatingType, 2, 1> Vec2;
typedef Eigen::Matrix<FloatingType, 3, 1> Vec3;
typedef Eigen::Matrix<FloatingType, 2, 2> Mat2;
typedef Eigen::Matrix<FloatingType, 3, 3> Mat3;
typedef Eigen::Matrix<FloatingType, 3, 2> Mat3x2;
enum class ClosestPointOnTriangleType { AtA, AtB, AtC, AtAB, AtBC, AtAC, AtInterior, NotFound };
#include "ref_collision_geometry_56_267.inc"

// positions: 3 x N with one float of padding (Vec3fa::loadu reads four); a column reads as a Vec3 and exposes its address
struct PosMat {
    std::vector<float> d;
    void set(const float* p, int n) { d.assign(3 * (size_t)n + 1, 0.f); for (size_t i = 0; i < 3 * (size_t)n; i++) d[i] = p[i]; }
    struct Col : Vec3 { const float* q; const float* data() const { return q; } };
    Col col(int j) const { Col c; c.q = &d[3 * (size_t)j]; for (int k = 0; k < 3; k++) c.d[k] = c.q[k]; return c; }
};
struct IntMat3 { std::vector<int32_t> v; int32_t operator()(int i, int j) const { return v[3 * (size_t)j + i]; } };
struct RefTopology { IntMat3 surfaceFacesTetMeshVIds; };      // the lines read it both as pTopology->surfaceFacesTetMeshVIds(i, f) and through the mesh
struct TetMeshFEM {
    PosMat mVertPos, prev;
    RefTopology topo;
    RefTopology* pTopology = &topo;
    const IntMat3& surfaceFacesTetMeshVIds() const { return topo.surfaceFacesTetMeshVIds; }
    Vec3 vertex(size_t i) const { return mVertPos.col((int)i); }
    Vec3 vertexPrevPos(size_t i) const { return prev.col((int)i); }
    void computeFaceOrientedVolume(int32_t surfaceFaceId, Vec3& orientedVolume);
    void computeFaceNormal(int32_t surfaceFaceId, Vec3& normal);
};
typedef TetMeshFEM VBDBaseTetMesh;
embree::Vec3fa loadVertexPos(TetMeshFEM* pTM, int32_t vId);
#include "ref_tetmesh_h_562_575.inc"

// Modules/CollisionDetector/CollisionDetertionParameters.h / Modules/VBD/VBDPhysics.h: the fields the lines read
struct CollidingPointInfo {
    bool shortestPathFound = true;
    int intersectedMeshId = 1, closestSurfaceFaceId = 0;
    Vec3 closestSurfacePt, closestSurfacePtBarycentrics, closestPointNormal;
};
struct VBDCollisionInfo { FloatingType penetrationDepth = 0.f; Vec3 collisionForce; Mat3 collisionHessian; };
template <class T> struct RefArray {
    std::vector<T> v;
    size_t size() const { return v.size(); }
    void insertN(size_t, size_t n) { v.resize(v.size() + n); }
    T& operator[](size_t i) { return v[i]; }
};
struct VBDCollisionDetectionResult {
    RefArray<VBDCollisionInfo> collisionForceAndHessian;
    std::vector<CollidingPointInfo> collidingPts;
    int idTMQuery = 0, idVQuery = 0;
    size_t numIntersections() const { return collidingPts.size(); }
};
struct RefObjectParams { float frictionDynamic = 0.f, frictionEpsV = 0.f; };
struct RefPhysicsParams { float dt = 0.f, collisionStiffness = 0.f; };
struct DebugCols { struct Col { void operator=(const Vec3&) {} void operator+=(const Vec3&) {} }; Col col(IdType) { return Col(); } };
template <class F> static void debugOperation(int, F&&) {}

struct VBDPhysics {
    std::vector<std::shared_ptr<VBDBaseTetMesh>> tMeshes;
    std::vector<RefObjectParams> obj;
    RefPhysicsParams p;
    std::vector<DebugCols> collisionForceAll;
    RefPhysicsParams& physicsParams() { return p; }
    RefObjectParams& getObjectParam(int i) { return obj[i]; }
    void updateCollisionInfo(VBDCollisionDetectionResult& collisionResult);
    void accumlateCollisionForceAndHessianPerCollision(int collisionId, int collisionVertexOrder, VBDCollisionDetectionResult& collisionResult, Vec3& force,
                                                       Mat3& hessian, IdType meshId, IdType vertexId, bool apply_friction);
    void computeVertexFriction(CFloatingType mu, CFloatingType lambda, const Mat3x2& T, const Vec2& u, CFloatingType epsU, Vec3& force, Mat3& hessian);
};
}  // namespace GAIA

using namespace GAIA;
#include "ref_dcd_h_102_105.inc"
#include "ref_vbdphysics_cpp_2563_2592.inc"
#include "ref_vbdphysics_cpp_2205_2311.inc"
#include "ref_vbdphysics_cpp_3005_3047.inc"
#undef GAIA

// ONE vertex-face contact: query vertex p (mesh 0, friction mu0 / epsV0) against the triangle tri (mesh 1, mu1 / epsV1);
// updateCollisionInfo, then the share of vertex `order` (0-2: the face's vertices, 3: the query vertex). Returns the penetration depth.
extern "C" float ref_contact_cpu(const float* p3, const float* pPrev3, const float* tri9, const float* triPrev9, int order, float k, float mu0, float mu1,
                                 float epsV0, float epsV1, float dt, float* f3, float* h9) {
    GAIA_CC::VBDPhysics ph;
    ph.p.dt = dt; ph.p.collisionStiffness = k;
    ph.obj.resize(2); ph.obj[0].frictionDynamic = mu0; ph.obj[0].frictionEpsV = epsV0; ph.obj[1].frictionDynamic = mu1; ph.obj[1].frictionEpsV = epsV1;
    ph.collisionForceAll.resize(2);
    auto m0 = std::make_shared<GAIA_CC::TetMeshFEM>(); auto m1 = std::make_shared<GAIA_CC::TetMeshFEM>();
    m0->mVertPos.set(p3, 1); m0->prev.set(pPrev3, 1);
    m1->mVertPos.set(tri9, 3); m1->prev.set(triPrev9, 3); m1->topo.surfaceFacesTetMeshVIds.v = {0, 1, 2};
    ph.tMeshes = {m0, m1};
    GAIA_CC::VBDCollisionDetectionResult res;
    res.collidingPts.resize(1);
    ph.updateCollisionInfo(res);
    GAIA_CC::Vec3 force; GAIA_CC::Mat3 H;
    ph.accumlateCollisionForceAndHessianPerCollision(0, order, res, force, H, order == 3 ? 0 : 1, order == 3 ? 0 : order, true);
    for (int i = 0; i < 3; i++) f3[i] = force.d[i];
    for (int i = 0; i < 9; i++) h9[i] = H.d[i];
    return res.collisionForceAndHessian[0].penetrationDepth;
}
