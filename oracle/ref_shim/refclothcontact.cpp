// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own cloth contact / boundary / friction code, compiled from where it lies under
// /root/reference (Simulator/Modules/VBDCloth/VBDClothPhysics.cpp):
//   VBDClothSimulationFramework::accumulateVFContactForceAndHessian      :1248-1321
//   ::accumulateEEContactForceAndHessian                                 :1323-1399
//   ::accumlateBoundaryForceAndHessian                                   :1401-1463
//   ::computeFriction                                                    :1500-1531
//   ::computeContactRepulsiveForce                                       :1533-1579
// oracle/Makefile extracts exactly lines 1248-1463 and 1500-1579 into the build directory (deleted after linking). They are
// members of a framework class that needs Eigen, MeshFrame2, Embree and oneTBB, so this file supplies a stand-in <Eigen/Dense>
// (ref_shim/stub/Eigen/Dense) and stand-in classes with the members those lines read, backed by flat arrays. The closest
// point of a vertex-face contact and the closest points of an edge-edge contact are INPUTS here (updateContactInfo = false):
// the caller computes them with the reference's own closestPointTriangle / get_closest_points_between_segments
// (ref_shim/refgeom.cpp). The whole file lives in namespace GAIA_CL (the extracted lines say GAIA::, renamed by a macro) so
// that its stand-in framework class does not collide with the one in refgeom.cpp.
#include <Eigen/Dense>
#include <cstdint>
#include <cstring>
#include <cstdio>
#include <memory>
#include <vector>

#define SQR(x) ((x) * (x))
#define CMP_EPSILON 0.00001f
#define GAIA GAIA_CL

namespace GAIA {
typedef float FloatingType;
typedef const float CFloatingType;
typedef int32_t IdType;
typedef Eigen::Matrix<FloatingType, 2, 1> Vec2;
typedef Eigen::Matrix<FloatingType, 3, 1> Vec3;
typedef const Eigen::Matrix<FloatingType, 3, 1> CVec3;
typedef Eigen::Matrix<FloatingType, 2, 2> Mat2;
typedef Eigen::Matrix<FloatingType, 3, 3> Mat3;
typedef Eigen::Matrix<FloatingType, 3, 2> Mat3x2;

struct EdgeInfo { int eV1, eV2, eV12Next, eV21Next, fId1, fId2; };
struct RefTopology { std::vector<EdgeInfo> edgeInfos; };
struct RefObjectParams { float frictionDynamic, frictionEpsV; };
struct TriMeshFEM {
    const float* pos = nullptr; const float* prev = nullptr; const int32_t* faces = nullptr;
    RefObjectParams obj{};
    RefObjectParams* pObjectParams = &obj;
    RefTopology topo;
    RefTopology* pTopology = &topo;
    Vec3 vertex(size_t i) const { return Vec3(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]); }
    Vec3 vertexPrevPos(size_t i) const { return Vec3(prev[3 * i], prev[3 * i + 1], prev[3 * i + 2]); }
    IdType facePosVId(int f, int k) const { return faces[3 * f + k]; }
};
typedef TriMeshFEM VBDBaseTriMesh;
// Simulator/Modules/CollisionDetector/ClothContactDetector.h:32-48, 86-103 (the fields the extracted lines read)
struct VFContactPointInfo {
    int contactVertexId = -1, contactVertexSideMeshId = -1, contactFaceId = -1, contactFaceSideMeshId = -1;
    Vec3 contactPoint, barycentrics;
};
struct EEContactPointInfo {
    int contactMeshId1 = -1, contactEdgeId1 = -1, contactMeshId2 = -1, contactEdgeId2 = -1;
    FloatingType mu1 = 0.f, mu2 = 0.f;
    Vec3 c1, c2;
};
struct ClothVFContactQueryResult { std::vector<VFContactPointInfo> contactPts; };
struct ClothEEContactQueryResult { std::vector<EEContactPointInfo> contactPts; };
struct RefClothPhysicsParams {
    float contactStiffness, contactRadius, thickness, dt, boundaryContactStiffness, boundaryFrictionDynamic, boundaryFrictionEpsValue;
    int contactEnergyTypes;
    Eigen::Matrix<FloatingType, 3, 2> worldBounds;
};
typedef std::vector<std::shared_ptr<TriMeshFEM>> MeshList;
// called only with updateContactInfo = true, which this shim never passes
static void updateVFContactPointInfo(MeshList&, VFContactPointInfo&) {}
static void updateEEContactPointInfo(MeshList&, EEContactPointInfo&) {}

struct VBDClothSimulationFramework {
    RefClothPhysicsParams p{};
    RefClothPhysicsParams& physicsParams() { return p; }
    MeshList triMeshesAll;
    CFloatingType accumulateVFContactForceAndHessian(IdType contactId, IdType contactVertexOrder, ClothVFContactQueryResult* pCollisionResult,
                                                     bool updateContactInfo, bool apply_friction, Vec3& force, Mat3& hessian);
    CFloatingType accumulateEEContactForceAndHessian(IdType contactId, IdType contactVertexOrder, ClothEEContactQueryResult* pCollisionResult,
                                                     bool updateContactInfo, bool apply_friction, Vec3& force, Mat3& hessian);
    void accumlateBoundaryForceAndHessian(VBDBaseTriMesh* pMesh, IdType iMesh, IdType vId, Vec3& force, Mat3& hessian);
    void computeFriction(CFloatingType mu, CFloatingType lambda, const Mat3x2& T, const Vec2& u, CFloatingType epsU, Vec3& force, Mat3& hessian);
    void computeContactRepulsiveForce(CFloatingType dis_, CFloatingType thickness, FloatingType& dEdD, FloatingType& d2E_dDdD);
};
}  // namespace GAIA

using namespace GAIA;
#include "ref_cloth_contact_1248_1463.inc"
#include "ref_cloth_contact_1500_1579.inc"
#undef GAIA

using namespace GAIA_CL;

namespace {
struct MeshIn { const float* pos; const float* prev; const int32_t* faces; int nE; const int32_t* edge6; float frictionDynamic, frictionEpsV; };
std::shared_ptr<TriMeshFEM> make_mesh(const MeshIn& in) {
    auto m = std::make_shared<TriMeshFEM>();
    m->pos = in.pos; m->prev = in.prev; m->faces = in.faces; m->obj.frictionDynamic = in.frictionDynamic; m->obj.frictionEpsV = in.frictionEpsV;
    m->topo.edgeInfos.resize(in.nE);
    for (int e = 0; e < in.nE; e++) {
        EdgeInfo& E = m->topo.edgeInfos[e];
        E.eV1 = in.edge6[6 * e]; E.eV2 = in.edge6[6 * e + 1]; E.eV12Next = in.edge6[6 * e + 2]; E.eV21Next = in.edge6[6 * e + 3];
        E.fId1 = in.edge6[6 * e + 4]; E.fId2 = in.edge6[6 * e + 5];
    }
    return m;
}
void out_fh(const Vec3& f, const Mat3& H, float* f3, float* h9) { for (int k = 0; k < 3; k++) f3[k] = f.d[k]; for (int k = 0; k < 9; k++) h9[k] = H.d[k]; }
}  // namespace

extern "C" {
// params8: contactStiffness, contactRadius, thickness, dt, boundaryContactStiffness, boundaryFrictionDynamic, boundaryFrictionEpsValue,
// contactEnergyTypes (as float). Mesh 0 = vertex side / edge 1, mesh 1 = face side / edge 2; fr4 = frictionDynamic, frictionEpsV of each.
static void set_params(VBDClothSimulationFramework& fw, const float* params8) {
    fw.p.contactStiffness = params8[0]; fw.p.contactRadius = params8[1]; fw.p.thickness = params8[2]; fw.p.dt = params8[3];
    fw.p.boundaryContactStiffness = params8[4]; fw.p.boundaryFrictionDynamic = params8[5]; fw.p.boundaryFrictionEpsValue = params8[6];
    fw.p.contactEnergyTypes = (int)params8[7];
}

float ref_cloth_vf_contact(const float* params8, const float* fr4, const float* posV, const float* prevV, int v, const float* posF, const float* prevF,
                           const int32_t* facesF, int f, const float* contactPoint3, const float* bary3, int vertexOrder, int applyFriction,
                           float* f3, float* h9) {
    VBDClothSimulationFramework fw;
    set_params(fw, params8);
    fw.triMeshesAll.push_back(make_mesh(MeshIn{posV, prevV, nullptr, 0, nullptr, fr4[0], fr4[1]}));
    fw.triMeshesAll.push_back(make_mesh(MeshIn{posF, prevF, facesF, 0, nullptr, fr4[2], fr4[3]}));
    ClothVFContactQueryResult res;
    res.contactPts.resize(1);
    VFContactPointInfo& c = res.contactPts[0];
    c.contactVertexSideMeshId = 0; c.contactVertexId = v; c.contactFaceSideMeshId = 1; c.contactFaceId = f;
    c.contactPoint = Vec3(contactPoint3[0], contactPoint3[1], contactPoint3[2]); c.barycentrics = Vec3(bary3[0], bary3[1], bary3[2]);
    Vec3 force; Mat3 H;
    const float d = fw.accumulateVFContactForceAndHessian(0, vertexOrder, &res, false, applyFriction != 0, force, H);
    out_fh(force, H, f3, h9);
    return d;
}

float ref_cloth_ee_contact(const float* params8, const float* fr4, const float* pos1, const float* prev1, int nE1, const int32_t* edges1, int e1,
                           const float* pos2, const float* prev2, int nE2, const int32_t* edges2, int e2, const float* c1, const float* c2, float mu1,
                           float mu2, int vertexOrder, int applyFriction, float* f3, float* h9) {
    VBDClothSimulationFramework fw;
    set_params(fw, params8);
    fw.triMeshesAll.push_back(make_mesh(MeshIn{pos1, prev1, nullptr, nE1, edges1, fr4[0], fr4[1]}));
    fw.triMeshesAll.push_back(make_mesh(MeshIn{pos2, prev2, nullptr, nE2, edges2, fr4[2], fr4[3]}));
    ClothEEContactQueryResult res;
    res.contactPts.resize(1);
    EEContactPointInfo& c = res.contactPts[0];
    c.contactMeshId1 = 0; c.contactEdgeId1 = e1; c.contactMeshId2 = 1; c.contactEdgeId2 = e2; c.mu1 = mu1; c.mu2 = mu2;
    c.c1 = Vec3(c1[0], c1[1], c1[2]); c.c2 = Vec3(c2[0], c2[1], c2[2]);
    Vec3 force; Mat3 H;
    const float d = fw.accumulateEEContactForceAndHessian(0, vertexOrder, &res, false, applyFriction != 0, force, H);
    out_fh(force, H, f3, h9);
    return d;
}

void ref_cloth_boundary(const float* params8, const float* boundsMin3, const float* boundsMax3, const float* pos, const float* prev, int v, float* f3, float* h9) {
    VBDClothSimulationFramework fw;
    set_params(fw, params8);
    for (int d = 0; d < 3; d++) { fw.p.worldBounds(d, 0) = boundsMin3[d]; fw.p.worldBounds(d, 1) = boundsMax3[d]; }
    auto m = make_mesh(MeshIn{pos, prev, nullptr, 0, nullptr, 0.f, 0.f});
    Vec3 force; Mat3 H;
    fw.accumlateBoundaryForceAndHessian(m.get(), 0, v, force, H);
    out_fh(force, H, f3, h9);
}
}  // extern "C"
