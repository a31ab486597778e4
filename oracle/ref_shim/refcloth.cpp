// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own cloth material code, compiled from where it lies under /root/reference:
//   VBDTriMeshStVk::greenStrain, getCotangent, ::intializeEdgeQuadraticForms     Simulator/Modules/VBDCloth/VBDTriMeshStVK.cpp:40-88
//   VBDTriMeshStVk::accumlateMaterialForceAndHessian, ::accumlateStVKForceAndHessian,
//   ::accumlateBendingForceAndHessian                                            Simulator/Modules/VBDCloth/VBDTriMeshStVK.cpp:715-982
//   assembleMembraneHessian                                                      Simulator/Modules/VBDCloth/VBDBaseTriMesh.h:149-160
// oracle/Makefile extracts exactly those line ranges into the build directory (ref_cloth_*.inc, deleted after linking); nothing
// of the reference is copied into the repo. The lines are member functions of classes that need Eigen 3.4 and MeshFrame2
// (both absent), so this file supplies (i) a stand-in <Eigen/Dense> (ref_shim/stub/Eigen/Dense: the fixed-size operations those
// lines use, written for this repository) and (ii) a stand-in class with the accessors they call, backed by flat arrays the
// caller passes: positions, faces, per-face DmInv / rest area, vertex -> face and vertex -> hinge tables, edge infos. What IS
// the reference's: every arithmetic statement of the membrane force / 6x6 Hessian / damping, of the bending force / Hessian,
// and of the quadratic bending forms. What is NOT: the mesh set-up that produces DmInv, rest areas and the tables
// (TriMeshFEM::initialize, Eigen + MeshFrame2) - those come from the oracle and stay restatements.
#include <Eigen/Dense>
#include <cstdint>
#include <cstring>
#include <vector>

#define SQR(x) ((x) * (x))

namespace GAIA {
typedef float FloatingType;
typedef const float CFloatingType;
typedef int32_t IdType;
typedef Eigen::Matrix<FloatingType, 2, 1> Vec2;
typedef Eigen::Matrix<FloatingType, 3, 1> Vec3;
typedef const Eigen::Matrix<FloatingType, 3, 1> CVec3;
typedef Eigen::Matrix<FloatingType, 4, 1> Vec4;
typedef Eigen::Matrix<FloatingType, 2, 2> Mat2;
typedef Eigen::Matrix<FloatingType, 3, 3> Mat3;
typedef Eigen::Matrix<FloatingType, 4, 4> Mat4;
typedef Eigen::Matrix<FloatingType, 6, 6> Mat6;
typedef Eigen::Matrix<FloatingType, 3, 2> Mat3x2;

// Simulator/Modules/TriMesh/TriMesh.h:51-60 (field names as the extracted lines use them)
struct EdgeInfo { int eV1, eV2, eV12Next, eV21Next, fId1, fId2; };
struct RefTopology { std::vector<EdgeInfo> edgeInfos; };
struct RefObjectParams { float lambda, miu, dampingStVK, bendingStiffness, dampingBending; };
struct RefPhysicsParams { float dt, degenerateTriangleThres; };

template <class F> static void cpu_parallel_for(int b, int e, F&& fn) { for (int i = b; i < e; i++) fn(i); }

#include "ref_cloth_assemble_149_160.inc"

struct VBDTriMeshStVk {
    // flat inputs
    int nV = 0, nF = 0;
    const float* pos = nullptr; const float* prev = nullptr;
    const int32_t* faces = nullptr;              // 3 per face
    const float* dmInv = nullptr;                // 4 per face, column-major 2x2
    const float* area = nullptr;
    const int32_t *vfOff = nullptr, *vfFace = nullptr, *vfOrder = nullptr;      // vertex -> incident faces
    const int32_t *vbOff = nullptr, *vbEdge = nullptr, *vbOrder = nullptr;      // vertex -> hinges it takes part in
    RefObjectParams obj{};
    RefPhysicsParams phys{};
    RefPhysicsParams* pPhysicsParams = &phys;
    RefTopology topo;
    RefTopology* pTopology = &topo;
    std::vector<Mat4> edgeLaplacianQuadraticForms;

    // accessors the extracted lines call (TriMeshFEM / VBDBaseTriMesh members in the reference)
    const RefObjectParams& objectParams() const { return obj; }
    Vec3 vertex(size_t i) const { return Vec3(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]); }
    Vec3 vertexPrevPos(size_t i) const { return Vec3(prev[3 * i], prev[3 * i + 1], prev[3 * i + 2]); }
    size_t numEdges() const { return topo.edgeInfos.size(); }
    size_t numNeiFaces(int iV) const { return (size_t)(vfOff[iV + 1] - vfOff[iV]); }
    int getVertexIthNeiFace(int iV, size_t i) const { return vfFace[vfOff[iV] + i]; }
    IdType getVertexIthNeiFaceOrder(int iV, size_t i) const { return vfOrder[vfOff[iV] + i]; }
    size_t numRelevantBendings(int iV) const { return (size_t)(vbOff[iV + 1] - vbOff[iV]); }
    IdType getVertexIthRelevantBending(int iV, size_t i) const { return vbEdge[vbOff[iV] + i]; }
    IdType getVertexIthRelevantBendingOrder(int iV, size_t i) const { return vbOrder[vbOff[iV] + i]; }
    const EdgeInfo& getEdgeInfo(IdType e) const { return topo.edgeInfos[e]; }
    FloatingType faceRestposeArea(int f) const { return area[f]; }
    Mat2 getDmInv(int f) const { Mat2 m; for (int k = 0; k < 4; k++) m.d[k] = dmInv[4 * f + k]; return m; }
    // TriMeshFEM::computeNormal (Simulator/Modules/TriMesh/TriMesh.h:677-690) and ::calculateDeformationGradient
    // (Simulator/Modules/TriMesh/TriMesh.cpp:200-211): four lines each, restated over the flat arrays
    Vec3 computeNormal(int f, bool normalize = true) const {
        Vec3 v1 = vertex(faces[3 * f + 1]) - vertex(faces[3 * f]);
        Vec3 v2 = vertex(faces[3 * f + 2]) - vertex(faces[3 * f]);
        Vec3 normal = v1.cross(v2);
        if (normalize) normal.normalize();
        return normal;
    }
    void calculateDeformationGradient(int f, Mat3x2& F) const {
        Vec3 v1 = vertex(faces[3 * f + 1]) - vertex(faces[3 * f]);
        Vec3 v2 = vertex(faces[3 * f + 2]) - vertex(faces[3 * f]);
        F.col(0) = v1;
        F.col(1) = v2;
        F = F * getDmInv(f);
    }
    // defined by the extracted lines
    Mat2 greenStrain(const Mat3x2& F);
    void intializeEdgeQuadraticForms();
    void accumlateMaterialForceAndHessian(int iV, Vec3& grad, Mat3& hessian);
    void accumlateStVKForceAndHessian(int iV, Vec3& force, Mat3& hessian);
    void accumlateBendingForceAndHessian(int vId, Vec3& force, Mat3& hessian);
};
}  // namespace GAIA

using namespace GAIA;
#include "ref_cloth_stvk_40_88.inc"
#include "ref_cloth_stvk_715_982.inc"

static void fill_edges(VBDTriMeshStVk& m, int nE, const int32_t* edge6) {
    m.topo.edgeInfos.resize(nE);
    for (int e = 0; e < nE; e++) {
        EdgeInfo& E = m.topo.edgeInfos[e];
        E.eV1 = edge6[6 * e]; E.eV2 = edge6[6 * e + 1]; E.eV12Next = edge6[6 * e + 2]; E.eV21Next = edge6[6 * e + 3];
        E.fId1 = edge6[6 * e + 4]; E.fId2 = edge6[6 * e + 5];
    }
}

extern "C" {

// VBDTriMeshStVk::intializeEdgeQuadraticForms on the REST positions: Q16 [nE][16] column-major (boundary edges stay zero)
void ref_cloth_bending_forms(int nV, const float* restPos, int nF, const int32_t* faces, int nE, const int32_t* edge6, float* Q16) {
    VBDTriMeshStVk m;
    m.nV = nV; m.nF = nF; m.pos = restPos; m.prev = restPos; m.faces = faces;
    fill_edges(m, nE, edge6);
    m.intializeEdgeQuadraticForms();
    for (int e = 0; e < nE; e++) std::memcpy(Q16 + 16 * (size_t)e, m.edgeLaplacianQuadraticForms[e].d, 16 * sizeof(float));
}

// VBDTriMeshStVk::accumlateMaterialForceAndHessian (membrane + bending, both with their damping) for vertex iV.
// Q16: the forms ref_cloth_bending_forms returned (the reference computes them once at set-up, as here)
void ref_cloth_material_vertex(int nV, const float* pos, const float* prev, int nF, const int32_t* faces, const float* dmInv4, const float* area,
                               const int32_t* vfOff, const int32_t* vfFace, const int32_t* vfOrder, int nE, const int32_t* edge6, const float* Q16,
                               const int32_t* vbOff, const int32_t* vbEdge, const int32_t* vbOrder, float miu, float lambda, float dampingStVK,
                               float bendingStiffness, float dampingBending, float dt, float degenerateTriangleThres, int iV, int which,
                               float* f3, float* h9) {
    VBDTriMeshStVk m;
    m.nV = nV; m.nF = nF; m.pos = pos; m.prev = prev; m.faces = faces; m.dmInv = dmInv4; m.area = area;
    m.vfOff = vfOff; m.vfFace = vfFace; m.vfOrder = vfOrder; m.vbOff = vbOff; m.vbEdge = vbEdge; m.vbOrder = vbOrder;
    m.obj.miu = miu; m.obj.lambda = lambda; m.obj.dampingStVK = dampingStVK; m.obj.bendingStiffness = bendingStiffness; m.obj.dampingBending = dampingBending;
    m.phys.dt = dt; m.phys.degenerateTriangleThres = degenerateTriangleThres;
    fill_edges(m, nE, edge6);
    m.edgeLaplacianQuadraticForms.resize(nE);
    for (int e = 0; e < nE; e++) std::memcpy(m.edgeLaplacianQuadraticForms[e].d, Q16 + 16 * (size_t)e, 16 * sizeof(float));
    Vec3 force; Mat3 H;
    if (which == 0) m.accumlateMaterialForceAndHessian(iV, force, H);
    else if (which == 1) m.accumlateStVKForceAndHessian(iV, force, H);
    else m.accumlateBendingForceAndHessian(iV, force, H);
    for (int k = 0; k < 3; k++) f3[k] = force.d[k];
    for (int k = 0; k < 9; k++) h9[k] = H.d[k];
}

}  // extern "C"
