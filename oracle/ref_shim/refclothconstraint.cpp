// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own cloth sewing and vertex-face attachment terms, compiled from where they lie under
// /root/reference (Simulator/Modules/VBDCloth):
//   VBDBaseTriMesh::accumlateSewingForceAndHessian                                      VBDBaseTriMesh.cpp:320-344
//   VertexFaceAttachmentConstraint, VertexConstraints                                   VBDTriMeshConstraints.h:31-49
//   the attachment set-up loop of VBDClothSimulationFramework::initialize              VBDClothPhysics.cpp:161-220
//   VBDClothSimulationFramework::accumulateConstraintForceAndHessian                    VBDClothPhysics.cpp:1464-1498
// oracle/Makefile extracts those line ranges into the build directory (deleted after linking); stand-ins: <Eigen/Dense> from
// ref_shim/stub and the members the lines read (meshes over flat position arrays). Everything lives in namespace GAIA_CN.
#include <Eigen/Dense>
#include <array>
#include <cstdint>
#include <cstdlib>
#include <memory>
#include <utility>
#include <vector>

#define GAIA GAIA_CN
#define RELEASE_ASSERT(expr) do { if (!(expr)) std::abort(); } while (0)

namespace GAIA {
typedef float FloatingType;
typedef const float CFloatingType;
typedef int32_t IdType;
typedef Eigen::Matrix<FloatingType, 3, 1> Vec3;
typedef Eigen::Matrix<FloatingType, 3, 3> Mat3;
typedef Eigen::Matrix<int, 3, 1> Vec3I;

#include "ref_triconstraints_h_31_49.inc"

struct RefObjectParams { FloatingType sewingStiffness = 0.f; };
struct VBDBaseTriMesh {
    std::vector<float> pos;                       // 3 * nV
    RefObjectParams obj;
    RefObjectParams* pObjectParams = &obj;
    std::vector<std::array<int, 3>> sewingVertices;
    std::vector<FloatingType> sewingRatio;
    std::vector<std::vector<std::pair<int, int>>> sewingVMap;
    std::vector<VertexConstraints> vertexConstraints;
    Vec3 vertex(size_t i) const { return Vec3(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]); }
    void accumlateSewingForceAndHessian(int iV, Vec3& force, Mat3& hessian);
};
typedef VBDBaseTriMesh TriMeshFEM;

struct RefAttachmentParams {
    std::vector<IdType> vertexConstraintVertexInfos, vertexConstraintTargetFacesInfos;
    std::vector<std::array<FloatingType, 3>> vertexConstraintBarys;
    std::vector<FloatingType> vertexConstraintStiffness;
};

struct VBDClothSimulationFramework {
    std::vector<std::shared_ptr<VBDBaseTriMesh>> meshes;
    int nSim = 0;
    size_t numSimulationMeshes() const { return (size_t)nSim; }
    VBDBaseTriMesh& getSimulatedMesh(size_t i) { return *meshes[i]; }
    TriMeshFEM& getMesh(size_t i) { return *meshes[i]; }
    void accumulateConstraintForceAndHessian(IdType iMesh, IdType vId, Vec3& force, Mat3& hessian);
    void setupAttachments(RefAttachmentParams& vfAttachmentConstraintParams) {
        size_t numConstraints = vfAttachmentConstraintParams.vertexConstraintVertexInfos.size() / 2;
#include "ref_clothphysics_cpp_161_220.inc"
    }
};
}  // namespace GAIA

using namespace GAIA;
#include "ref_basetrimesh_cpp_320_344.inc"
#include "ref_clothphysics_cpp_1464_1498.inc"
#undef GAIA

// sewing of one mesh: seams [nSeams][3], ratios; force / Hessian of vertex v at positions pos
extern "C" void ref_cloth_sewing(int nV, const float* pos, int nSeams, const int32_t* verts3, const float* ratios, float stiffness, int v, float* f3, float* h9) {
    GAIA_CN::VBDBaseTriMesh m;
    m.pos.assign(pos, pos + 3 * (size_t)nV);
    m.obj.sewingStiffness = stiffness;
    m.sewingRatio.assign(ratios, ratios + nSeams);
    m.sewingVMap.resize(nV);
    for (int i = 0; i < nSeams; i++) {        // TriMeshFEM::generate_sewing_filters, M/TriMesh/TriMesh.h:623-632 (needs the mesh class: restated)
        m.sewingVertices.push_back({verts3[3 * i], verts3[3 * i + 1], verts3[3 * i + 2]});
        for (int o = 0; o < 3; o++) m.sewingVMap[verts3[3 * i + o]].push_back({i, o});
    }
    GAIA_CN::Vec3 force; GAIA_CN::Mat3 H;
    m.accumlateSewingForceAndHessian(v, force, H);
    for (int k = 0; k < 3; k++) f3[k] = force.d[k];
    for (int k = 0; k < 9; k++) h9[k] = H.d[k];
}

// attachments: nMesh meshes (the first nSim simulated; positions concatenated, vOff [nMesh + 1]); nCon constraints given the
// way the reference's JSON gives them: vInfos [nCon][2] (mesh, vertex), fInfos [nCon][4] (mesh or -1, three face vertices),
// barys [nCon][3], stiffness [nCon] (or NULL: the struct's default). Output for every (mesh < nSim, vertex): force 3 + Hessian 9,
// and the number of constraint entries that vertex ended up with.
extern "C" void ref_cloth_attachments(int nMesh, int nSim, const int32_t* vOff, const float* pos, int nCon, const int32_t* vInfos, const int32_t* fInfos,
                                      const float* barys, const float* stiffness, float* out12, int32_t* nEntries) {
    GAIA_CN::VBDClothSimulationFramework fw;
    fw.nSim = nSim;
    for (int i = 0; i < nMesh; i++) {
        auto m = std::make_shared<GAIA_CN::VBDBaseTriMesh>();
        m->pos.assign(pos + 3 * (size_t)vOff[i], pos + 3 * (size_t)vOff[i + 1]);
        m->vertexConstraints.resize(vOff[i + 1] - vOff[i]);
        fw.meshes.push_back(m);
    }
    GAIA_CN::RefAttachmentParams P;
    P.vertexConstraintVertexInfos.assign(vInfos, vInfos + 2 * (size_t)nCon);
    P.vertexConstraintTargetFacesInfos.assign(fInfos, fInfos + 4 * (size_t)nCon);
    for (int i = 0; i < nCon; i++) P.vertexConstraintBarys.push_back({barys[3 * i], barys[3 * i + 1], barys[3 * i + 2]});
    if (stiffness) P.vertexConstraintStiffness.assign(stiffness, stiffness + nCon);
    fw.setupAttachments(P);
    for (int im = 0; im < nSim; im++)
        for (int v = 0; v < vOff[im + 1] - vOff[im]; v++) {
            GAIA_CN::Vec3 force; GAIA_CN::Mat3 H;
            fw.accumulateConstraintForceAndHessian(im, v, force, H);
            float* o = out12 + 12 * (size_t)(vOff[im] + v);
            for (int k = 0; k < 3; k++) o[k] = force.d[k];
            for (int k = 0; k < 9; k++) o[3 + k] = H.d[k];
            nEntries[vOff[im] + v] = (int)fw.meshes[im]->vertexConstraints[v].attachmentConstraints.size();
        }
}
