// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own cloth deformers, compiled from where they lie under /root/reference:
//   VBDClothBaseDeformer, ClothDeformerRotator, ClothDeformerTranslater      Simulator/Modules/VBDCloth/VBDClothDeformer.h:29-41, 43-199
// oracle/Makefile extracts the line ranges into the build directory (deleted after linking). Stand-ins as in refdeformer.cpp;
// setFixedPointCloth (:5-27) reduced to recording. Everything lives in namespace GAIA_CD.
#include <Eigen/Dense>
#include <cmath>
#include <cstdint>
#include <memory>
#include <vector>

#define GAIA GAIA_CD
namespace Eigen { typedef Matrix<float, 3, 3> Matrix3f; }

namespace GAIA {
typedef float FloatingType;
typedef Eigen::Matrix<FloatingType, 3, 1> Vec3;
struct Vec3Block {
    float* p;
    operator Vec3() const { return Vec3(p[0], p[1], p[2]); }
    float& operator[](int i) { return p[i]; }
    Vec3Block& operator=(const Vec3& v) { p[0] = v.d[0]; p[1] = v.d[1]; p[2] = v.d[2]; return *this; }
    Vec3Block& operator+=(const Vec3& v) { p[0] += v.d[0]; p[1] += v.d[1]; p[2] += v.d[2]; return *this; }
};
struct TriMeshFEM { std::vector<float> pos; std::vector<int> fixedByDeformer; Vec3Block vertex(int i) { return Vec3Block{&pos[3 * (size_t)i]}; } };
struct VBDClothSimulationFramework { std::vector<std::shared_ptr<TriMeshFEM>> baseTriMeshesForSimulation; };
inline void setFixedPointCloth(VBDClothSimulationFramework& physics, const std::vector<int>& selectedMeshes, const std::vector<std::vector<int>>& selectedVerts) {
    for (size_t i = 0; i < selectedMeshes.size(); ++i) for (int v : selectedVerts[i]) physics.baseTriMeshesForSimulation[selectedMeshes[i]]->fixedByDeformer.push_back(v);
}
#include "ref_clothdeformer_h_29_41.inc"
#include "ref_clothdeformer_h_43_199.inc"
}  // namespace GAIA
#undef GAIA

// One cloth, one deformer on `verts` (type 0 rotator: centre c3, axis a3, angular velocity w, end time t1; type 1 translator: speed
// a3 between t0 and t1), created at pos0, then applied once per substep at times[k] with step dt; out [nTimes][nSel][3]
extern "C" void ref_cloth_deformer(int type, int nV, const float* pos0, int nSel, const int32_t* verts, const float* c3, const float* a3, double w, float t0, float t1,
                                   int nTimes, const float* times, float dt, float* out) {
    GAIA_CD::VBDClothSimulationFramework ph;
    auto m = std::make_shared<GAIA_CD::TriMeshFEM>();
    m->pos.assign(pos0, pos0 + 3 * (size_t)nV);
    ph.baseTriMeshesForSimulation.push_back(m);
    std::vector<int> meshes{0};
    std::vector<std::vector<int>> sel{std::vector<int>(verts, verts + nSel)};
    std::shared_ptr<GAIA_CD::VBDClothBaseDeformer> d;
    if (type == 0) d = std::make_shared<GAIA_CD::ClothDeformerRotator>(GAIA_CD::Vec3(c3[0], c3[1], c3[2]), GAIA_CD::Vec3(a3[0], a3[1], a3[2]), w, ph, meshes, sel, (double)t1);
    else d = std::make_shared<GAIA_CD::ClothDeformerTranslater>(GAIA_CD::Vec3(a3[0], a3[1], a3[2]), ph, meshes, sel, t0, t1);
    for (int t = 0; t < nTimes; t++) {
        (*d)(ph, times[t], 0, 0, 0, dt);
        for (int i = 0; i < nSel; i++) for (int k = 0; k < 3; k++) out[(t * (size_t)nSel + i) * 3 + k] = m->pos[3 * (size_t)verts[i] + k];
    }
}
