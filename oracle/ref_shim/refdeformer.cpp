// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own tet-path deformers, compiled from where they lie under /root/reference:
//   VBDBaseDeformer, DeformerRotator, DeformerTranslater            Simulator/Modules/VBD/VBDDeformer.h:6-16, 45-179
// oracle/Makefile extracts the line ranges into the build directory (deleted after linking). Stand-ins: <Eigen/Dense> from
// ref_shim/stub, a mesh whose vertex(i) is a view into a flat position array, and setFixed (VBDDeformer.h:20-43 touches the GPU
// buffers of the real mesh class) reduced to recording which vertices it fixes. Everything lives in namespace GAIA_DF.
#include <Eigen/Dense>
#include <cmath>
#include <cstdint>
#include <memory>
#include <vector>

#define GAIA GAIA_DF
namespace Eigen { typedef Matrix<float, 3, 3> Matrix3f; }

namespace GAIA {
typedef float FloatingType;
typedef Eigen::Matrix<FloatingType, 3, 1> Vec3;
struct Vec3Block {      // Eigen::Block of the position matrix: reads as a Vec3, assigns through
    float* p;
    operator Vec3() const { return Vec3(p[0], p[1], p[2]); }
    float& operator[](int i) { return p[i]; }
    Vec3Block& operator=(const Vec3& v) { p[0] = v.d[0]; p[1] = v.d[1]; p[2] = v.d[2]; return *this; }
};
struct RefMesh { std::vector<float> pos; std::vector<int> fixedByDeformer; Vec3Block vertex(int i) { return Vec3Block{&pos[3 * (size_t)i]}; } };
typedef RefMesh VBDBaseTetMesh;
typedef RefMesh TetMeshFEM;
struct VBDPhysics { std::vector<std::shared_ptr<RefMesh>> tMeshes; };
inline void setFixed(VBDPhysics& physics, const std::vector<int>& selectedMeshes, const std::vector<std::vector<int>>& selectedVerts) {
    for (size_t i = 0; i < selectedMeshes.size(); ++i) for (int v : selectedVerts[i]) physics.tMeshes[selectedMeshes[i]]->fixedByDeformer.push_back(v);
}
#include "ref_vbddeformer_h_6_16.inc"
#include "ref_vbddeformer_h_45_179.inc"
}  // namespace GAIA
#undef GAIA

// One mesh, one deformer on `verts` (type 0: rotator about `a3` = axis through `c3` = centre with angular velocity w; type 1: translator
// with speed a3), created at positions pos0, then applied at nTimes times; out [nTimes][nSel][3] = the selected vertices after each
extern "C" void ref_deformer(int type, int nV, const float* pos0, int nSel, const int32_t* verts, const float* c3, const float* a3, double w, double endTime,
                             int nTimes, const double* times, float* out) {
    GAIA_DF::VBDPhysics ph;
    auto m = std::make_shared<GAIA_DF::RefMesh>();
    m->pos.assign(pos0, pos0 + 3 * (size_t)nV);
    ph.tMeshes.push_back(m);
    std::vector<int> meshes{0};
    std::vector<std::vector<int>> sel{std::vector<int>(verts, verts + nSel)};
    std::shared_ptr<GAIA_DF::VBDBaseDeformer> d;
    if (type == 0) d = std::make_shared<GAIA_DF::DeformerRotator>(GAIA_DF::Vec3(c3[0], c3[1], c3[2]), GAIA_DF::Vec3(a3[0], a3[1], a3[2]), w, ph, meshes, sel, endTime);
    else d = std::make_shared<GAIA_DF::DeformerTranslater>(GAIA_DF::Vec3(a3[0], a3[1], a3[2]), ph, meshes, sel, endTime);
    for (int t = 0; t < nTimes; t++) {
        (*d)(ph, times[t], 0, 0, 0, 0.0);
        for (int i = 0; i < nSel; i++) for (int k = 0; k < 3; k++) out[(t * (size_t)nSel + i) * 3 + k] = m->pos[3 * (size_t)verts[i] + k];
    }
}
