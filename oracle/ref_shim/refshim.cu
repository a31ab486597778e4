// ORACLE — TEST INFRASTRUCTURE ONLY.
// Thin extern "C" shim over the REFERENCE's own __host__ __device__ VBD math. The reference translation unit
// (Simulator/Modules/VBD/VBDPhysicsCompute.cu) is #included from where it lies under /root/reference (path injected
// by oracle/Makefile as REF_VBD_CU); this file only builds the pointer structs the reference functions expect
// (its "run GPU code on CPU" debug mode, M/VBD/VBDPhysics.cpp:653-683) and forwards the calls. Output: oracle/_ref/.
#include REF_VBD_CU

using namespace GAIA;

extern "C" {

// GAIA::accumulateMaterialForceAndHessianForVertex_NeoHookean (VBDPhysicsCompute.cu:1694-1709)
void ref_neohookean_vertex(int nV, int nT, float* pos, float* prevPos, int* tets, float* DmInvs, float* restVol,
                           int* nbrInfos, int* nbrTets, int* nbrOrder, float miu, float lmbd, float dampH, float dampD,
                           float dt, int iV, float* force3, float* h9) {
    TetMeshTopologyGPU topo;
    topo.tetVIds = tets; topo.vertexNeighborTets = nbrTets; topo.vertexNeighborTets_vertexOrder = nbrOrder;
    topo.vertexNeighborTets_infos = nbrInfos; topo.nVerts = nV; topo.nTets = nT;
    VBDTetMeshNeoHookeanGPU m;
    memset(&m, 0, sizeof(m));
    m.pTopology = &topo; m.DmInvs = DmInvs; m.tetRestVolume = restVol; m.vertPos = pos; m.vertPrevPos = prevPos;
    m.miu = miu; m.lmbd = lmbd; m.dampingHydrostatic = dampH; m.dampingDeviatoric = dampD;
    for (int i = 0; i < 3; i++) force3[i] = 0.f;
    for (int i = 0; i < 9; i++) h9[i] = 0.f;
    accumulateMaterialForceAndHessianForVertex_NeoHookean(&m, iV, force3, h9, dt);
}

// GAIA::accumlateCollisionForceAndHessianPerCollision_v2 (VBDPhysicsCompute.cu:1812-1930) for ONE v-f contact:
// the query vertex lives in mesh 0 (vertex 0), the face in mesh 1 (vertices 0,1,2)
void ref_collision_per_collision(const float* p, const float* pPrev, const float* tri /*9*/, const float* triPrev /*9*/,
                                 int order, float k, float mu0, float mu1, float epsV0, float epsV1, float dt,
                                 float* force3, float* h9) {
    VBDBaseTetMeshGPU m0, m1;
    memset(&m0, 0, sizeof(m0)); memset(&m1, 0, sizeof(m1));
    m0.vertPos = (float*)p; m0.vertPrevPos = (float*)pPrev; m0.frictionDynamic = mu0; m0.frictionEpsV = epsV0;
    m1.vertPos = (float*)tri; m1.vertPrevPos = (float*)triPrev; m1.frictionDynamic = mu1; m1.frictionEpsV = epsV1;
    VBDBaseTetMeshGPU* meshes[2] = {&m0, &m1};
    VBDPhysicsDataGPU phys;
    memset(&phys, 0, sizeof(phys));
    phys.tetMeshes = meshes; phys.numMeshes = 2; phys.collisionStiffness = k; phys.dt = dt;
    CollisionDataGPU cd;
    memset(&cd, 0, sizeof(cd));
    cd.closestSurfaceFaceVIds[0] = 0; cd.closestSurfaceFaceVIds[1] = 1; cd.closestSurfaceFaceVIds[2] = 2;
    cd.intersectedTMeshId = 1; cd.activeColliding = 1;
    for (int i = 0; i < 3; i++) force3[i] = 0.f;
    for (int i = 0; i < 9; i++) h9[i] = 0.f;
    accumlateCollisionForceAndHessianPerCollision_v2(&phys, &m0, cd, 0, order, force3, h9);
}

// GAIA::accumulateBoundaryForceAndHessianGPU (VBD_GeneralCompute.h:87-166); GPU-variant semantics: friction once
void ref_boundary(const float* p, const float* pPrev, const float* bounds6, float kB, float muB, float epsVB, float dt,
                  float* force3, float* h9) {
    VBDBaseTetMeshGPU m0;
    memset(&m0, 0, sizeof(m0));
    m0.vertPos = (float*)p; m0.vertPrevPos = (float*)pPrev;
    VBDPhysicsDataGPU phys;
    memset(&phys, 0, sizeof(phys));
    for (int i = 0; i < 6; i++) phys.worldBounds[i] = bounds6[i];
    phys.boundaryCollisionStiffness = kB; phys.boundaryFrictionDynamic = muB; phys.boundaryFrictionEpsV = epsVB; phys.dt = dt;
    for (int i = 0; i < 3; i++) force3[i] = 0.f;
    for (int i = 0; i < 9; i++) h9[i] = 0.f;
    accumulateBoundaryForceAndHessianGPU(&phys, &m0, 0, force3, h9);
}

// CuMatrix::solve3x3_psd_stable (CuMatrix.h:235-267)
int ref_solve3x3(const float* m9, const float* b3, float* out3) { return CuMatrix::solve3x3_psd_stable(m9, b3, out3) ? 1 : 0; }

// GAIA::dampVelocity (VBDPhysicsCompute.cu:1397-1413); GPU-variant clamp rule
void ref_damp_velocity(float* v3, float maxV, float expD, float constD) { dampVelocity(v3, maxV, expD, constD); }

// GAIA::evaluateNeoHookeanEnergy (VBDPhysicsCompute.cu:1775-1808): the per-tet elastic energy (minus its rest value) behind
// evaluateElasticEnergy_kernel - the same A (mu Phi_D + lmbd Phi_H) that evaluateVertexMeritEnergy sums over a vertex's tets
float ref_neohookean_tet_energy(float miu, float lmbd, float A, int tetId, const int* tetVIds4, float* vs, const float* DmInvs) {
    return evaluateNeoHookeanEnergy(miu, lmbd, A, tetId, tetVIds4, vs, DmInvs);
}

int refshim_ok() { return 1; }
}
