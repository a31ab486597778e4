// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h, oracle_cloth.h).
// The cloth VBD step: VBDClothSimulationFramework::runStep_CDEveryIter and everything it calls per vertex.
#include <algorithm>
#include <cmath>
#include "oracle_cloth.h"

namespace orc {

static const float CMP_EPSILON = 1e-5f;
static const int GRAVITY_AXIS = 1;

static inline void add_outer(M3& H, float s, V3 n) {   // H += s * n n^T
    const float v[3] = {n.x, n.y, n.z};
    for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) H(r, c) += s * v[r] * v[c];
}
static inline V3 mul(const M3& A, V3 x) {
    return V3(A(0, 0) * x.x + A(0, 1) * x.y + A(0, 2) * x.z, A(1, 0) * x.x + A(1, 1) * x.y + A(1, 2) * x.z, A(2, 0) * x.x + A(2, 1) * x.y + A(2, 2) * x.z);
}
static inline V3 normalized(V3 v) { const float n2 = sqnorm(v); return n2 > 0.f ? v / std::sqrt(n2) : v; }   // Eigen normalized()

// TriMeshFEM::calculateDeformationGradient (M/TriMesh/TriMesh.cpp:200-211): F = [x1-x0, x2-x0] * DmInv
static inline void deformation_gradient(const TriMesh& m, int f, float F[6]) {   // F column-major 3x2
    const V3 x0 = m.pos[m.faces[3 * f]];
    const V3 v1 = m.pos[m.faces[3 * f + 1]] - x0, v2 = m.pos[m.faces[3 * f + 2]] - x0;
    const float* I = &m.DmInv[4 * (size_t)f];
    for (int r = 0; r < 3; r++) { F[r] = v1[r] * I[0] + v2[r] * I[1]; F[3 + r] = v1[r] * I[2] + v2[r] * I[3]; }
}

// VBDTriMeshStVk::accumlateStVKForceAndHessian (M/VBDCloth/VBDTriMeshStVK.cpp:721-911) + assembleMembraneHessian
// (M/VBDCloth/VBDBaseTriMesh.h:149-160)
static void stvk_force_hessian(const ClothScene& s, const TriMesh& m, int iV, V3& force, M3& hessian) {
    const M3 old = hessian;
    const float lmbd = m.obj.lambda, miu = m.obj.miu;
    for (int i = m.vNbrFOff[iV]; i < m.vNbrFOff[iV + 1]; i++) {
        const int faceId = m.vNbrF[i], order = m.vNbrFOrder[i];
        float F[6]; deformation_gradient(m, faceId, F);
        const float F1_1 = F[0], F2_1 = F[1], F3_1 = F[2], F1_2 = F[3], F2_2 = F[4], F3_2 = F[5];
        // G = 0.5 (F^T F - I)
        const float G00 = 0.5f * ((F1_1 * F1_1 + F2_1 * F2_1 + F3_1 * F3_1) - 1.f), G11 = 0.5f * ((F1_2 * F1_2 + F2_2 * F2_2 + F3_2 * F3_2) - 1.f);
        const float G01 = 0.5f * (F1_1 * F1_2 + F2_1 * F2_2 + F3_1 * F3_2);
        const float A = m.restArea[faceId];
        const float tr = G00 + G11;
        const float S00 = 2.f * miu * G00 + lmbd * tr, S11 = 2.f * miu * G11 + lmbd * tr, S01 = 2.f * miu * G01;
        const float* I = &m.DmInv[4 * (size_t)faceId];
        const float Dm_inv1_1 = I[0], Dm_inv2_1 = I[1], Dm_inv1_2 = I[2], Dm_inv2_2 = I[3];
        // F12 = -A * F * S * DmInv^T   (3x2)
        float FS[6];
        for (int r = 0; r < 3; r++) { FS[r] = F[r] * S00 + F[3 + r] * S01; FS[3 + r] = F[r] * S01 + F[3 + r] * S11; }
        float F12[6];
        for (int r = 0; r < 3; r++) { F12[r] = -A * (FS[r] * Dm_inv1_1 + FS[3 + r] * Dm_inv1_2); F12[3 + r] = -A * (FS[r] * Dm_inv2_1 + FS[3 + r] * Dm_inv2_2); }
        const float e_uu = G00, e_vv = G11, e_uv = G01, e_sum = e_uu + e_vv;
        float D[6][6];
        D[0][0] = F1_1 * (F1_1 * lmbd + 2 * F1_1 * miu) + 2 * miu * e_uu + lmbd * e_sum + F1_2 * F1_2 * miu;
        D[0][1] = D[1][0] = F1_1 * (F2_1 * lmbd + 2 * F2_1 * miu) + F1_2 * F2_2 * miu;
        D[0][2] = D[2][0] = F1_1 * (F3_1 * lmbd + 2 * F3_1 * miu) + F1_2 * F3_2 * miu;
        D[0][3] = D[3][0] = 2 * miu * e_uv + F1_1 * F1_2 * lmbd + F1_1 * F1_2 * miu;
        D[0][4] = D[4][0] = F1_1 * F2_2 * lmbd + F1_2 * F2_1 * miu;
        D[0][5] = D[5][0] = F1_1 * F3_2 * lmbd + F1_2 * F3_1 * miu;
        D[1][1] = F2_1 * (F2_1 * lmbd + 2 * F2_1 * miu) + 2 * miu * e_uu + lmbd * e_sum + F2_2 * F2_2 * miu;
        D[1][2] = D[2][1] = F2_1 * (F3_1 * lmbd + 2 * F3_1 * miu) + F2_2 * F3_2 * miu;
        D[1][3] = D[3][1] = F1_2 * F2_1 * lmbd + F1_1 * F2_2 * miu;
        D[1][4] = D[4][1] = 2 * miu * e_uv + F2_1 * F2_2 * lmbd + F2_1 * F2_2 * miu;
        D[1][5] = D[5][1] = F2_1 * F3_2 * lmbd + F2_2 * F3_1 * miu;
        D[2][2] = F3_1 * (F3_1 * lmbd + 2 * F3_1 * miu) + 2 * miu * e_uu + lmbd * e_sum + F3_2 * F3_2 * miu;
        D[2][3] = D[3][2] = F1_2 * F3_1 * lmbd + F1_1 * F3_2 * miu;
        D[2][4] = D[4][2] = F2_2 * F3_1 * lmbd + F2_1 * F3_2 * miu;
        D[2][5] = D[5][2] = 2 * miu * e_uv + F3_1 * F3_2 * lmbd + F3_1 * F3_2 * miu;
        D[3][3] = F1_2 * (F1_2 * lmbd + 2 * F1_2 * miu) + 2 * miu * e_vv + lmbd * e_sum + F1_1 * F1_1 * miu;
        D[3][4] = D[4][3] = F1_2 * (F2_2 * lmbd + 2 * F2_2 * miu) + F1_1 * F2_1 * miu;
        D[3][5] = D[5][3] = F1_2 * (F3_2 * lmbd + 2 * F3_2 * miu) + F1_1 * F3_1 * miu;
        D[4][4] = F2_2 * (F2_2 * lmbd + 2 * F2_2 * miu) + 2 * miu * e_vv + lmbd * e_sum + F2_1 * F2_1 * miu;
        D[4][5] = D[5][4] = F2_2 * (F3_2 * lmbd + 2 * F3_2 * miu) + F2_1 * F3_1 * miu;
        D[5][5] = F3_2 * (F3_2 * lmbd + 2 * F3_2 * miu) + 2 * miu * e_vv + lmbd * e_sum + F3_1 * F3_1 * miu;
        for (int a = 0; a < 6; a++) for (int b = 0; b < 6; b++) D[a][b] *= A;
        float m1 = 0.f, m2 = 0.f;
        switch (order) {
            case 0: m1 = -Dm_inv1_1 - Dm_inv2_1; m2 = -Dm_inv1_2 - Dm_inv2_2; force += V3(-F12[0] - F12[3], -F12[1] - F12[4], -F12[2] - F12[5]); break;
            case 1: m1 = Dm_inv1_1; m2 = Dm_inv1_2; force += V3(F12[0], F12[1], F12[2]); break;
            case 2: m1 = Dm_inv2_1; m2 = Dm_inv2_2; force += V3(F12[3], F12[4], F12[5]); break;
            default: break;
        }
        float HL[3][6];
        for (int r = 0; r < 3; r++) for (int c = 0; c < 6; c++) HL[r][c] = D[r][c] * m1 + D[r + 3][c] * m2;
        for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++) hessian(r, c) += HL[r][c] * m1 + HL[r][c + 3] * m2;
    }
    // Rayleigh-style damping on the membrane block (:905-910)
    M3 dampingH;
    for (int k = 0; k < 9; k++) dampingH.m[k] = (hessian.m[k] - old.m[k]) * (m.obj.dampingStVK / s.dt);
    const V3 displacement = m.pos[iV] - m.prev[iV];
    force -= mul(dampingH, displacement);
    for (int k = 0; k < 9; k++) hessian.m[k] += dampingH.m[k];
}

// VBDTriMeshStVk::accumlateBendingForceAndHessian (:913-982)
static void bending_force_hessian(const ClothScene& s, const TriMesh& m, int vId, V3& force, M3& hessian) {
    const float kB = m.obj.bendingStiffness, damping = m.obj.dampingBending, dt = s.dt;
    const float thres2 = s.phys.degenerateTriangleThres * s.phys.degenerateTriangleThres;
    for (int i = m.vBendOff[vId]; i < m.vBendOff[vId + 1]; i++) {
        const int eId = m.vBend[i], order = m.vBendOrder[i];
        const ClothEdge& E = m.edges[eId];
        if (E.fId2 == -1) continue;
        const V3 X[4] = {m.pos[E.eV1], m.pos[E.eV2], m.pos[E.eV12Next], m.pos[E.eV21Next]};
        const V3 n1 = cross(X[1] - X[0], X[2] - X[0]), n2 = cross(X[2] - X[3], X[1] - X[3]);
        if (sqnorm(n1) < thres2 || sqnorm(n2) < thres2) continue;
        const float* Q = m.Q[eId].data();
        V3 dE;   // row `order` of kB * Q * Xs
        for (int k = 0; k < 4; k++) dE += (kB * Q[4 * order + k]) * X[k];
        const float h = kB * Q[4 * order + order];
        force -= dE + (h * (m.pos[vId] - m.prev[vId])) * (damping / dt);
        const float hh = h * (1.f + damping / dt);
        hessian(0, 0) += hh; hessian(1, 1) += hh; hessian(2, 2) += hh;
    }
}

void cloth_material_force_hessian(const ClothScene& s, const TriMesh& m, int v, V3& force, M3& hessian) {
    stvk_force_hessian(s, m, v, force, hessian);
    bending_force_hessian(s, m, v, force, hessian);
}

// VBDClothSimulationFramework::computeFriction (M/VBDCloth/VBDClothPhysics.cpp:1500-1531); T = [t0 t1]
static void cloth_friction(float mu, float lambda, V3 t0, V3 t1, float u0, float u1, float epsU, V3& force, M3& hessian) {
    const float uNorm = std::sqrt(u0 * u0 + u1 * u1);
    force = V3(); hessian = M3();
    if (uNorm > 0) {
        float f1;
        if (uNorm > epsU) f1 = 1 / uNorm;
        else f1 = (-uNorm / epsU + 2) / epsU;
        force = (-mu * lambda) * ((f1 * u0) * t0 + (f1 * u1) * t1);
        add_outer(hessian, mu * lambda * f1, t0);
        add_outer(hessian, mu * lambda * f1, t1);
    }
}

// computeContactRepulsiveForce (:1533-1579)
void cloth_contact_repulsion(const ClothScene& s, float dis_, float& dEdD, float& d2E) {
    const float thickness = s.phys.thickness, contactRadius = s.phys.contactRadius, k = s.phys.contactStiffness;
    const float disAdjusted = dis_ - thickness;
    const float penetrationDepth = contactRadius - disAdjusted;
    dEdD = 0.f; d2E = 0.f;
    switch (s.phys.contactEnergyTypes) {
        case 0: dEdD = -k * penetrationDepth; d2E = k; break;
        case 1: {
            const float tau = contactRadius * 0.5f;
            if (disAdjusted < tau && disAdjusted > CMP_EPSILON) {
                const float k2 = 0.5f * (tau * tau) * k;
                dEdD = -k2 / disAdjusted; d2E = k2 / (disAdjusted * disAdjusted);
            } else { dEdD = -k * penetrationDepth; d2E = k; }
        } break;
        case 2: {
            const float dr = disAdjusted - contactRadius;
            dEdD = (-k * (dr * dr) / disAdjusted - 2 * k * std::log(disAdjusted / contactRadius) * dr);
            d2E = k * (((dr * dr) / (disAdjusted * disAdjusted)) - (4 * disAdjusted - 4 * contactRadius) / disAdjusted - 2 * std::log(disAdjusted / contactRadius));
        } break;
        default: break;
    }
}

// accumulateVFContactForceAndHessian (:1248-1321) with updateContactInfo = true (updateVFContactPointInfo,
// M/CollisionDetector/ClothContactDetector.cpp:1197-1228) and apply_friction = true
static float vf_contact_force_hessian(const ClothScene& s, const VFContact& c0, int vertexOrder, V3& force, M3& hessian) {
    VFContact c = c0;   // the reference refreshes the stored contact in place (same values from every caller); a local copy here
    const TriMesh& mv = s.meshes[c.vMesh];
    const TriMesh& mf = s.meshes[c.fMesh];
    const int t0 = mf.faces[3 * c.f], t1 = mf.faces[3 * c.f + 1], t2 = mf.faces[3 * c.f + 2];
    const V3 x = mv.pos[c.v];
    c.cp = closest_point_triangle(x, mf.pos[t0], mf.pos[t1], mf.pos[t2], c.bary, c.closestType);
    c.d = norm(x - c.cp);
    const V3 diff = x - c.cp;
    const V3 n = normalized(diff);
    const float dis = dot(diff, n);
    if (dis < s.phys.contactRadius + s.phys.thickness) {
        const float bs[4] = {-c.bary.x, -c.bary.y, -c.bary.z, 1.f};
        const float b = bs[vertexOrder];
        float dEdD, d2E; cloth_contact_repulsion(s, dis, dEdD, d2E);
        const float lambda = -dEdD;
        force += (b * lambda) * n;
        add_outer(hessian, d2E * b * b, n);
        // friction
        const V3 dx3 = x - mv.prev[c.v];
        const V3 cpPrev = c.bary.x * mf.prev[t0] + c.bary.y * mf.prev[t1] + c.bary.z * mf.prev[t2];
        const V3 dx = dx3 - (c.cp - cpPrev);
        const V3 e0 = normalized(mf.pos[t1] - mf.pos[t0]);
        const V3 e1 = normalized(cross(e0, n));
        const float mu = (mf.obj.frictionDynamic + mv.obj.frictionDynamic) * 0.5f;
        const float epsV = (mf.obj.frictionEpsV + mv.obj.frictionEpsV) * 0.5f;
        V3 ff; M3 fh;
        cloth_friction(mu, lambda, e0, e1, dot(e0, dx), dot(e1, dx), epsV * s.dt, ff, fh);
        force += b * ff;
        for (int k = 0; k < 9; k++) hessian.m[k] += b * b * fh.m[k];
    }
    return dis;
}

// accumulateEEContactForceAndHessian (:1323-1399) with updateEEContactPointInfo (ClothContactDetector.cpp:1230-1251)
static float ee_contact_force_hessian(const ClothScene& s, const EEContact& c0, int vertexOrder, V3& force, M3& hessian) {
    EEContact c = c0;
    const TriMesh& m1 = s.meshes[c.m1];
    const TriMesh& m2 = s.meshes[c.m2];
    const ClothEdge& E1 = m1.edges[c.e1];
    const ClothEdge& E2 = m2.edges[c.e2];
    closest_points_segments(m1.pos[E1.eV1], m1.pos[E1.eV2], m2.pos[E2.eV1], m2.pos[E2.eV2], c.c1, c.c2, c.mu1, c.mu2);
    const V3 diff = c.c1 - c.c2;
    const float dis = norm(diff);
    const V3 v1 = m1.pos[E1.eV2] - m1.pos[E1.eV1], v2 = m2.pos[E2.eV2] - m2.pos[E2.eV1];
    const float parallelEps = sqnorm(cross(v1, v2)) / (norm(v1) * norm(v2));
    if (dis < s.phys.contactRadius + s.phys.thickness && parallelEps > CMP_EPSILON) {
        const V3 n = diff / dis;
        const float bs[4] = {1.f - c.mu1, c.mu1, -1.f + c.mu2, -c.mu2};
        const float b = bs[vertexOrder];
        float dEdD, d2E; cloth_contact_repulsion(s, dis, dEdD, d2E);
        const float lambda = -dEdD;
        force += (b * lambda) * n;
        add_outer(hessian, d2E * b * b, n);
        const V3 c1Prev = (1.f - c.mu1) * m1.prev[E1.eV1] + c.mu1 * m1.prev[E1.eV2];
        const V3 c2Prev = (1.f - c.mu2) * m2.prev[E2.eV1] + c.mu2 * m2.prev[E2.eV2];
        const V3 dx = (c.c1 - c1Prev) - (c.c2 - c2Prev);
        const V3 t0 = normalized(v1);
        const V3 t1 = normalized(cross(t0, n));
        const float mu = (m1.obj.frictionDynamic + m2.obj.frictionDynamic) * 0.5f;
        const float epsV = (m1.obj.frictionEpsV + m2.obj.frictionEpsV) * 0.5f;
        V3 ff; M3 fh;
        cloth_friction(mu, lambda, t0, t1, dot(t0, dx), dot(t1, dx), epsV * s.dt, ff, fh);
        force += b * ff;
        for (int k = 0; k < 9; k++) hessian.m[k] += b * b * fh.m[k];
    }
    return dis;
}

// accumlateBoundaryForceAndHessian (:1401-1463)
static void boundary_force_hessian(const ClothScene& s, const TriMesh& m, int vId, V3& force, M3& hessian) {
    const float kb = s.phys.boundaryContactStiffness, mu = s.phys.boundaryFrictionDynamic, epsV = s.phys.boundaryFrictionEpsValue;
    for (int d = 0; d < 3; d++) {
        const float lower = s.phys.worldBoundsMin[d], upper = s.phys.worldBoundsMax[d];
        const float xd = m.pos[vId][d];
        float depth; float sign;
        if (xd < lower) { depth = lower - xd; sign = 1.f; }
        else if (xd > upper) { depth = xd - upper; sign = -1.f; }
        else continue;
        force[d] += sign * (depth * kb);
        hessian(d, d) += kb;
        const V3 dx = m.pos[vId] - m.prev[vId];
        V3 t0, t1; t0[(d + 1) % 3] = 1.f; t1[(d + 2) % 3] = 1.f;
        V3 ff; M3 fh;
        cloth_friction(mu, depth * kb, t0, t1, dot(t0, dx), dot(t1, dx), epsV * s.dt, ff, fh);
        force += ff;
        for (int k = 0; k < 9; k++) hessian.m[k] += fh.m[k];
    }
}

// parity probes (tests/test_oracle_vs_ref.py): one contact / the world box of one vertex in isolation
float cloth_probe_vf(const ClothScene& s, int vMesh, int v, int fMesh, int f, int vertexOrder, V3& force, M3& hessian) {
    VFContact c{}; c.vMesh = vMesh; c.v = v; c.fMesh = fMesh; c.f = f;
    force = V3(); hessian = M3();
    return vf_contact_force_hessian(s, c, vertexOrder, force, hessian);
}
float cloth_probe_ee(const ClothScene& s, int m1, int e1, int m2, int e2, int vertexOrder, V3& force, M3& hessian) {
    EEContact c{}; c.m1 = m1; c.e1 = e1; c.m2 = m2; c.e2 = e2;
    force = V3(); hessian = M3();
    return ee_contact_force_hessian(s, c, vertexOrder, force, hessian);
}
void cloth_probe_boundary(const ClothScene& s, int mesh, int v, V3& force, M3& hessian) {
    force = V3(); hessian = M3();
    boundary_force_hessian(s, s.meshes[mesh], v, force, hessian);
}

// evaluateVertexForceAndHessian / the first part of VBDStepWithExistingCollisions (:1026-1104, 1156-1221)
void cloth_vertex_force_hessian(ClothScene& s, int iMesh, int vId, V3& force, M3& hessian) {
    TriMesh& m = s.meshes[iMesh];
    force = V3(); hessian = M3();
    for (auto& c : m.vf[vId]) vf_contact_force_hessian(s, c, 3, force, hessian);
    for (int i = m.vNbrFOff[vId]; i < m.vNbrFOff[vId + 1]; i++) {
        const int fId = m.vNbrF[i];
        for (auto& link : m.faceLinks[fId]) vf_contact_force_hessian(s, s.meshes[link.vMesh].vf[link.v][link.contactId], m.vNbrFOrder[i], force, hessian);
    }
    for (int i = m.vNbrEOff[vId]; i < m.vNbrEOff[vId + 1]; i++)
        for (auto& c : m.ee[m.vNbrE[i]]) ee_contact_force_hessian(s, c, m.vNbrEOrder[i], force, hessian);
    // accumlateInertiaForceAndHessian (M/VBDCloth/VBDBaseTriMesh.cpp:211-218)
    force += m.mass[vId] * (m.inertia[vId] - m.pos[vId]) * s.dtSqrReciprocal;
    const float hi = m.mass[vId] * s.dtSqrReciprocal;
    hessian(0, 0) += hi; hessian(1, 1) += hi; hessian(2, 2) += hi;
    cloth_material_force_hessian(s, m, vId, force, hessian);
    if (!m.sewingVertices.empty()) cloth_sewing_force_hessian(m, vId, force, hessian);      // :1112-1114
    boundary_force_hessian(s, m, vId, force, hessian);
    cloth_constraint_force_hessian(s, iMesh, vId, force, hessian);                           // :1118
    // (external forces are not part of the path: vertexExternalForces is zero, M/VBDCloth/VBDBaseTriMesh.cpp:185-188)
}

// VBDBaseTriMesh::accumlateSewingForceAndHessian (M/VBDCloth/VBDBaseTriMesh.cpp:320-344)
void cloth_sewing_force_hessian(const TriMesh& m, int iV, V3& force, M3& hessian) {
    if (m.sewingVMap.empty()) return;
    const float k = m.sewingStiffness;
    for (const auto& p : m.sewingVMap[iV]) {
        const int sid = p.first, order = p.second;
        float w = 1.f;
        const float lbda = m.sewingRatio[sid];
        if (order == 1) w = -(1 - lbda);
        else if (order == 2) w = -lbda;
        const V3 x0 = m.pos[m.sewingVertices[sid][0]], x1 = m.pos[m.sewingVertices[sid][1]], x2 = m.pos[m.sewingVertices[sid][2]];
        const V3 x = x0 - (1 - lbda) * x1 - lbda * x2;
        force -= (w * k) * x;
        const float h = w * w * k;
        hessian(0, 0) += h; hessian(1, 1) += h; hessian(2, 2) += h;
    }
}

// VBDClothSimulationFramework::accumulateConstraintForceAndHessian (M/VBDCloth/VBDClothPhysics.cpp:1464-1498). Note :1472: the
// position is read from the V-SIDE mesh with the id of the vertex being solved, also for the face-side entries.
void cloth_constraint_force_hessian(const ClothScene& s, int iMesh, int vId, V3& force, M3& hessian) {
    const TriMesh& m = s.meshes[iMesh];
    if (m.attachments.empty()) return;
    for (const Attachment& a : m.attachments[vId]) {
        const V3 vertPos = s.meshes[a.meshV].pos[vId];
        const float bs[4] = {-a.attachedPos[0], -a.attachedPos[1], -a.attachedPos[2], 1.f};
        V3 attachedPos;
        if (a.meshF != -1) {
            const TriMesh& t = s.meshes[a.meshF];
            attachedPos = a.attachedPos[0] * t.pos[a.face[0]] + a.attachedPos[1] * t.pos[a.face[1]] + a.attachedPos[2] * t.pos[a.face[2]];
        } else attachedPos = a.attachedPos;
        const V3 diff = vertPos - attachedPos;
        const float b = bs[a.vertexOrder];
        force -= (a.stiffness * b) * diff;
        const float h = a.stiffness * b * b;
        hessian(0, 0) += h; hessian(1, 1) += h; hessian(2, 2) += h;
    }
}

// TriMeshFEM::generate_sewing_filters, the sewingVMap part (M/TriMesh/TriMesh.h:596-633)
void cloth_set_sewing(ClothScene& s, int mesh, int nSeams, const int32_t* verts3, const float* ratios, float stiffness) {
    TriMesh& m = s.meshes[mesh];
    m.sewingVertices.clear(); m.sewingRatio.assign(ratios, ratios + nSeams); m.sewingStiffness = stiffness;
    m.sewingVMap.assign(m.nV, {});
    for (int i = 0; i < nSeams; i++) {
        m.sewingVertices.push_back({verts3[3 * i], verts3[3 * i + 1], verts3[3 * i + 2]});
        for (int o = 0; o < 3; o++) m.sewingVMap[verts3[3 * i + o]].push_back({i, o});
    }
}

// TriMeshFEM::generate_sewing_filters, the filteredPairsPath part (M/TriMesh/TriMesh.h:596-622): {"vf": [[v, f], ...], "ee": [[e, e'], ...]};
// the vertex-face pairs go one way, the edge pairs both ways
void cloth_set_filtered_pairs(ClothScene& s, int mesh, int nVF, const int32_t* vfPairs, int nEE, const int32_t* eePairs) {
    TriMesh& m = s.meshes[mesh];
    m.sewingVF.assign(m.nV, {}); m.sewingEE.assign(m.nE, {});
    for (int i = 0; i < nVF; i++) m.sewingVF[vfPairs[2 * i]].push_back(vfPairs[2 * i + 1]);
    for (int i = 0; i < nEE; i++) { m.sewingEE[eePairs[2 * i]].push_back(eePairs[2 * i + 1]); m.sewingEE[eePairs[2 * i + 1]].push_back(eePairs[2 * i]); }
    for (auto& l : m.sewingVF) { std::sort(l.begin(), l.end()); l.erase(std::unique(l.begin(), l.end()), l.end()); }
    for (auto& l : m.sewingEE) { std::sort(l.begin(), l.end()); l.erase(std::unique(l.begin(), l.end()), l.end()); }
}

// the set-up loop of VBDClothSimulationFramework::initialize (M/VBDCloth/VBDClothPhysics.cpp:161-219)
int cloth_add_attachment(ClothScene& s, int meshV, int vertex, int meshF, const int32_t faceVerts[3], const float baryOrPos[3], float stiffness) {
    TriMesh& mv = s.meshes[meshV];
    if (mv.attachments.empty()) mv.attachments.assign(mv.nV, {});
    Attachment a;
    a.vertexOrder = 3; a.meshF = meshF; a.meshV = meshV; a.vertexId = vertex; a.stiffness = stiffness;
    for (int k = 0; k < 3; k++) a.face[k] = meshF >= 0 ? faceVerts[k] : -1;
    a.attachedPos = V3(baryOrPos[0], baryOrPos[1], baryOrPos[2]);
    mv.attachments[vertex].push_back(a);
    if (meshF >= 0 && s.meshes[meshF].simulated) {      // (simulated meshes come first: meshF < numSimulationMeshes())
        TriMesh& mf = s.meshes[meshF];
        if (mf.attachments.empty()) mf.attachments.assign(mf.nV, {});
        for (int iFV = 0; iFV < 3; iFV++) {
            if (a.attachedPos[iFV] == 0.f) continue;
            Attachment f = a;
            f.vertexOrder = iFV;
            mf.attachments[faceVerts[iFV]].push_back(f);
        }
    }
    return GAIA_OK;
}

// applyConservativeBoundTruncation (:1223-1246)
static bool truncate_to_bound(const TriMesh& m, int vId, V3& newPos) {
    V3 acc = newPos - m.posAtCD[vId];
    const float size = norm(acc);
    const float bound = m.bounds[vId];
    if (size > bound) {
        acc = acc * (bound / size);
        newPos = m.posAtCD[vId] + acc;
        return true;
    }
    return false;
}

// VBDBaseTriMesh::applyInitialStep + forwardStep_intertia (M/VBDCloth/VBDBaseTriMesh.cpp:60-182)
static void apply_initial_step(ClothScene& s, TriMesh& m) {
    const float dt = s.dt;
    m.prev = m.pos;
    if (m.hasVelocitiesPrev) {
        for (int i = 0; i < m.nV; i++) m.accel[i] = (m.vel[i] - m.velPrev[i]) / dt;
        m.hasApproxAcceleration = true;
    } else m.hasVelocitiesPrev = true;
    m.velPrev = m.vel;
    const V3 g(s.phys.gravity[0], s.phys.gravity[1], s.phys.gravity[2]);
    // external forces are zero: velocities += dt * 0 * invMass leaves them unchanged
    if (s.phys.associateGravityWithInertia) for (auto& v : m.vel) v += dt * g;
    for (int fp : m.fixedPoints) m.vel[fp] = V3();
    for (int i = 0; i < m.nV; i++) m.inertia[i] = m.pos[i] + dt * m.vel[i];
    const float gravNorm = norm(g);
    const V3 gravDir = g / gravNorm;
    switch (m.obj.initializationType) {
        case 0: m.pos = m.inertia; break;
        case 1:
            for (int i = 0; i < m.nV; i++) {
                if (m.hasApproxAcceleration) {
                    float ac = dot(m.accel[i], gravDir);
                    ac = ac < gravNorm ? ac : gravNorm;
                    ac = ac > 1e-5f ? ac : 0.f;
                    m.pos[i] = m.prev[i] + dt * m.velPrev[i] + ((dt * dt) * gravDir) * ac;
                } else m.pos[i] = m.inertia[i];
            }
            break;
        case 2: m.pos = m.prev; break;
        case 3:
            for (int i = 0; i < m.nV; i++) {
                const float bound = m.bounds[i];
                V3 step = m.inertia[i] - m.prev[i];
                const float size = norm(step);
                if (s.phys.handleCollision && size > bound && size) step = step * (bound / size);
                m.pos[i] = m.prev[i] + step;
            }
            break;
        default: break;
    }
}

// ChebysevAccelerator::getAcceleratorOmega (same recurrence as M/VBD/VBDPhysics.cpp:3558-3574)
float cloth_accelerator_omega(float rho, int order, float prevOmega) {      // ChebysevAccelerator::getAcceleratorOmega (M/SolverUtils/ChebysevAccelerator.h:15-35)
    switch (order) {
        case 1: return 1.f;
        case 2: return 2.f / (2.f - rho * rho);
        default: return 4.f / (4.f - rho * rho * prevOmega);
    }
}

// one iteration of runStep_CDEveryIter (:399-509)
// setFixedPointCloth (M/VBDCloth/VBDClothDeformer.h:5-21)
static void set_fixed_points(TriMesh& m, const std::vector<int>& verts) {
    for (int v : verts) {
        m.fixedMask[v] = 1;
        if (std::find(m.fixedPoints.begin(), m.fixedPoints.end(), v) == m.fixedPoints.end()) m.fixedPoints.push_back(v);
    }
}

// ClothDeformerRotator constructor (:44-69)
void cloth_add_rotator(ClothScene& s, int mesh, const int32_t* verts, int n, const float center[3], const float axis[3],
                       double angularVelocity, double rotationEnd) {
    ClothDeformer d;
    d.type = 0; d.mesh = mesh; d.verts.assign(verts, verts + n);
    d.o = V3(center[0], center[1], center[2]); d.axis = V3(axis[0], axis[1], axis[2]);
    d.angularVelocity = angularVelocity; d.rotationEnd = rotationEnd;
    TriMesh& m = s.meshes[mesh];
    set_fixed_points(m, d.verts);
    for (int v : d.verts) {
        const V3 P = m.pos[v];
        const V3 OP = P - d.o;
        const V3 root = dot(OP, d.axis) * d.axis;
        d.roots.push_back(root);
        d.rootPs.push_back(P - root);
    }
    s.deformers.push_back(d);
}

// ClothDeformerTranslater constructor (:135-149)
void cloth_add_translator(ClothScene& s, int mesh, const int32_t* verts, int n, const float speed[3], float tStart, float tEnd) {
    ClothDeformer d;
    d.type = 1; d.mesh = mesh; d.verts.assign(verts, verts + n);
    d.speed = V3(speed[0], speed[1], speed[2]);
    d.deformationStartTime = tStart; d.deformationEndTime = tEnd;
    set_fixed_points(s.meshes[mesh], d.verts);
    s.deformers.push_back(d);
}

// VBDClothSimulationFramework::applyDeformers (M/VBDCloth/VBDClothPhysics.cpp:2464-2470) with the two operator()s
// (VBDClothDeformer.h:71-117, 151-191)
static void apply_deformers(ClothScene& s) {
    for (const ClothDeformer& d : s.deformers) {
        TriMesh& m = s.meshes[d.mesh];
        if (d.type == 0) {
            float curTime = s.curTime;
            if (curTime >= d.rotationEnd && d.rotationEnd > 0) curTime = (float)d.rotationEnd;
            const double theta = curTime * d.angularVelocity;
            const double ux = d.axis.x, uy = d.axis.y, uz = d.axis.z;
            const float R[9] = {(float)(cos(theta) + ux * ux * (1 - cos(theta))), (float)(ux * uy * (1 - cos(theta)) - uz * sin(theta)), (float)(ux * uz * (1 - cos(theta)) + uy * sin(theta)),
                                (float)(uy * ux * (1 - cos(theta)) + uz * sin(theta)), (float)(cos(theta) + uy * uy * (1 - cos(theta))), (float)(uy * uz * (1 - cos(theta)) - ux * sin(theta)),
                                (float)(uz * ux * (1 - cos(theta)) - uy * sin(theta)), (float)(uz * uy * (1 - cos(theta)) + ux * sin(theta)), (float)(cos(theta) + uz * uz * (1 - cos(theta)))};
            for (size_t i = 0; i < d.verts.size(); i++) {
                const V3& q = d.rootPs[i];
                const V3 rot((R[0] * q.x + R[1] * q.y) + R[2] * q.z, (R[3] * q.x + R[4] * q.y) + R[5] * q.z, (R[6] * q.x + R[7] * q.y) + R[8] * q.z);
                m.pos[d.verts[i]] = d.roots[i] + rot;
            }
        } else if (s.curTime >= d.deformationStartTime && (s.curTime < d.deformationEndTime || d.deformationEndTime < 0)) {
            for (int v : d.verts) m.pos[v] += d.speed * s.dt;
        }
    }
}

// DynamicCollider::updateColliderMeshes -> ColliderTrimeshSequence::update / set_state
// (M/SpatialQuery/ColiiderTriMeshSequence.h:59-129); true when a collider moved
static bool update_colliders(ClothScene& s) {
    bool any = false;
    const int numSubsteps = s.phys.numSubsteps;
    for (ColliderSequence& c : s.colliderSeqs) {
        TriMesh& m = s.meshes[c.mesh];
        const int frameId = s.frameId;
        const std::vector<int>& kf = c.keyFrames;
        auto lerp = [&](size_t k0, size_t k1, float t) {
            const float w = 1 - t;
            for (int v = 0; v < m.nV; v++) m.pos[v] = c.keyPos[k0][v] * w + c.keyPos[k1][v] * t;
        };
        bool updated = false;
        if (c.firstTime) {
            c.firstTime = false;
            if (frameId < kf.front()) lerp(0, 0, 0.f);
            else if (frameId >= kf.back()) lerp(kf.size() - 1, kf.size() - 1, 0.f);
            else {
                const size_t hi = std::upper_bound(kf.begin(), kf.end(), frameId) - kf.begin();
                lerp(hi - 1, hi, float(frameId - kf[hi - 1]) / (kf[hi] - kf[hi - 1]));
            }
            updated = true;
        } else if (!(frameId < kf.front() || frameId >= kf.back())) {
            const size_t hi = std::upper_bound(kf.begin(), kf.end(), frameId) - kf.begin();
            const int prevF = kf[hi - 1], numFrames = kf[hi] - prevF;
            if (c.interpolateSubstep) {
                if (s.iIter < c.interpolateIter) {
                    float t = float(frameId - prevF) / numFrames;
                    t += float(s.substep) / ((size_t)numSubsteps * numFrames);
                    t += float(s.iIter) / ((size_t)c.interpolateIter * numSubsteps * numFrames);
                    lerp(hi - 1, hi, t);
                    updated = true;
                }
            } else if (s.substep == 0 && s.iIter == 0) {
                lerp(hi - 1, hi, float(frameId - prevF) / numFrames);
                updated = true;
            }
        }
        any = any || updated;
    }
    return any;
}

bool cloth_probe_update_colliders(ClothScene& s, int frameId, int substep, int iIter) {      // parity probe: one call of updateCollider at the given place of the step
    s.frameId = frameId; s.substep = substep; s.iIter = iIter;
    return update_colliders(s);
}

static void cloth_iteration(ClothScene& s, int iIter, float& omega) {
    s.iIter = iIter;
    if (iIter == 0) apply_deformers(s);
    if (iIter == 0) for (int i = 0; i < s.numSim; i++) std::fill(s.meshes[i].activeCollisionMask.begin(), s.meshes[i].activeCollisionMask.end(), 0);
    if (!s.colliderSeqs.empty() && update_colliders(s)) s.collisionDetectionRequired = true;     // VBDClothPhysics.cpp:413-418
    if (s.collisionDetectionRequired || iIter == 0) {
        if (s.phys.handleCollision) cloth_detect(s);
        else for (int i = 0; i < s.numSim; i++) std::fill(s.meshes[i].bounds.begin(), s.meshes[i].bounds.end(), std::numeric_limits<float>::max());
    }
    if (s.phys.applyAcceleration) for (int i = 0; i < s.numSim; i++) s.meshes[i].prevIter = s.meshes[i].pos;
    if (iIter == 0) for (int i = 0; i < s.numSim; i++) apply_initial_step(s, s.meshes[i]);
    std::atomic<int> truncated(0);
    for (auto& group : s.groups) {
        const int64_t n = (int64_t)group.size() / 2;
        s.pool->parallel_for(n, [&](int64_t b, int64_t e) {
            for (int64_t i = b; i < e; i++) {
                const int iMesh = group[2 * i], vId = group[2 * i + 1];
                TriMesh& m = s.meshes[iMesh];
                if (m.fixedMask[vId]) continue;
                // VBDStepWithExistingCollisions (:1026-1154)
                V3 force; M3 hessian;
                cloth_vertex_force_hessian(s, iMesh, vId, force, hessian);
                float dx[3];
                const float f3[3] = {force.x, force.y, force.z};
                if (solve3x3_psd_stable(hessian.m, f3, dx)) {
                    V3 posNew = m.pos[vId] + s.phys.stepSize * V3(dx[0], dx[1], dx[2]);
                    if (truncate_to_bound(m, vId, posNew)) truncated = 1;
                    m.next[vId] = posNew;
                }   // "Solver failed": the vertex keeps vertexPositionNext as it was
            }
        }, 64);
        for (int64_t i = 0; i < n; i++) {
            TriMesh& m = s.meshes[group[2 * i]];
            const int vId = group[2 * i + 1];
            if (!m.fixedMask[vId]) m.pos[vId] = m.next[vId];
        }
        s.stats.colourSweeps++;
    }
    for (int i = 0; i < s.numSim; i++) { const TriMesh& m = s.meshes[i]; for (int v = 0; v < m.nV; v++) if (!m.fixedMask[v]) s.stats.vertexSolves++; }
    if (truncated) s.collisionDetectionRequired = true;     // rule R6
    if (s.phys.applyAcceleration) {
        omega = cloth_accelerator_omega(s.phys.acceleratorRho, iIter + 1, omega);
        // applyAcceleration (:893-915)
        if (omega > 1.f)
            for (int i = 0; i < s.numSim; i++) {
                TriMesh& m = s.meshes[i];
                for (int v = 0; v < m.nV; v++) {
                    if (m.activeCollisionMask[v]) continue;
                    V3 posNew = omega * (m.pos[v] - m.prevIter[v]) + m.prevIter[v];
                    if (truncate_to_bound(m, v, posNew)) s.collisionDetectionRequired = true;
                    m.pos[v] = posNew;
                }
            }
    }
}

// updateVelocity (:2318-2359)
static void update_velocity(ClothScene& s) {
    for (int i = 0; i < s.numSim; i++) {
        TriMesh& m = s.meshes[i];
        for (int v = 0; v < m.nV; v++) {
            V3 vel = (m.pos[v] - m.prev[v]) / s.dt;
            const float vMag = norm(vel);
            const float maxV = m.obj.maxVelocityMagnitude;
            if (m.obj.hasNoGravZone && maxV > 0 && vel[GRAVITY_AXIS] > m.obj.noGravZoneThreshold) {
                if (vMag > maxV) vel = vel * (maxV / vMag);
            } else if (vMag > 1e-6f) {
                float vNew = vMag * m.obj.exponentialVelDamping - m.obj.constantVelDamping;
                vNew = vNew > 1e-6f ? vNew : 0.f;
                vel = vel * (vNew / vMag);
            } else vel = vel * 0.f;
            m.vel[v] = vel;
        }
        for (int fp : m.fixedPoints) m.vel[fp] = V3();
    }
}

// parity probes (tests/test_oracle_vs_ref.py): the pieces of a substep around the solver, one at a time
void cloth_probe_initial_step(ClothScene& s) { for (int i = 0; i < s.numSim; i++) apply_initial_step(s, s.meshes[i]); }
void cloth_probe_update_velocity(ClothScene& s) { update_velocity(s); }
bool cloth_probe_truncate(const ClothScene& s, int mesh, int v, V3& newPos) { return truncate_to_bound(s.meshes[mesh], v, newPos); }

void cloth_run_substeps(ClothScene& s, int nSubsteps) {
    for (s.substep = 0; s.substep < nSubsteps; s.substep++) {
        s.curTime += s.dt;
        float omega = 1.f;
        for (int it = 0; it < s.phys.iterations; it++) cloth_iteration(s, it, omega);
        update_velocity(s);
        s.stats.substeps++;
    }
    s.stats.frames++;
    s.frameId++;
}

void cloth_run_iterations(ClothScene& s, int nIters) {
    s.substep = 0;
    s.curTime += s.dt;
    float omega = 1.f;
    for (int it = 0; it < nIters; it++) cloth_iteration(s, it, omega);
}

// merit-energy pieces for the finite-difference tests (evaluateVertexMeritEnergy, M/VBDCloth/VBDTriMeshStVK.cpp:690-712;
// bending energy 0.5 * kB * X^T Q X per hinge)
float cloth_elastic_energy(const ClothScene& s, const TriMesh& m) {
    double E = 0;
    for (int f = 0; f < m.nF; f++) {
        float F[6]; deformation_gradient(m, f, F);
        const double uu = 0.5 * ((double)F[0] * F[0] + (double)F[1] * F[1] + (double)F[2] * F[2] - 1.0), vv = 0.5 * ((double)F[3] * F[3] + (double)F[4] * F[4] + (double)F[5] * F[5] - 1.0);
        const double uv = 0.5 * ((double)F[0] * F[3] + (double)F[1] * F[4] + (double)F[2] * F[5]);
        E += m.restArea[f] * (0.5 * m.obj.lambda * (uu + vv) * (uu + vv) + m.obj.miu * (uu * uu + vv * vv + 2 * uv * uv));
    }
    for (int e = 0; e < m.nE; e++) {
        const ClothEdge& Ed = m.edges[e];
        if (Ed.fId2 == -1) continue;
        const V3 X[4] = {m.pos[Ed.eV1], m.pos[Ed.eV2], m.pos[Ed.eV12Next], m.pos[Ed.eV21Next]};
        for (int d = 0; d < 3; d++) {
            double q = 0;
            for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) q += (double)X[r][d] * m.Q[e][4 * r + c] * X[c][d];
            E += 0.5 * m.obj.bendingStiffness * q;
        }
    }
    (void)s;
    return (float)E;
}

}  // namespace orc
