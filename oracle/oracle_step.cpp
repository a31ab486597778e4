// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// The VBD step, restating the reference CPU/TBB path:
//   VBDPhysics::runStep_hybridCollisionHandling   M/VBD/VBDPhysics.cpp:1634-1707
//   VBDBaseTetMesh::applyInitialStep/computeInertia/forwardStep
//                                                 M/VBD/VBD_BaseMaterial.h:201-259, VBD_BaseMaterial.cpp:7-132
//   VBDTetMeshNeoHookean::accumlateMaterialForceAndHessian2 + assembleVertexVForceAndHessian
//                                                 M/VBD/VBD_NeoHookean.cpp:379-581, 126-152
//   accumlateInertiaForceAndHessian               M/VBD/VBD_BaseMaterial.h:359-366
//   accumlateBoundaryForceAndHessian / computeVertexFriction
//                                                 M/VBD/VBDPhysics.cpp:2821-2954, 2563-2592
//   accumlateCollisionForceAndHessian[PerCollision]  M/VBD/VBDPhysics.cpp:2956-3047
//   VBDStepWithCollision                          M/VBD/VBDPhysics.cpp:2402-2483
//   CuMatrix::solve3x3_psd_stable                 3P/CuMatrix/CuMatrix/MatrixOps/CuMatrix.h:235-267
//   updateVelocities / dampVelocities             M/VBD/VBDPhysics.cpp:2486-2561
//   enableModels (VBD override)                   M/VBD/VBDPhysics.cpp:685-709
//
// Determinism note: the reference updates vertex(v) in place inside a colour (VBDPhysics.cpp:2480). Collision
// force/Hessian of a colour are computed BEFORE the colour's vertex sweep (:1675-1680) and cached, elastic terms
// only read vertices of other colours, so the in-place sweep is order-independent; the oracle keeps it in place.
#include "vbd_oracle.h"

namespace orc {

static const float CMP_EPSILON = 1e-5f;    // M/Types/Types.h:20
static const float CMP_EPSILON2 = 1e-10f;  // M/Types/Types.h:21
static const int GRAVITY_AXIS = 1;         // M/Types/Types.h:9

bool solve3x3_psd_stable(const float* m, const float* b, float* out) {
    const float a11 = m[0], a12 = m[3], a13 = m[6];
    const float a21 = m[1], a22 = m[4], a23 = m[7];
    const float a31 = m[2], a32 = m[5], a33 = m[8];
    const float i11 = a33 * a22 - a32 * a23;
    const float i12 = -(a33 * a12 - a32 * a13);
    const float i13 = a23 * a12 - a22 * a13;
    const float det = (a11 * i11 + a21 * i12 + a31 * i13);
    if (std::fabs(det) < CMP_EPSILON * (std::fabs(a11 * i11) + std::fabs(a21 * i12) + std::fabs(a31 * i13))) {
        out[0] = b[0]; out[1] = b[1]; out[2] = b[2];
        return false;
    }
    const float deti = (float)(1.0 / det);   // `1.0 / det` with DType=float: double divide, stored to float
    const float i21 = -(a33 * a21 - a31 * a23);
    const float i22 = a33 * a11 - a31 * a13;
    const float i23 = -(a23 * a11 - a21 * a13);
    const float i31 = a32 * a21 - a31 * a22;
    const float i32 = -(a32 * a11 - a31 * a12);
    const float i33 = a22 * a11 - a21 * a12;
    out[0] = deti * (i11 * b[0] + i12 * b[1] + i13 * b[2]);
    out[1] = deti * (i21 * b[0] + i22 * b[1] + i23 * b[2]);
    out[2] = deti * (i31 * b[0] + i32 * b[1] + i33 * b[2]);
    return true;
}

// computeVertexFriction (VBDPhysics.cpp:2563-2592). T is 3x2 column-major.
void vertex_friction(float mu, float lambda, const float T[6], const float u[2], float epsU, V3& force, M3& hessian) {
    const float uNorm = std::sqrt(u[0] * u[0] + u[1] * u[1]);
    if (uNorm > 0) {
        float f1_SF_over_x;
        if (uNorm > epsU) f1_SF_over_x = 1 / uNorm;
        else f1_SF_over_x = (-uNorm / epsU + 2) / epsU;
        // force = -mu * lambda * T * f1_SF_over_x * u   (Eigen: ((-mu*lambda) * T) * f1 ... evaluated as scalar*(T*u))
        const float s = -mu * lambda * f1_SF_over_x;
        force = V3(s * (T[0] * u[0] + T[3] * u[1]), s * (T[1] * u[0] + T[4] * u[1]), s * (T[2] * u[0] + T[5] * u[1]));
        const float k = mu * lambda * f1_SF_over_x;
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++)
            hessian(r, c) = k * (T[r] * T[c] + T[3 + r] * T[3 + c]);
    } else {
        force = V3();
        hessian = M3();
    }
}

// VBDTetMeshNeoHookean::accumlateMaterialForce (M/VBD/VBD_NeoHookean.cpp:781-870): the elastic force alone, no damping; the same
// F, det F, cofactors and vertex weights as the force half of accumlateMaterialForceAndHessian below
V3 neohookean_force(const Mesh& m, int iV) {
    const int start = m.vNbrInfos[2 * iV], num = m.vNbrInfos[2 * iV + 1];
    const float miu = m.obj.miu, lmbd = m.obj.lmbd;
    const float a = 1 + miu / lmbd;
    V3 force;
    for (int iNei = 0; iNei < num; iNei++) {
        const int tetId = m.vNbrTets[start + iNei];
        const float A = m.restVol[tetId];
        const float* Dm = &m.DmInv[9 * (size_t)tetId];
        const int* tv = &m.tets[4 * (size_t)tetId];
        float Ds[9];
        const V3 v0 = m.pos[tv[0]];
        for (int c = 0; c < 3; c++) { const V3 d = m.pos[tv[c + 1]] - v0; Ds[3 * c] = d.x; Ds[3 * c + 1] = d.y; Ds[3 * c + 2] = d.z; }
        float F[9];
        for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++)
            F[r + 3 * c] = Ds[r] * Dm[3 * c] + Ds[r + 3] * Dm[3 * c + 1] + Ds[r + 6] * Dm[3 * c + 2];
        const float F1_1 = F[0], F2_1 = F[1], F3_1 = F[2], F1_2 = F[3], F2_2 = F[4], F3_2 = F[5], F1_3 = F[6], F2_3 = F[7], F3_3 = F[8];
        const float detF = F1_1 * (F2_2 * F3_3 - F2_3 * F3_2) - F1_2 * (F2_1 * F3_3 - F2_3 * F3_1) + F1_3 * (F2_1 * F3_2 - F2_2 * F3_1);
        const float dd[9] = {F2_2 * F3_3 - F2_3 * F3_2, F1_3 * F3_2 - F1_2 * F3_3, F1_2 * F2_3 - F1_3 * F2_2,
                             F2_3 * F3_1 - F2_1 * F3_3, F1_1 * F3_3 - F1_3 * F3_1, F1_3 * F2_1 - F1_1 * F2_3,
                             F2_1 * F3_2 - F2_2 * F3_1, F1_2 * F3_1 - F1_1 * F3_2, F1_1 * F2_2 - F1_2 * F2_1};
        const float lk = lmbd * (detF - a);
        float dE_dF[9];
        for (int q = 0; q < 9; q++) dE_dF[q] = A * (miu * F[q] + lk * dd[q]);
        float m1, m2, m3;
        switch (m.vNbrOrder[start + iNei]) {
            case 0: m1 = -Dm[0] - Dm[1] - Dm[2]; m2 = -Dm[3] - Dm[4] - Dm[5]; m3 = -Dm[6] - Dm[7] - Dm[8]; break;
            case 1: m1 = Dm[0]; m2 = Dm[3]; m3 = Dm[6]; break;
            case 2: m1 = Dm[1]; m2 = Dm[4]; m3 = Dm[7]; break;
            default: m1 = Dm[2]; m2 = Dm[5]; m3 = Dm[8]; break;
        }
        force -= V3(dE_dF[0] * m1 + dE_dF[3] * m2 + dE_dF[6] * m3, dE_dF[1] * m1 + dE_dF[4] * m2 + dE_dF[7] * m3, dE_dF[2] * m1 + dE_dF[5] * m2 + dE_dF[8] * m3);
    }
    return force;
}

void neohookean_force_hessian(const Scene& s, const Mesh& m, int iV, V3& force, M3& hessian) {
    const int start = m.vNbrInfos[2 * iV], num = m.vNbrInfos[2 * iV + 1];
    const float miu = m.obj.miu, lmbd = m.obj.lmbd;
    const float a = 1 + miu / lmbd;
    const V3 displacement = m.pos[iV] - m.prev[iV];
    for (int iNei = 0; iNei < num; iNei++) {
        const int tetId = m.vNbrTets[start + iNei];
        const float A = m.restVol[tetId];
        const float* Dm = &m.DmInv[9 * (size_t)tetId];
        const int* tv = &m.tets[4 * (size_t)tetId];
        // Ds, F = Ds * DmInv
        float Ds[9];
        {
            V3 v0 = m.pos[tv[0]];
            for (int c = 0; c < 3; c++) { V3 d = m.pos[tv[c + 1]] - v0; Ds[3 * c] = d.x; Ds[3 * c + 1] = d.y; Ds[3 * c + 2] = d.z; }
        }
        float F[9];
        for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++)
            F[r + 3 * c] = Ds[r] * Dm[3 * c] + Ds[r + 3] * Dm[3 * c + 1] + Ds[r + 6] * Dm[3 * c + 2];
        const float F1_1 = F[0], F2_1 = F[1], F3_1 = F[2], F1_2 = F[3], F2_2 = F[4], F3_2 = F[5], F1_3 = F[6], F2_3 = F[7], F3_3 = F[8];
        const float detF = F1_1 * (F2_2 * F3_3 - F2_3 * F3_2) - F1_2 * (F2_1 * F3_3 - F2_3 * F3_1) + F1_3 * (F2_1 * F3_2 - F2_2 * F3_1);
        float dd[9] = {F2_2 * F3_3 - F2_3 * F3_2, F1_3 * F3_2 - F1_2 * F3_3, F1_2 * F2_3 - F1_3 * F2_2,
                       F2_3 * F3_1 - F2_1 * F3_3, F1_1 * F3_3 - F1_3 * F3_1, F1_3 * F2_1 - F1_1 * F2_3,
                       F2_1 * F3_2 - F2_2 * F3_1, F1_2 * F3_1 - F1_1 * F3_2, F1_1 * F2_2 - F1_2 * F2_1};
        float H[81];  // 9x9 column-major; symmetric
        for (int c = 0; c < 9; c++) for (int r = 0; r < 9; r++) H[r + 9 * c] = dd[r] * dd[c];
        const float k = detF - a;
        auto addsym = [&](int r, int c, float v) { H[r + 9 * c] += v; H[c + 9 * r] += v; };
        addsym(0, 4, k * F3_3); addsym(0, 5, k * -F2_3); addsym(0, 7, k * -F3_2); addsym(0, 8, k * F2_2);
        addsym(1, 3, k * -F3_3); addsym(1, 5, k * F1_3); addsym(1, 6, k * F3_2); addsym(1, 8, k * -F1_2);
        addsym(2, 3, k * F2_3); addsym(2, 4, k * -F1_3); addsym(2, 6, k * -F2_2); addsym(2, 7, k * F1_2);
        addsym(3, 7, k * F3_1); addsym(3, 8, k * -F2_1);
        addsym(4, 6, k * -F3_1); addsym(4, 8, k * F1_1);
        addsym(5, 6, k * F2_1); addsym(5, 7, k * -F1_1);
        for (int q = 0; q < 81; q++) { H[q] *= lmbd; H[q] *= A; }
        float dE_dF[9];
        {
            const float lk = lmbd * (detF - a);
            for (int q = 0; q < 9; q++) dE_dF[q] = A * (miu * F[q] + lk * dd[q]);
        }
        const int ord = m.vNbrOrder[start + iNei];
        float m1, m2, m3;
        switch (ord) {
            case 0: m1 = -Dm[0] - Dm[1] - Dm[2]; m2 = -Dm[3] - Dm[4] - Dm[5]; m3 = -Dm[6] - Dm[7] - Dm[8]; break;
            case 1: m1 = Dm[0]; m2 = Dm[3]; m3 = Dm[6]; break;
            case 2: m1 = Dm[1]; m2 = Dm[4]; m3 = Dm[7]; break;
            default: m1 = Dm[2]; m2 = Dm[5]; m3 = Dm[8]; break;
        }
        // assembleVertexVForceAndHessian (VBD_NeoHookean.cpp:126-152)
        V3 dE_dxi(dE_dF[0] * m1 + dE_dF[3] * m2 + dE_dF[6] * m3,
                  dE_dF[1] * m1 + dE_dF[4] * m2 + dE_dF[7] * m3,
                  dE_dF[2] * m1 + dE_dF[5] * m2 + dE_dF[8] * m3);
        float HL[27];  // 3x9 column-major: HL(r,c) = HL[r + 3*c]
        for (int c = 0; c < 9; c++) for (int r = 0; r < 3; r++)
            HL[r + 3 * c] = H[r + 9 * c] * m1 + H[(r + 3) + 9 * c] * m2 + H[(r + 6) + 9 * c] * m3;
        M3 d2;
        for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++)
            d2(r, c) = HL[r + 3 * c] * m1 + HL[r + 3 * (c + 3)] * m2 + HL[r + 3 * (c + 6)] * m3;
        M3 dampingH;
        for (int q = 0; q < 9; q++) dampingH.m[q] = d2.m[q] * m.obj.dampingHydrostatic;
        float tmp = (m1 * m1 + m2 * m2 + m3 * m3) * miu * A;
        d2(0, 0) += tmp; d2(1, 1) += tmp; d2(2, 2) += tmp;
        tmp *= m.obj.dampingDeviatoric;
        dampingH(0, 0) += tmp; dampingH(1, 1) += tmp; dampingH(2, 2) += tmp;
        for (int q = 0; q < 9; q++) dampingH.m[q] /= s.dt;
        V3 dampingForce(dampingH(0, 0) * displacement.x + dampingH(0, 1) * displacement.y + dampingH(0, 2) * displacement.z,
                        dampingH(1, 0) * displacement.x + dampingH(1, 1) * displacement.y + dampingH(1, 2) * displacement.z,
                        dampingH(2, 0) * displacement.x + dampingH(2, 1) * displacement.y + dampingH(2, 2) * displacement.z);
        force -= dE_dxi + dampingForce;
        for (int q = 0; q < 9; q++) hessian.m[q] += d2.m[q] + dampingH.m[q];
    }
}

static void boundary_force_hessian(const Scene& s, const Mesh& m, int iV, V3& force, M3& hessian, bool apply_friction) {
    const float kB = s.phys.boundaryCollisionStiffness;
    if (s.phys.useBowlGround) {
        // the inside of a sphere as ground (VBDPhysics.cpp:2824-2864), restated as written - including the Hessian's sign
        const V3 center = s.bowl.center;
        const float radius = s.bowl.radius;
        const V3 pos = m.pos[iV];
        if (!s.bowl.cap && (pos.y - center.y > 0)) return;
        const float dist = norm(pos - center);
        if (dist < radius) return;
        const V3 dir = (pos - center) / dist;
        const float penetrationDepth = dist - radius;
        force += (-penetrationDepth * kB) * dir;
        const float rd = radius / dist;
        const float d[3] = {dir.x, dir.y, dir.z};
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) {
            const float I = r == c ? 1.f : 0.f, dd = d[r] * d[c];
            hessian(r, c) += kB * ((-I + dd) + rd * (I - dd));
        }
        if (apply_friction) {
            const V3 dx = m.pos[iV] - m.prev[iV];
            const V3 c0 = cross(dir, V3(0, 0, 1));
            const V3 t0 = c0 / norm(c0);
            const V3 c1 = cross(dir, t0);
            const V3 t1 = c1 / norm(c1);
            float T[6] = {t0.x, t0.y, t0.z, t1.x, t1.y, t1.z};
            float u[2] = {T[0] * dx.x + T[1] * dx.y + T[2] * dx.z, T[3] * dx.x + T[4] * dx.y + T[5] * dx.z};
            V3 ff; M3 fh;
            vertex_friction(s.phys.boundaryFrictionDynamic, penetrationDepth * kB, T, u, s.phys.boundaryFrictionEpsV * s.dt, ff, fh);
            force += ff;
            for (int q = 0; q < 9; q++) hessian.m[q] += fh.m[q];
        }
        return;
    }
    if (!s.phys.usePlaneGround) return;
    for (int iDim = 0; iDim < 3; iDim++) {
        const float dt = s.dt;
        const float lowerBound = s.phys.worldBoundsMin[iDim], upperBound = s.phys.worldBoundsMax[iDim];
        const float x = m.pos[iV][iDim];
        float penetrationDepth; bool hit = false;
        if (x < lowerBound) {
            penetrationDepth = lowerBound - x;
            force[iDim] += penetrationDepth * kB;
            hessian(iDim, iDim) += kB;
            hit = true;
        } else if (x > upperBound) {
            penetrationDepth = x - upperBound;
            force[iDim] -= penetrationDepth * kB;
            hessian(iDim, iDim) += kB;
            hit = true;
        }
        if (hit && apply_friction) {
            V3 dx = m.pos[iV] - m.prev[iV];
            float T[6] = {0, 0, 0, 0, 0, 0};
            T[(iDim + 1) % 3] = 1; T[3 + (iDim + 2) % 3] = 1;
            float u[2] = {T[0] * dx.x + T[1] * dx.y + T[2] * dx.z, T[3] * dx.x + T[4] * dx.y + T[5] * dx.z};
            const float lambda = penetrationDepth * kB;
            const float mu = s.phys.boundaryFrictionDynamic;
            const float epsU = s.phys.boundaryFrictionEpsV * dt;
            V3 ff; M3 fh;
            vertex_friction(mu, lambda, T, u, epsU, ff, fh);
            force += ff;
            for (int q = 0; q < 9; q++) hessian.m[q] += fh.m[q];
        }
    }
}

static void collision_per_collision(const ColResult& r, int order, V3& force, M3& hessian) {
    if (!r.shortestPathFound) return;
    if (r.penetrationDepth > 0) {
        float b = 0.f;
        switch (order) {
            case 0: b = -r.closestSurfacePtBarycentrics[0]; break;
            case 1: b = -r.closestSurfacePtBarycentrics[1]; break;
            case 2: b = -r.closestSurfacePtBarycentrics[2]; break;
            case 3: b = 1.0f; break;
        }
        if (b != 0) {
            force += r.collisionForce * b;
            for (int q = 0; q < 9; q++) hessian.m[q] += r.collisionHessian.m[q] * b * b;
        }
    }
}

static void collision_force_hessian(Scene& s, Mesh& m, int iV, V3& force, M3& hessian) {
    int sv = m.tetVertToSurf[iV];
    if (sv >= 0) {
        ColResult& r = m.colResults[sv];
        if (r.numIntersections) collision_per_collision(r, 3, force, hessian);
        for (const ColRelation& rel : m.colRelations[iV]) {
            ColResult& o = s.meshes[rel.meshId].colResults[rel.surfaceVertexId];
            collision_per_collision(o, rel.collisionVertexOrder, force, hessian);
        }
    }
}

// evaluateVertexMeritEnergy (M/VBD/VBD_NeoHookean.cpp:90-124): inertia + Neo-Hookean energy of the incident tets, with the
// vertex at `x` (the reference writes the trial position into the mesh and evaluates; here it is passed in)
void vertex_merit_energy_parts(const Scene& s, const Mesh& m, int iV, const V3& x, float& meInertiaOut, float& meElasticOut);
static float vertex_merit_energy(const Scene& s, const Mesh& m, int iV, const V3& x) {
    float a, b;
    vertex_merit_energy_parts(s, m, iV, x, a, b);
    return a + b;
}
void vertex_merit_energy_parts(const Scene& s, const Mesh& m, int iV, const V3& x, float& meInertiaOut, float& meElasticOut) {
    // 0.5 * vertexMass * squaredNorm * dtSqrReciprocal: the literal is a double, so the product is formed in double
    const V3 d = x - m.inertia[iV];
    const float meInertia = (float)(0.5 * (double)m.mass[iV] * (double)sqnorm(d) * (double)s.dtSqrReciprocal);
    float meElastic = 0.f;
    const int start = m.vNbrInfos[2 * iV], num = m.vNbrInfos[2 * iV + 1];
    const float miu = m.obj.miu, lmbd = m.obj.lmbd;
    const float a = 1 + miu / lmbd;
    for (int iNei = 0; iNei < num; iNei++) {
        const int tetId = m.vNbrTets[start + iNei];
        const float A = m.restVol[tetId];
        const float* Dm = &m.DmInv[9 * (size_t)tetId];
        const int* tv = &m.tets[4 * (size_t)tetId];
        V3 p[4];
        for (int k = 0; k < 4; k++) p[k] = tv[k] == iV ? x : m.pos[tv[k]];
        float Ds[9];
        for (int c = 0; c < 3; c++) { V3 e = p[c + 1] - p[0]; Ds[3 * c] = e.x; Ds[3 * c + 1] = e.y; Ds[3 * c + 2] = e.z; }
        float F[9];
        for (int c = 0; c < 3; c++) for (int r = 0; r < 3; r++)
            F[r + 3 * c] = Ds[r] * Dm[3 * c] + Ds[r + 3] * Dm[3 * c + 1] + Ds[r + 6] * Dm[3 * c + 2];
        float sum = 0.f;
        for (int q = 0; q < 9; q++) sum += F[q] * F[q];
        const float PhiD = 0.5f * (sum - 3.f);
        const float detF = F[0] * (F[4] * F[8] - F[7] * F[5]) - F[3] * (F[1] * F[8] - F[7] * F[2]) + F[6] * (F[1] * F[5] - F[4] * F[2]);
        const float PhiH = 0.5f * ((detF - a) * (detF - a));
        meElastic += A * (miu * PhiD + lmbd * PhiH);
    }
    meInertiaOut = meInertia; meElasticOut = meElastic;
}

// backTracingLineSearchVBD (M/VBD/VBD_NeoHookean.h:115-153) as called from VBDStepWithCollision (M/VBD/VBDPhysics.cpp:2467-2474):
// c = 0, first trial alpha = stepSize; the vertex is left at the LAST trial position (accepted or not), as in the reference
static int line_search(Scene& s, Mesh& m, int iV, const V3& dx, float alpha, float tau) {
    const V3 org = m.pos[iV];
    const float E0 = vertex_merit_energy(s, m, iV, org);
    const float c = 0.f, mm = sqnorm(dx);
    int shrinks = 0;
    for (int it = 0; it < s.lineSearch.maxIters; it++) {
        const V3 x = alpha * dx + org;
        m.pos[iV] = x;
        const float e = vertex_merit_energy(s, m, iV, x);
        if (e < E0 - alpha * c * mm) break;
        alpha = alpha * tau;
        shrinks++;
    }
    return shrinks;
}

// what VBDStepWithCollision does once the terms are summed (M/VBD/VBDPhysics.cpp:2430-2462; useDouble3x3 = false, the default): nothing
// below the force threshold; else the 3x3 solve, or - when it fails - the force itself with stepSizeGD.
// Returns 0 (no move), 1 (move by stepSize * dir), -1 (NaN direction: the reference exits).
int vertex_step_direction(float stepSizeNewton, float stepSizeGD, V3 force, const M3& h, V3& dir, float& stepSize, bool& solverOk) {
    solverOk = true; stepSize = stepSizeNewton; dir = V3();
    if (!(sqnorm(force) > CMP_EPSILON2)) return 0;
    float d[3];
    const float f[3] = {force.x, force.y, force.z};
    solverOk = solve3x3_psd_stable(h.m, f, d);
    if (!solverOk) { stepSize = stepSizeGD; d[0] = f[0]; d[1] = f[1]; d[2] = f[2]; }
    if (std::isnan(d[0]) || std::isnan(d[1]) || std::isnan(d[2])) return -1;
    dir = V3(d[0], d[1], d[2]);
    return 1;
}

// VBDBaseTetMesh::accumlateInertiaForceAndHessian (M/VBD/VBD_BaseMaterial.h:359-366)
void inertia_force_hessian(float mass, V3 inertia, V3 x, float dtSqrReciprocal, V3& force, M3& h) {
    force += (mass * (inertia - x)) * dtSqrReciprocal;
    const float hh = 1.f * mass * dtSqrReciprocal;
    h(0, 0) += hh; h(1, 1) += hh; h(2, 2) += hh;
}

static void vbd_step_with_collision(Scene& s, Mesh& m, int iV, bool apply_friction) {
    M3 h; V3 force;
    neohookean_force_hessian(s, m, iV, force, h);
    inertia_force_hessian(m.mass[iV], m.inertia[iV], m.pos[iV], s.dtSqrReciprocal, force, h);
    boundary_force_hessian(s, m, iV, force, h, apply_friction);
    collision_force_hessian(s, m, iV, force, h);
    // external force: zero (VBDBaseTetMesh::evaluateExternalForce, VBD_BaseMaterial.h:308-311)
    V3 dir; float stepSize; bool ok;
    const int rc = vertex_step_direction(s.phys.stepSize, s.phys.stepSizeGD, force, h, dir, stepSize, ok);
    if (rc < 0) { s.err = "descentDirection has NaN"; return; }
    if (rc > 0) {
        if (s.lineSearch.enabled) {
            const int k = line_search(s, m, iV, dir, stepSize, ok ? s.lineSearch.tau : s.lineSearch.tauGD);
            if (k) __atomic_fetch_add(&s.lineSearchShrinks, (int64_t)k, __ATOMIC_RELAXED);
        } else
            m.pos[iV] += stepSize * dir;
    }
}

static void apply_initial_step(Scene& s, Mesh& m) {
    const float dt = s.dt;
    const int nV = m.nV;
    m.prev = m.pos;
    if (m.hasVelocitiesPrev) {
        for (int v = 0; v < nV; v++) m.accel[v] = (m.vel[v] - m.velPrev[v]) / dt;
        m.hasApproxAcceleration = true;
    } else m.hasVelocitiesPrev = true;
    m.velPrev = m.vel;
    // computeInertia: external forces are zero; gravity associated with inertia
    if (s.phys.associateGravityWithInertia) {
        V3 dg = dt * V3(s.phys.gravity[0], s.phys.gravity[1], s.phys.gravity[2]);
        for (int v = 0; v < nV; v++) m.vel[v] += dg;
    }
    for (int fp : m.fixedPoints) m.vel[fp] = V3();
    for (int v = 0; v < nV; v++) m.inertia[v] = m.pos[v] + dt * m.vel[v];
    // forwardStep
    V3 g(s.phys.gravity[0], s.phys.gravity[1], s.phys.gravity[2]);
    const float gravNorm = norm(g);
    const V3 gravDir = g / gravNorm;
    switch (m.obj.initializationType) {
        case 0: m.pos = m.inertia; break;
        case 1: break;
        case 2: for (int v = 0; v < nV; v++) m.pos[v] = m.prev[v] + dt * m.velPrev[v]; break;
        case 3: for (int v = 0; v < nV; v++) m.pos[v] = 0.5f * m.pos[v] + 0.5f * m.inertia[v]; break;
        case 4:
            for (int v = 0; v < nV; v++) {
                m.pos[v] = m.pos[v] + dt * m.velPrev[v];
                m.pos[v] += (dt * dt * m.obj.initRatio_g) * g;
            }
            break;
        case 5: {
            // forwardStepSimplecticEuler (M/VBD/VBD_BaseMaterial.cpp:134-147): the internal forces of ALL vertices first, then
            // velocity = velocitiesPrev + dt (external 0 + internal) / mass and position += velocity dt - for every vertex, the
            // fixed ones included (the reference has no guard there). The velocity it stores is overwritten by updateVelocities.
            std::vector<V3> f(nV);
            for (int v = 0; v < nV; v++) f[v] = neohookean_force(m, v);
            for (int v = 0; v < nV; v++) {
                const V3 vel = m.velPrev[v] + (dt * (V3() + f[v])) * (1.f / m.mass[v]);
                m.vel[v] = vel;
                m.pos[v] += vel * dt;
            }
        } break;
        case 6:
            for (int v = 0; v < nV; v++) {
                if (m.hasApproxAcceleration) {
                    float ac = dot(m.accel[v], gravDir);
                    ac = ac < gravNorm ? ac : gravNorm;
                    ac = ac > 1e-5f ? ac : 0.f;
                    m.pos[v] = m.prev[v] + dt * m.velPrev[v] + (dt * dt) * gravDir * ac;
                }
            }
            break;
        default: break;
    }
}

static void update_velocities(Scene& s) {
    for (auto& m : s.meshes) {
        if (!m.activeForMaterialSolve) continue;
        s.pool->parallel_for(m.nV, [&](int64_t b, int64_t e) {
            for (int64_t v = b; v < e; v++) m.vel[v] = (m.pos[v] - m.prev[v]) / s.dt;
        });
    }
    // dampVelocities
    for (auto& m : s.meshes) {
        if (!m.activeForMaterialSolve) continue;
        s.pool->parallel_for(m.nV, [&](int64_t b, int64_t e) {
            for (int64_t v = b; v < e; v++) {
                const float vMag = norm(m.vel[v]);
                const float maxV = m.obj.maxVelocityMagnitude;
                if (m.obj.hasNoGravZone && maxV > 0 && m.pos[v][GRAVITY_AXIS] > m.obj.noGravZoneThreshold) {
                    if (vMag > maxV) m.vel[v] = m.vel[v] * (maxV / vMag);
                } else if (vMag > 1e-6f) {
                    float vMagNew = vMag * m.obj.exponentialVelDamping - m.obj.constantVelDamping;
                    vMagNew = vMagNew > 1e-6f ? vMagNew : 0.f;
                    m.vel[v] = m.vel[v] * (vMagNew / vMag);
                } else {
                    m.vel[v] = m.vel[v] * 0.f;
                }
            }
        });
        for (int fp : m.fixedPoints) m.vel[fp] = V3();
    }
}

// Chebyshev accelerator: a GPU-path feature of the reference (runStepGPU, VBDPhysics.cpp:869-907; omega
// VBDPhysics.cpp:3558-3574; apply kernel VBDPhysicsCompute.cu:910-944), restated here so the oracle covers it.
// the accelerator's position update (VBDSolveParallelGroup_applyAcceleration_kernel, M/VBD/VBDPhysicsCompute.cu:932-935)
V3 accelerate_position(V3 y, V3 yPrevPrev, float omega) { return yPrevPrev + omega * (y - yPrevPrev); }
float accelerator_omega(int order, float rho, float prevOmega) {      // VBDPhysics::getAcceleratorOmega (M/VBD/VBDPhysics.cpp:3558-3574)
    switch (order) {
        case 1: return 1;
        case 2: return 2 / (2 - rho * rho);
        default: return 4.f / (4 - rho * rho * prevOmega);
    }
}

void phase_solve_group(Scene& s, int g, bool apply_friction) {
    const std::vector<int>& grp = s.vertexParallelGroups[g];
    const std::vector<int>& act = s.activeCollisionsEachGroup[g];
    const int64_t nAct = (int64_t)act.size() / 2;
    s.pool->parallel_for(nAct, [&](int64_t b, int64_t e) {
        for (int64_t i = b; i < e; i++) {
            Mesh& m = s.meshes[act[2 * i]];
            s.ev(6, act[2 * i], act[2 * i + 1]);
            update_collision_info(s, m.colResults[m.tetVertToSurf[act[2 * i + 1]]]);
        }
    }, 16);
    const int64_t n = (int64_t)grp.size() / 2;
    std::atomic<int64_t> solved{0};
    s.pool->parallel_for(n, [&](int64_t b, int64_t e) {
        int64_t cnt = 0;
        for (int64_t i = b; i < e; i++) {
            Mesh& m = s.meshes[grp[2 * i]];
            const int v = grp[2 * i + 1];
            if (!m.fixedMask[v] && m.activeForMaterialSolve) { s.ev(7, grp[2 * i], v, apply_friction ? 1 : 0); vbd_step_with_collision(s, m, v, apply_friction); cnt++; }
        }
        solved += cnt;
    }, 64);
    s.stats.vertexSolves += solved.load();
    s.stats.colourSweeps++;
}

static void trace_active_lists(Scene& s) {
    if (!s.trace) return;
    for (size_t g = 0; g < s.activeCollisionsEachGroup.size(); g++)
        for (size_t i = 0; i + 1 < s.activeCollisionsEachGroup[g].size(); i += 2) s.ev(11, (int)g, s.activeCollisionsEachGroup[g][i], s.activeCollisionsEachGroup[g][i + 1]);
}

static void solve_iterations(Scene& s, int nIters, bool withIntermediateCD) {
    const bool accel = s.phys.useAccelerator != 0;
    std::vector<std::vector<V3>> pp, pIter;
    float omega = 1.f;
    if (accel) { for (auto& m : s.meshes) { pp.push_back(m.pos); pIter.push_back(m.pos); } }
    for (s.iIter = 0; s.iIter < nIters; s.iIter++) {
        const bool apply_friction = s.iIter >= s.phys.frictionStartIter;
        if (accel) {
            omega = accelerator_omega(s.iIter + 1, s.phys.acceleratorRho, omega);
            for (size_t im = 0; im < s.meshes.size(); im++) pIter[im] = s.meshes[im].pos;
        }
        for (size_t g = 0; g < s.vertexParallelGroups.size(); g++) phase_solve_group(s, (int)g, apply_friction);
        if (accel) {
            for (size_t im = 0; im < s.meshes.size(); im++) {
                Mesh& m = s.meshes[im];
                if (!m.activeForMaterialSolve) continue;
                for (int v = 0; v < m.nV; v++) {
                    if (m.fixedMask[v]) continue;
                    if (!m.activeCollisionMask[v]) m.pos[v] = accelerate_position(m.pos[v], pp[im][v], omega);
                    pp[im][v] = pIter[im][v];
                }
            }
        }
        const int ici = s.phys.intermediateCollisionIterations;
        if (withIntermediateCD && ici > 0 && s.iIter % ici == ici - 1) {
            s.ev(8);
            dcd(s, true);
            ccd(s, true);
            s.ev(5);
            prepare_collision_data(s);
            trace_active_lists(s);
        }
    }
}

static void apply_deformers(Scene& s) {
    for (Deformer& d : s.deformers) {
        Mesh& m = s.meshes[d.mesh];
        double t = (double)s.curTime;
        if (t >= d.endTime && d.endTime > 0) t = d.endTime;
        if (d.type == 0) {
            const double theta = t * d.angularVelocity;
            const double ux = d.axis.x, uy = d.axis.y, uz = d.axis.z;
            const double c = std::cos(theta), sn = std::sin(theta);
            const float R[9] = {(float)(c + ux * ux * (1 - c)), (float)(ux * uy * (1 - c) - uz * sn), (float)(ux * uz * (1 - c) + uy * sn),
                                (float)(uy * ux * (1 - c) + uz * sn), (float)(c + uy * uy * (1 - c)), (float)(uy * uz * (1 - c) - ux * sn),
                                (float)(uz * ux * (1 - c) - uy * sn), (float)(uz * uy * (1 - c) + ux * sn), (float)(c + uz * uz * (1 - c))};
            for (size_t i = 0; i < d.verts.size(); i++) {
                const V3 q = d.rootPs[i];
                const V3 rot(R[0] * q.x + R[1] * q.y + R[2] * q.z, R[3] * q.x + R[4] * q.y + R[5] * q.z, R[6] * q.x + R[7] * q.y + R[8] * q.z);
                m.pos[d.verts[i]] = d.roots[i] + rot;
            }
        } else {
            const float tf = (float)t;
            for (size_t i = 0; i < d.verts.size(); i++) m.pos[d.verts[i]] = d.speed * tf + d.roots[i];
        }
    }
}

void enable_models(Scene& s) {
    // VBD override tests frameToAppear <= frameId (VBDPhysics.cpp:691)
    for (auto& m : s.meshes) {
        if (m.obj.frameToAppear <= s.frameId) {
            m.activeForCollision = true;
            m.activeForMaterialSolve = true;
        }
    }
}

void phase_begin_substep(Scene& s) {
    s.curTime += s.dt;
    dcd(s, false);
    for (auto& m : s.meshes) if (m.activeForMaterialSolve) apply_initial_step(s, m);
    apply_deformers(s);
}
void phase_prepare(Scene& s) { ccd(s); prepare_collision_data(s); }
void phase_intermediate_cd(Scene& s) { dcd(s, true); ccd(s, true); prepare_collision_data(s); }   // the block at the end of an iteration above
void phase_end_substep(Scene& s) { update_velocities(s); s.stats.substeps++; }

void run_step(Scene& s) {
    for (s.substep = 0; s.substep < s.phys.numSubsteps; s.substep++) {
        s.curTime += s.dt;
        s.ev(1);
        dcd(s, false);
        for (size_t im = 0; im < s.meshes.size(); im++) if (s.meshes[im].activeForMaterialSolve) { s.ev(2, (int)im); apply_initial_step(s, s.meshes[im]); }
        s.ev(3);
        apply_deformers(s);
        s.ev(4);
        ccd(s);
        s.ev(5);
        prepare_collision_data(s);
        trace_active_lists(s);
        solve_iterations(s, s.phys.iterations, true);
        s.ev(9);
        update_velocities(s);
        s.stats.substeps++;
    }
    s.stats.frames++;
}

// probes for tests/test_oracle_vs_ref.py (pin the restatement against the reference's own compiled math)
void probe_boundary(const Scene& s, const Mesh& m, int iV, V3& force, M3& hessian, bool apply_friction) {
    boundary_force_hessian(s, m, iV, force, hessian, apply_friction);
}
void probe_collision_per_collision(const ColResult& r, int order, V3& force, M3& hessian) { collision_per_collision(r, order, force, hessian); }

void run_iterations(Scene& s, int nIters) {
    s.substep = 0;
    s.curTime += s.dt;
    dcd(s, false);
    for (auto& m : s.meshes) if (m.activeForMaterialSolve) apply_initial_step(s, m);
    apply_deformers(s);
    ccd(s);
    prepare_collision_data(s);
    solve_iterations(s, nIters, false);
}

}  // namespace orc
