// Device-side geometric predicates of the collision path. Every function restates a reference routine with the SAME
// sequence of IEEE single operations (this header is only included from translation units compiled with -fmad=false),
// so that accept/reject decisions — and therefore contact sets — are bit-identical to the CPU path:
//   pointInTet / closestPointTriangle                M/CollisionDetector/CollisionGeometry.h:56-157
//   checkFeasibleRegion (edge / vertex)              M/CollisionDetector/DiscreteCollisionDetector.cpp:664-815
//   tetrahedralTraverseTo + helpers                  M/TetMesh/TetMeshFEM.cpp:503-742, TetMeshFEM.h:423-560,
//                                                    3P/CuMatrix/CuMatrix/Geometry/Geometry.h:183-206 (orthonormal basis)
//   cy::IntersectContinuousTriPoint, CubicRoots      M/CollisionDetector/CCDSolver.h:103-202, 3P/CyCodeBase/cyPolynomial.h:382-456,
//                                                    596-618, 799-902 (generic non-SSE paths, xError = 6e-4 for float)
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace gb {

#define GDEV __device__ __forceinline__

struct V3 { float x, y, z; };
GDEV V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
GDEV V3 v3(float4 p) { return v3(p.x, p.y, p.z); }
GDEV V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
GDEV V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
GDEV V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
GDEV V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
GDEV V3 operator*(float s, V3 a) { return v3(a.x * s, a.y * s, a.z * s); }
GDEV V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
GDEV float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
GDEV V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
GDEV float sqnorm(V3 a) { return dot(a, a); }
GDEV float norm(V3 a) { return sqrtf(dot(a, a)); }
GDEV float comp(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

enum ClosestType { AtA = 0, AtB, AtC, AtAB, AtBC, AtAC, AtInterior, NotFound };

__device__ const int kTetFaces[4][3] = {{1, 2, 3}, {2, 0, 3}, {0, 1, 3}, {1, 0, 2}};        // TetMeshFEM.cpp:22
__device__ const int kExit3[3][3] = {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}};                     // TetMeshFEM.cpp:25

GDEV int4 ld4(const int4* p) { return __ldg(p); }
GDEV int comp4(int4 v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w)); }

GDEV bool point_in_tet(V3 p, V3 a, V3 b, V3 c, V3 d) {
    const V3 AB = b - a, AC = c - a, AD = d - a;
    const float vol = dot(AB, cross(AC, AD));
    if (fabsf(vol) < 1e-5f) return false;
    const V3 vs[4] = {a, b, c, d};
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const V3 v1 = vs[kTetFaces[i][0]], v2 = vs[kTetFaces[i][1]], v3_ = vs[kTetFaces[i][2]];
        const V3 v = p - v1;
        if (dot(v, cross(v2 - v1, v3_ - v2)) * vol >= 0) return false;
    }
    return true;
}

GDEV V3 closest_point_triangle(V3 p, V3 a, V3 b, V3 c, V3& bary, int& type) {
    const V3 ab = b - a, ac = c - a, ap = p - a;
    const float d1 = dot(ab, ap), d2 = dot(ac, ap);
    if (d1 <= 0.f && d2 <= 0.f) { type = AtA; bary = v3(1, 0, 0); return a; }
    const V3 bp = p - b;
    const float d3 = dot(ab, bp), d4 = dot(ac, bp);
    if (d3 >= 0.f && d4 <= d3) { type = AtB; bary = v3(0, 1, 0); return b; }
    const V3 cp = p - c;
    const float d5 = dot(ab, cp), d6 = dot(ac, cp);
    if (d6 >= 0.f && d5 <= d6) { type = AtC; bary = v3(0, 0, 1); return c; }
    const float vc = d1 * d4 - d3 * d2;
    if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) {
        const float v = d1 / (d1 - d3);
        type = AtAB; bary = v3(1.0f - v, v, 0.f);
        return a + v * ab;
    }
    const float vb = d5 * d2 - d1 * d6;
    if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) {
        const float v = d2 / (d2 - d6);
        type = AtAC; bary = v3(1.0f - v, 0.f, v);
        return a + v * ac;
    }
    const float va = d3 * d6 - d5 * d4;
    if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) {
        type = AtBC;
        const float v = (d4 - d3) / ((d4 - d3) + (d5 - d6));
        bary = v3(0.f, 1.f - v, v);
        return b + v * (c - b);
    }
    const float denom = 1.f / (va + vb + vc);
    const float v = vb * denom, w = vc * denom;
    type = AtInterior; bary = v3(1.f - v - w, v, w);
    return (a + v * ab) + w * ac;
}

GDEV void build_orthonormal_basis(V3 n, V3& b1, V3& b2) {
    if (n.z < 0.f) {
        const float a = 1.0f / (1.0f - n.z);
        const float b = n.x * n.y * a;
        b1 = v3(1.0f - n.x * n.x * a, -b, n.x);
        b2 = v3(b, n.y * n.y * a - 1.0f, -n.y);
    } else {
        const float a = 1.0f / (1.0f + n.z);
        const float b = -n.x * n.y * a;
        b1 = v3(1.0f - n.x * n.x * a, b, -n.x);
        b2 = v3(b, 1.0f - n.y * n.y * a, -n.y);
    }
}

struct P2 { float x, y; };
GDEV float cross2(P2 a, P2 b) { return a.x * b.y - a.y * b.x; }

GDEV void exit_face_selection(const P2 p[4], bool possible[3], float eps) {
    P2 v1, v2;
    v1.x = p[1].x - p[0].x; v1.y = p[1].y - p[0].y; v2.x = p[2].x - p[0].x; v2.y = p[2].y - p[0].y;
    const float area = cross2(v1, v2);
    const float sgn = copysignf(1.0f, area);
    const float d0 = cross2(p[3], p[0]), d1 = cross2(p[3], p[1]), d2 = cross2(p[3], p[2]);
    possible[0] = (d1 * sgn >= -eps && d2 * sgn <= eps);
    possible[1] = (d2 * sgn >= -eps && d0 * sgn <= eps);
    possible[2] = (d0 * sgn >= -eps && d1 * sgn <= eps);
}

GDEV void origin_barycentric_2d(P2 a, P2 b, P2 c, float barys[3]) {
    P2 v0, v1, v2;
    v0.x = b.x - a.x; v0.y = b.y - a.y; v1.x = c.x - a.x; v1.y = c.y - a.y; v2.x = -a.x; v2.y = -a.y;
    const float den = v0.x * v1.y - v1.x * v0.y;
    barys[1] = (v2.x * v1.y - v1.x * v2.y) / den;
    barys[2] = (v0.x * v2.y - v2.x * v0.y) / den;
    barys[0] = 1.0f - barys[1] - barys[2];
}

// ---------------------------------------------------------------- continuous vertex-triangle test
namespace ccd {
GDEV bool is_different_sign(float a, float b) { return (a < 0) != (b < 0); }
GDEV float mult_sign(float v, float sign) { return v * (sign < 0 ? -1.f : 1.f); }
GDEV float eval3(const float c[4], float x) { float r = c[3]; r = r * x + c[2]; r = r * x + c[1]; r = r * x + c[0]; return r; }
GDEV float eval2(const float c[3], float x) { float r = c[2]; r = r * x + c[1]; r = r * x + c[0]; return r; }
GDEV float fmin2(float a, float b) { return a <= b ? a : b; }
GDEV float fmax2(float a, float b) { return a >= b ? a : b; }

static __device__ __noinline__ float find_closed3(const float coef[4], const float deriv[4], float x0, float x1, float y0, float xError) {
    const float ep2 = 2 * xError;
    float xr = (x0 + x1) / 2;
    if (x1 - x0 <= ep2) return xr;
    {
        const float xr0 = xr;
        for (int k = 0; k < 16; ++k) {
            float xn = xr - eval3(coef, xr) / eval2(deriv, xr);
            xn = fmin2(x1, fmax2(x0, xn));
            if (fabsf(xr - xn) <= xError) return xn;
            xr = xn;
        }
        if (!isfinite(xr)) xr = xr0;
    }
    float yr = eval3(coef, xr);
    float xb0 = x0, xb1 = x1;
    for (;;) {
        const bool side = is_different_sign(y0, yr);
        if (side) xb1 = xr; else xb0 = xr;
        const float dy = eval2(deriv, xr);
        const float dx = yr / dy;
        const float xn = xr - dx;
        if (xn > xb0 && xn < xb1) {
            const float stepsize = fabsf(xr - xn);
            xr = xn;
            if (stepsize > xError) yr = eval3(coef, xr);
            else break;
        } else {
            xr = (xb0 + xb1) / 2;
            if (xr == xb0 || xr == xb1 || xb1 - xb0 <= ep2) break;
            yr = eval3(coef, xr);
        }
    }
    return xr;
}

GDEV int quadratic_roots(float roots[2], const float coef[3], float x0, float x1) {
    const float c = coef[0], b = coef[1], a = coef[2];
    const float delta = b * b - 4 * a * c;
    if (delta > 0) {
        const float d = sqrtf(delta);
        const float q = -0.5f * (b + mult_sign(d, b));
        const float rv0 = q / a, rv1 = c / q;
        const float r0 = fmin2(rv0, rv1), r1 = fmax2(rv0, rv1);
        const int r0i = (r0 >= x0) & (r0 <= x1);
        const int r1i = (r1 >= x0) & (r1 <= x1);
        roots[0] = r0;
        roots[r0i] = r1;
        return r0i + r1i;
    } else if (delta < 0) return 0;
    const float r0 = -0.5f * b / a;
    roots[0] = r0;
    return (r0 >= x0) & (r0 <= x1);
}

GDEV void deflate3(float def[3], const float coef[4], float root) {
    def[2] = coef[3];
    def[1] = coef[2] + root * def[2];
    def[0] = coef[1] + root * def[1];
}

// roots has room for 4 floats (roots + 1 receives up to two quadratic roots, one of which may be discarded)
static __device__ __noinline__ int cubic_roots(float roots[4], const float coef[4], float x0, float x1) {
    const float xError = 6e-4f;
    const float y0 = eval3(coef, x0), y1 = eval3(coef, x1);
    const float a = coef[3] * 3, b_2 = coef[2], c = coef[1];
    const float deriv[4] = {c, 2 * b_2, a, 0};
    const float delta_4 = b_2 * b_2 - a * c;
    if (delta_4 > 0) {
        const float d_2 = sqrtf(delta_4);
        const float q = -(b_2 + mult_sign(d_2, b_2));
        const float rv0 = q / a, rv1 = c / q;
        const float xa = fmin2(rv0, rv1), xb = fmax2(rv0, rv1);
        if (is_different_sign(y0, y1)) {
            if (xa >= x1 || xb <= x0 || (xa <= x0 && xb >= x1)) {
                roots[0] = find_closed3(coef, deriv, x0, x1, y0, xError);
                return 1;
            }
        } else {
            if ((xa >= x1 || xb <= x0) || (xa <= x0 && xb >= x1)) return 0;
        }
        if (xa > x0) {
            const float ya = eval3(coef, xa);
            if (is_different_sign(y0, ya)) {
                roots[0] = find_closed3(coef, deriv, x0, xa, y0, xError);
                if (is_different_sign(ya, y1) || (xb < x1 && is_different_sign(ya, eval3(coef, xb)))) {
                    float def[3];
                    deflate3(def, coef, roots[0]);
                    return quadratic_roots(roots + 1, def, xa, x1) + 1;
                } else return 1;
            }
            if (xb < x1) {
                const float yb = eval3(coef, xb);
                if (is_different_sign(ya, yb)) {
                    roots[0] = find_closed3(coef, deriv, xa, xb, ya, xError);
                    if (is_different_sign(yb, y1)) {
                        float def[3];
                        deflate3(def, coef, roots[0]);
                        return quadratic_roots(roots + 1, def, xb, x1) + 1;
                    } else return 1;
                }
                if (is_different_sign(yb, y1)) {
                    roots[0] = find_closed3(coef, deriv, xb, x1, yb, xError);
                    return 1;
                }
            } else {
                if (is_different_sign(ya, y1)) {
                    roots[0] = find_closed3(coef, deriv, xa, x1, ya, xError);
                    return 1;
                }
            }
        } else {
            const float yb = eval3(coef, xb);
            if (is_different_sign(y0, yb)) {
                roots[0] = find_closed3(coef, deriv, x0, xb, y0, xError);
                if (is_different_sign(yb, y1)) {
                    float def[3];
                    deflate3(def, coef, roots[0]);
                    return quadratic_roots(roots + 1, def, xb, x1) + 1;
                } else return 1;
            }
            if (is_different_sign(yb, y1)) {
                roots[0] = find_closed3(coef, deriv, xb, x1, yb, xError);
                return 1;
            }
        }
        return 0;
    } else {
        if (is_different_sign(y0, y1)) {
            roots[0] = find_closed3(coef, deriv, x0, x1, y0, xError);
            return 1;
        }
        return 0;
    }
}

GDEV void barycentric(V3 p, V3 a, V3 b, V3 c, V3& out) {
    const V3 v0 = b - a, v1 = c - a, v2 = p - a;
    const float d00 = dot(v0, v0), d01 = dot(v0, v1), d11 = dot(v1, v1), d20 = dot(v2, v0), d21 = dot(v2, v1);
    const float denom = d00 * d11 - d01 * d01;
    out.y = (d11 * d20 - d01 * d21) / denom;
    out.z = (d00 * d21 - d01 * d20) / denom;
    out.x = 1.0f - out.y - out.z;
}

// x0[k], x1[k]: triangle vertex k at t=0 / t=1; p0, p1: the point. Returns the first coplanar time in [0,1] at which the
// point lies inside the triangle.
static __device__ __noinline__ bool tri_point(float& tt, const V3 x0[3], const V3 x1[3], V3 p0, V3 p1, V3& barys) {
    const V3 a0[4] = {x0[0], x0[1], x0[2], p0};
    const V3 a1[4] = {x1[0], x1[1], x1[2], p1};
    float t[4];
    int numt;
    {
        const V3 dx0 = a1[0] - a0[0], dx1 = a1[1] - a0[1], dx2 = a1[2] - a0[2], dx3 = a1[3] - a0[3];
        const V3 e1 = a0[1] - a0[0], e2 = a0[2] - a0[0], e3 = a0[3] - a0[0];
        const V3 de0 = dx3 - dx0, de1 = dx1 - dx0, de2 = dx2 - dx0;
        const V3 e1xe2 = cross(e1, e2);
        const V3 de1xde2 = cross(de1, de2);
        const V3 dde1xe2 = cross(e1, de2) + cross(de1, e2);
        float poly[4];
        poly[0] = dot(e3, e1xe2);
        poly[1] = dot(de0, e1xe2) + dot(e3, dde1xe2);
        poly[2] = dot(e3, de1xde2) + dot(de0, dde1xe2);
        poly[3] = dot(de0, de1xde2);
        numt = cubic_roots(t, poly, 0.f, 1.f);
    }
    for (int i = 0; i < numt; ++i) {
        V3 xx[4];
#pragma unroll
        for (int k = 0; k < 4; k++) xx[k] = a0[k] + (a1[k] - a0[k]) * t[i];
        const V3 nrm = cross(xx[1] - xx[0], xx[2] - xx[0]);
        const float ax = fabsf(nrm.x), ay = fabsf(nrm.y), az = fabsf(nrm.z);
        const int xy = ax < ay, xz = ax < az, yz = ay < az;
        const int maxCmp = (xy | xz) + (xz & yz);
        P2 y[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (maxCmp == 0) { y[k].x = xx[k].y; y[k].y = xx[k].z; }
            else if (maxCmp == 1) { y[k].x = xx[k].x; y[k].y = xx[k].z; }
            else { y[k].x = xx[k].x; y[k].y = xx[k].y; }
        }
#pragma unroll
        for (int k = 0; k < 3; k++) { y[k].x -= y[3].x; y[k].y -= y[3].y; }
        // Vec2::Cross: r = (-y, x); r.Dot(p)
        const bool s01 = signbit((-y[0].y) * y[1].x + y[0].x * y[1].y);
        const bool s12 = signbit((-y[1].y) * y[2].x + y[1].x * y[2].y);
        const bool s20 = signbit((-y[2].y) * y[0].x + y[2].x * y[0].y);
        if (s01 == s12 && s01 == s20) {
            tt = t[i];
            barycentric(p0 + (p1 - p0) * tt, x0[0] + (x1[0] - x0[0]) * tt, x0[1] + (x1[1] - x0[1]) * tt, x0[2] + (x1[2] - x0[2]) * tt, barys);
            return true;
        }
    }
    return false;
}
}  // namespace ccd

}  // namespace gb
