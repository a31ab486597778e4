// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle_common.h).
// Vertex-triangle continuous collision test, float, restating
//   cy::IntersectContinuousTriPoint / MovingPointsPlanar / InterpolatedPoints2D / Barycentric
//                                         M/CollisionDetector/CCDSolver.h:103-202
//   cy::CubicRoots<float,false,RootFinderNewton>(roots, coef, x0, x1, xError=6e-4f)
//                                         3P/CyCodeBase/cyPolynomial.h:799-902
//   RootFinderNewton::FindClosed<3,float,false>   cyPolynomial.h:382-456
//   QuadraticRoots (generic, non-SSE path) cyPolynomial.h:596-618   [the `_INCLUDED_IMM` SSE specialisation is an
//                                         MSVC-only path; GCC builds of the reference take this generic one]
//   PolynomialEval / PolynomialDeflate     cyPolynomial.h:69, 117-121
// Pinned against the reference's own code compiled into oracle/_ref/libgaia_ref_ccd.so (tests/test_oracle_vs_ref.py).
#pragma once
#include <cmath>

namespace orc { namespace ccdm {

struct Vec3 {
    float x, y, z;
    Vec3() : x(0), y(0), z(0) {}
    Vec3(float a, float b, float c) : x(a), y(b), z(c) {}
};
inline Vec3 operator+(Vec3 a, Vec3 b) { return Vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline Vec3 operator-(Vec3 a, Vec3 b) { return Vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline Vec3 operator*(Vec3 a, float s) { return Vec3(a.x * s, a.y * s, a.z * s); }
inline Vec3 cross(Vec3 a, Vec3 p) { return Vec3(a.y * p.z - a.z * p.y, a.z * p.x - a.x * p.z, a.x * p.y - a.y * p.x); }
inline float dot(Vec3 a, Vec3 p) { return a.x * p.x + a.y * p.y + a.z * p.z; }
struct Vec2 { float x, y; };
inline float cross2(Vec2 a, Vec2 p) { return (-a.y) * p.x + a.x * p.y; }   // Vec2 r(-y,x); r.Dot(p)

inline bool is_different_sign(float a, float b) { return (a < 0) != (b < 0); }
inline float mult_sign(float v, float sign) { return v * (sign < 0 ? -1.f : 1.f); }
inline float eval3(const float c[4], float x) { float r = c[3]; for (int i = 2; i >= 0; --i) r = r * x + c[i]; return r; }
inline float eval2(const float c[3], float x) { float r = c[2]; for (int i = 1; i >= 0; --i) r = r * x + c[i]; return r; }
inline float fmin2(float a, float b) { return a <= b ? a : b; }
inline float fmax2(float a, float b) { return a >= b ? a : b; }

inline float find_closed3(const float coef[4], const float deriv[4], float x0, float x1, float y0, float y1, float xError) {
    (void)y1;
    const float ep2 = 2 * xError;
    float xr = (x0 + x1) / 2;
    if (x1 - x0 <= ep2) return xr;
    {
        const float xr0 = xr;
        for (int k = 0; k < 16; ++k) {
            float xn = xr - eval3(coef, xr) / eval2(deriv, xr);
            xn = fmin2(x1, fmax2(x0, xn));   // Clamp(xn, x0, x1) = Min(maxVal, Max(minVal, v))
            if (std::fabs(xr - xn) <= xError) return xn;
            xr = xn;
        }
        if (!std::isfinite(xr)) xr = xr0;
    }
    float yr = eval3(coef, xr);
    float xb0 = x0, xb1 = x1;
    for (;;) {
        const int side = is_different_sign(y0, yr);
        if (side) xb1 = xr; else xb0 = xr;
        const float dy = eval2(deriv, xr);
        const float dx = yr / dy;
        const float xn = xr - dx;
        if (xn > xb0 && xn < xb1) {
            const float stepsize = std::fabs(xr - xn);
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

inline int quadratic_roots(float roots[2], const float coef[3], float x0, float x1) {
    const float c = coef[0], b = coef[1], a = coef[2];
    const float delta = b * b - 4 * a * c;
    if (delta > 0) {
        const float d = std::sqrt(delta);
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

inline void deflate3(float def[3], const float coef[4], float root) {
    def[2] = coef[3];
    for (int i = 2; i > 0; --i) def[i - 1] = coef[i] + root * def[i];
}

inline int cubic_roots(float roots[3], const float coef[4], float x0, float x1, float xError = 6e-4f) {
    const float y0 = eval3(coef, x0), y1 = eval3(coef, x1);
    const float a = coef[3] * 3, b_2 = coef[2], c = coef[1];
    const float deriv[4] = {c, 2 * b_2, a, 0};
    const float delta_4 = b_2 * b_2 - a * c;
    if (delta_4 > 0) {
        const float d_2 = std::sqrt(delta_4);
        const float q = -(b_2 + mult_sign(d_2, b_2));
        const float rv0 = q / a, rv1 = c / q;
        const float xa = fmin2(rv0, rv1), xb = fmax2(rv0, rv1);
        if (is_different_sign(y0, y1)) {
            if (xa >= x1 || xb <= x0 || (xa <= x0 && xb >= x1)) {
                roots[0] = find_closed3(coef, deriv, x0, x1, y0, y1, xError);
                return 1;
            }
        } else {
            if ((xa >= x1 || xb <= x0) || (xa <= x0 && xb >= x1)) return 0;
        }
        int numRoots = 0;
        if (xa > x0) {
            const float ya = eval3(coef, xa);
            if (is_different_sign(y0, ya)) {
                roots[0] = find_closed3(coef, deriv, x0, xa, y0, ya, xError);
                if (is_different_sign(ya, y1) || (xb < x1 && is_different_sign(ya, eval3(coef, xb)))) {
                    float def[4];
                    deflate3(def, coef, roots[0]);
                    return quadratic_roots(roots + 1, def, xa, x1) + 1;
                } else return 1;
            }
            if (xb < x1) {
                const float yb = eval3(coef, xb);
                if (is_different_sign(ya, yb)) {
                    roots[0] = find_closed3(coef, deriv, xa, xb, ya, yb, xError);
                    if (is_different_sign(yb, y1)) {
                        float def[4];
                        deflate3(def, coef, roots[0]);
                        return quadratic_roots(roots + 1, def, xb, x1) + 1;
                    } else return 1;
                }
                if (is_different_sign(yb, y1)) {
                    roots[0] = find_closed3(coef, deriv, xb, x1, yb, y1, xError);
                    return 1;
                }
            } else {
                if (is_different_sign(ya, y1)) {
                    roots[0] = find_closed3(coef, deriv, xa, x1, ya, y1, xError);
                    return 1;
                }
            }
        } else {
            const float yb = eval3(coef, xb);
            if (is_different_sign(y0, yb)) {
                roots[0] = find_closed3(coef, deriv, x0, xb, y0, yb, xError);
                if (is_different_sign(yb, y1)) {
                    float def[4];
                    deflate3(def, coef, roots[0]);
                    return quadratic_roots(roots + 1, def, xb, x1) + 1;
                } else return 1;
            }
            if (is_different_sign(yb, y1)) {
                roots[0] = find_closed3(coef, deriv, xb, x1, yb, y1, xError);
                return 1;
            }
        }
        return numRoots;
    } else {
        if (is_different_sign(y0, y1)) {
            roots[0] = find_closed3(coef, deriv, x0, x1, y0, y1, xError);
            return 1;
        }
        return 0;
    }
}

inline int moving_points_planar(float t[3], const Vec3 x0[4], const Vec3 x1[4]) {
    const Vec3 dx0 = x1[0] - x0[0], dx1 = x1[1] - x0[1], dx2 = x1[2] - x0[2], dx3 = x1[3] - x0[3];
    const Vec3 e1 = x0[1] - x0[0], e2 = x0[2] - x0[0], e3 = x0[3] - x0[0];
    const Vec3 de0 = dx3 - dx0, de1 = dx1 - dx0, de2 = dx2 - dx0;
    const Vec3 e1xe2 = cross(e1, e2);
    const Vec3 de1xde2 = cross(de1, de2);
    const Vec3 dde1xe2 = cross(e1, de2) + cross(de1, e2);
    float poly[4];
    poly[0] = dot(e3, e1xe2);
    poly[1] = dot(de0, e1xe2) + dot(e3, dde1xe2);
    poly[2] = dot(e3, de1xde2) + dot(de0, dde1xe2);
    poly[3] = dot(de0, de1xde2);
    return cubic_roots(t, poly, 0.f, 1.f);
}

inline void interpolated_points_2d(Vec2 y[4], float t, const Vec3 x0[4], const Vec3 x1[4]) {
    Vec3 xx[4];
    for (int i = 0; i < 4; i++) xx[i] = x0[i] + (x1[i] - x0[i]) * t;
    const Vec3 nrm = cross(xx[1] - xx[0], xx[2] - xx[0]);
    const float ax = std::fabs(nrm.x), ay = std::fabs(nrm.y), az = std::fabs(nrm.z);
    const int xy = ax < ay, xz = ax < az, yz = ay < az;
    const int maxCmp = (xy | xz) + (xz & yz);
    for (int i = 0; i < 4; i++) {
        switch (maxCmp) {
            case 0: y[i].x = xx[i].y; y[i].y = xx[i].z; break;
            case 1: y[i].x = xx[i].x; y[i].y = xx[i].z; break;
            default: y[i].x = xx[i].x; y[i].y = xx[i].y; break;
        }
    }
}

inline void barycentric(Vec3 p, Vec3 a, Vec3 b, Vec3 c, Vec3& out) {
    const Vec3 v0 = b - a, v1 = c - a, v2 = p - a;
    const float d00 = dot(v0, v0), d01 = dot(v0, v1), d11 = dot(v1, v1), d20 = dot(v2, v0), d21 = dot(v2, v1);
    const float denom = d00 * d11 - d01 * d01;
    out.y = (d11 * d20 - d01 * d21) / denom;
    out.z = (d00 * d21 - d01 * d20) / denom;
    out.x = 1.0f - out.y - out.z;
}

inline bool IntersectContinuousTriPoint(float& tt, const Vec3 x[2][3], const Vec3 p[2], Vec3& barys) {
    const Vec3 xx0[4] = {x[0][0], x[0][1], x[0][2], p[0]};
    const Vec3 xx1[4] = {x[1][0], x[1][1], x[1][2], p[1]};
    float t[3];
    const int numt = moving_points_planar(t, xx0, xx1);
    for (int i = 0; i < numt; ++i) {
        Vec2 y[4];
        interpolated_points_2d(y, t[i], xx0, xx1);
        for (int k = 0; k < 3; k++) { y[k].x -= y[3].x; y[k].y -= y[3].y; }
        const bool y01p = std::signbit(cross2(y[0], y[1]));
        const bool y12p = std::signbit(cross2(y[1], y[2]));
        const bool y20p = std::signbit(cross2(y[2], y[0]));
        if (y01p == y12p && y01p == y20p) {
            tt = t[i];
            barycentric(p[0] + (p[1] - p[0]) * tt, x[0][0] + (x[1][0] - x[0][0]) * tt, x[0][1] + (x[1][1] - x[0][1]) * tt,
                        x[0][2] + (x[1][2] - x[0][2]) * tt, barys);
            return true;
        }
    }
    return false;
}

}}  // namespace orc::ccdm
