// FP64 device functions that follow the reference's operation order literally.
//
// They are (a) the minimal solver, one thread per hypothesis, (b) the "exact" residuals used for the
// guard-band re-check, for hypotheses outside the FP32 path's proven range and for the final
// re-scoring of near-best candidates.  This translation unit is compiled with -fmad=false so that,
// like the reference build (CMakeLists.txt:7-9,30-35: no FMA contraction), a*b+c stays two roundings.
// Remaining differences to the CPU reference are the last-ulp behaviour of CUDA's vs glibc's
// sin/cos/pow/log/atan2 (fmod, sqrt, + - * / are exact/IEEE on both).
//
// Reference locations (relative to /root/reference/src/pygcransac/include/):
//   math_utils.hpp:78-109,138-321   model.h:128-204   estimators/solver_rectifying_homography_two_sift.hpp
#pragma once
#include <cfloat>
#include <cstdint>

namespace gcr
{

constexpr double kPi = 3.14159265358979323846;
constexpr double kPi2 = 1.57079632679489661923;

// math_utils.hpp:78-88
__device__ __forceinline__ double clip_angle(double a)
{
    const double two_pi = 2.0 * kPi;
    a = fmod(a, two_pi);
    if (a < 0.0) a += two_pi;
    return a;
}

// math_utils.hpp:90-95
__device__ __forceinline__ double min_angle_diff(double a, double b)
{
    const double two_pi = 2.0 * kPi;
    const double d = fabs(clip_angle(a) - clip_angle(b));
    return fmin(d, two_pi - d);
}

// math_utils.hpp:97-102
__device__ __forceinline__ double lines_angles_diff(double a, double b)
{
    return fmin(min_angle_diff(a, b), min_angle_diff(a, b - kPi));
}

struct Model4
{
    double h7, h8, alpha, phi;
};

// two_sift.hpp:621-645 with the identity normaliser (model.h:42-55, S6); model.h:194-204
// `original`: the residual of three_sift_original.hpp:293-317, |log(rectified / alpha^3)| (its alpha is the reciprocal
// class scale); otherwise three_sift.hpp:293-317 == two_sift.hpp:621-645, |log(alpha^3 * rectified)|.
// (not inlined: one copy per kernel keeps the cold-start instruction footprint of the tail small)
__device__ __noinline__ double scale_residual(double x, double y, double s, const Model4& m, bool original = false)
{
    const double rectified = s * pow(-m.h7 * x - m.h8 * y + 1.0, -3.0);
    if (rectified < 1e-9) return DBL_MAX;
    const double alpha_cube = m.alpha * m.alpha * m.alpha;
    return original ? fabs(log(rectified / alpha_cube)) : fabs(log(alpha_cube * rectified));
}

// two_sift.hpp:647-665; model.h:156-165
__device__ __noinline__ double orientation_residual(double x, double y, double t, const Model4& m)
{
    double st, ct;
    sincos(t, &st, &ct);
    const double numer = (-x * st + y * ct) * m.h7 + st;
    const double denom = (x * st - y * ct) * m.h8 + ct;
    const double a = clip_angle(atan2(numer, denom));
    return fmin(lines_angles_diff(m.phi, a), lines_angles_diff(clip_angle(m.phi + kPi2), a));
}

// ---------------------------------------------------------------------------------------------
// minimal solver

struct P2
{
    double x, y;
};

__device__ __forceinline__ bool p2_less(const P2& a, const P2& b) // math_utils.hpp:121-124
{
    return a.x < b.x || (a.x == b.x && a.y < b.y);
}

__device__ __forceinline__ double cross2(const P2& o, const P2& p, const P2& q) // math_utils.hpp:229-232
{
    return (p.x - o.x) * (q.y - o.y) - (p.y - o.y) * (q.x - o.x);
}

// math_utils.hpp:104-109
__device__ __forceinline__ void line_from_point_angle(double x, double y, double t, double l[3])
{
    double s, c;
    sincos(t, &s, &c);
    l[0] = s;
    l[1] = -c;
    l[2] = y * c - x * s;
}

// the same line from a precomputed (sin, cos) pair
__device__ __forceinline__ void line_from_sincos(double x, double y, double s, double c, double l[3])
{
    l[0] = s;
    l[1] = -c;
    l[2] = y * c - x * s;
}

__device__ __forceinline__ void cross3(const double a[3], const double b[3], double r[3])
{
    r[0] = a[1] * b[2] - a[2] * b[1];
    r[1] = a[2] * b[0] - a[0] * b[2];
    r[2] = a[0] * b[1] - a[1] * b[0];
}

// math_utils.hpp:138-155: SIGNED distance (S7)
__device__ __forceinline__ bool are_collinear(double x1, double y1, double x2, double y2, double x3, double y3,
                                              double tol)
{
    const double p1[3] = {x1, y1, 1.0}, p2[3] = {x2, y2, 1.0};
    double l[3];
    cross3(p1, p2, l);
    const double n = sqrt(l[0] * l[0] + l[1] * l[1]);
    l[0] = l[0] / n;
    l[1] = l[1] / n;
    l[2] = l[2] / n;
    const double dist = (l[0] * x3 + l[1] * y3) + l[2] * 1.0;
    return dist < tol;
}

// two_sift.hpp:141-215 (S8).  Convex hull of the 4 sample points: math_utils.hpp:239-284, point test :291-321.
// vp_out = l1 x l2 (unnormalised), reused by the solver (two_sift.hpp:309-311 recomputes the same value).
// sc0 / sc1: (sin, cos) of the two orientations as sincos returns them (hoisted to feature preparation)
__device__ inline bool is_valid_sample(const double s0[3], const double s1[3], const double o0[3],
                                       const double o1[3], const double sc0[2], const double sc1[2], double vp_out[3])
{
    double l1[3], l2[3], vp[3];
    line_from_sincos(o0[0], o0[1], sc0[0], sc0[1], l1);
    line_from_sincos(o1[0], o1[1], sc1[0], sc1[1], l2);
    cross3(l1, l2, vp);
    vp_out[0] = vp[0];
    vp_out[1] = vp[1];
    vp_out[2] = vp[2];
    if (fabs(vp[0]) < 1e-6 && fabs(vp[1]) < 1e-6 && fabs(vp[2]) < 1e-6) return false;
    if (fabs(vp[2]) < 1e-6) return true;
    const double z = vp[2];
    const double vx = vp[0] / z, vy = vp[1] / z;
    if (are_collinear(s0[0], s0[1], s1[0], s1[1], vx, vy, 1.0)) return false;

    P2 pts[4] = {{s0[0], s0[1]}, {s1[0], s1[1]}, {o0[0], o0[1]}, {o1[0], o1[1]}};
    // lexicographic sort of 4 points (std::sort; equal points are indistinguishable)
#pragma unroll
    for (int i = 1; i < 4; i++)
    {
#pragma unroll
        for (int j = i; j > 0; j--)
        {
            if (p2_less(pts[j], pts[j - 1]))
            {
                const P2 t = pts[j];
                pts[j] = pts[j - 1];
                pts[j - 1] = t;
            }
        }
    }
    P2 hull[8];
    int k = 0;
    for (int i = 0; i < 4; ++i)
    {
        while (k >= 2 && cross2(hull[k - 2], hull[k - 1], pts[i]) <= 0) k--;
        hull[k++] = pts[i];
    }
    const int t = k + 1;
    for (int i = 3; i > 0; --i)
    {
        while (k >= t && cross2(hull[k - 2], hull[k - 1], pts[i - 1]) <= 0) k--;
        hull[k++] = pts[i - 1];
    }
    int n = k - 1;
    if (n == 2)
    {
        if (fabs(hull[0].x - hull[1].x) < 1e-9 && fabs(hull[0].y - hull[1].y) < 1e-9) n = 1;
    }
    if (n >= 3)
    {
        const P2 q = {vx, vy};
        bool pos = false, neg = false, inside = true;
        for (int i = 0; i < n; i++)
        {
            const double cp = cross2(hull[i], hull[(i + 1) % n], q);
            if (cp > 0) pos = true;
            else if (cp < 0) neg = true;
            if (pos && neg)
            {
                inside = false;
                break;
            }
        }
        if (inside) return false;
    }
    return true;
}

// math_utils.hpp:167-221 gaussElimination<3> (S9): up-front pivotisation, no zero-pivot guard.
__device__ inline void gauss3(double m[3][4], double x[3])
{
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int k = i + 1; k < 3; k++)
            if (fabs(m[i][i]) < fabs(m[k][i]))
            {
#pragma unroll
                for (int j = 0; j <= 3; j++)
                {
                    const double t = m[i][j];
                    m[i][j] = m[k][j];
                    m[k][j] = t;
                }
            }
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int k = i + 1; k < 3; k++)
        {
            const double f = m[k][i] / m[i][i];
#pragma unroll
            for (int j = 0; j <= 3; j++) m[k][j] = m[k][j] - f * m[i][j];
        }
#pragma unroll
    for (int i = 0; i < 3; i++)
    {
        const int r = 2 - i;
        x[r] = m[r][3];
#pragma unroll
        for (int c = r + 1; c < 3; c++) x[r] = x[r] - m[r][c] * x[c];
        x[r] = x[r] / m[r][r];
    }
}

// two_sift.hpp:261-344 (S9) given vp = l1 x l2 and pw0 / pw1 = pow(s, 1/3) of the two scale features
__device__ inline bool estimate_minimal(const double s0[3], const double s1[3], double pw0, double pw1, const double vp_in[3],
                                        Model4& out)
{
    double m[3][4];
    m[0][0] = s0[0]; m[0][1] = s0[1]; m[0][2] = pw0; m[0][3] = 1.0;
    m[1][0] = s1[0]; m[1][1] = s1[1]; m[1][2] = pw1; m[1][3] = 1.0;
    m[2][0] = vp_in[0]; m[2][1] = vp_in[1]; m[2][2] = 0.0; m[2][3] = vp_in[2];
    double sol[3];
    gauss3(m, sol);
    if (sol[0] != sol[0] || sol[1] != sol[1] || sol[2] != sol[2]) return false;
    if (sol[2] < 1e-9) return false;
    const double z = -sol[0] * vp_in[0] - sol[1] * vp_in[1] + vp_in[2]; // model.h:128-131
    if (fabs(z) > 1e-9) return false;
    out.h7 = sol[0];
    out.h8 = sol[1];
    out.alpha = sol[2];
    out.phi = clip_angle(atan2(vp_in[1], vp_in[0]));
    return true;
}

// Scale-only problems (three_sift.hpp:111-193, three_sift_original.hpp): the sample is valid unless its three positions
// are "collinear" in the reference's signed sense (S7); rows [x, y, s^(1/3), 1] (original: [x, y, -s^(-1/3), -1]).
// pw: pow(s, 1/3) of the three features (original: pow(s, -1/3))
__device__ inline bool estimate_minimal_scale3(const double s0[3], const double s1[3], const double s2[3], const double pw[3],
                                               bool original, Model4& out)
{
    double m[3][4];
    const double* f[3] = {s0, s1, s2};
#pragma unroll
    for (int i = 0; i < 3; i++)
    {
        m[i][0] = f[i][0];
        m[i][1] = f[i][1];
        m[i][2] = original ? -pw[i] : pw[i];
        m[i][3] = original ? -1.0 : 1.0;
    }
    double sol[3];
    gauss3(m, sol);
    if (sol[0] != sol[0] || sol[1] != sol[1] || sol[2] != sol[2]) return false;
    if (sol[2] < 1e-9) return false;
    out.h7 = sol[0];
    out.h8 = sol[1];
    out.alpha = sol[2];
    out.phi = 0.0;
    return true;
}

// two_sift.hpp:45-61 (S10)
__device__ __forceinline__ bool is_valid_model(const Model4& m)
{
    return !(fmax(fabs(m.h7), fabs(m.h8)) >= 1e-3);
}

} // namespace gcr
