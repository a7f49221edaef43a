/*
 * TEST INFRASTRUCTURE ONLY.  CPU restatement (plain C, double precision) of the hypothesis stage
 * of yuvalnis/graph-cut-ransac's hybrid rectifying-homography RANSAC:
 *     minimal sample  ->  degeneracy test  ->  closed-form minimal solver  ->  multi-class MSAC
 *     score + inlier masks  ->  model validity  ->  best-model selection rule.
 * It is the checker for the CUDA path in graph-cut-ransac_b200/csrc.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it; the
 * product never does (it fails loudly when its CUDA library is missing).
 *
 * PARITY PIN: the reference's own tests (tests/unit_tests.cpp) pin only the helper math
 * (angles, collinearity, hull, point-in-polygon, rectify/unrectify round trip); they are re-hosted
 * in tests/test_oracle_known_answers.py against this file.  The solver / residual / MSAC / selection
 * functions are pinned against the reference ITSELF: its unmodified headers compiled into
 * oracle/_ref/libgcr_ref.so (oracle/Makefile; Eigen/OpenCV replaced by the storage-only stand-ins in
 * oracle/shim/), run in the build container by tests/golden/make_golden.py, whose outputs are
 * committed under tests/golden/ and compared bit-for-bit by tests/test_oracle_golden.py.
 *
 * Each function cites the reference file:line it follows (paths relative to
 * /root/reference/src/pygcransac/include/).  Floating-point order is the reference's; build with
 * -ffp-contract=off and no fast-math so no FMA is formed.
 */
#include <math.h>
#include <float.h>
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define GCR_O_PI 3.14159265358979323846
#define GCR_O_PI_2 1.57079632679489661923

enum { GCR_O_SAMPLE_VALID = 1, GCR_O_SOLVED = 2, GCR_O_MODEL_VALID = 4 };

/* ---------------------------------------------------------------- angle helpers */

/* math_utils.hpp:78-88 clipAngle */
double gcr_o_clip_angle(double a)
{
    const double two_pi = 2.0 * GCR_O_PI;
    a = fmod(a, two_pi);
    if (a < 0.0) a += two_pi;
    return a;
}

/* math_utils.hpp:90-95 minAngleDiff */
double gcr_o_min_angle_diff(double a, double b)
{
    const double two_pi = 2.0 * GCR_O_PI;
    const double d = fabs(gcr_o_clip_angle(a) - gcr_o_clip_angle(b));
    return fmin(d, two_pi - d);
}

/* math_utils.hpp:97-102 linesAnglesDiff */
double gcr_o_lines_angles_diff(double a, double b)
{
    const double d1 = gcr_o_min_angle_diff(a, b);
    const double d2 = gcr_o_min_angle_diff(a, b - GCR_O_PI);
    return fmin(d1, d2);
}

/* math_utils.hpp:66-76 deg2rad / rad2deg */
double gcr_o_deg2rad(double a) { return a * (GCR_O_PI / 180.0); }
double gcr_o_rad2deg(double a) { return a * (0.318309886183790671538 * 180.0); }

/* math_utils.hpp:57-64 nChoose2 */
uint64_t gcr_o_nchoose2(uint64_t n) { return n == 0 ? 0 : (n * (n - 1)) / 2; }

/* ---------------------------------------------------------------- geometry helpers */

/* math_utils.hpp:104-109 lineFromPointAndAngle: l = (sin t, -cos t, y cos t - x sin t) */
void gcr_o_line_from_point_angle(double x, double y, double t, double l[3])
{
    const double c = cos(t), s = sin(t);
    l[0] = s;
    l[1] = -c;
    l[2] = y * c - x * s;
}

/* Eigen cross3 as used at two_sift.hpp:175,311 and math_utils.hpp:147 */
static void cross3(const double a[3], const double b[3], double r[3])
{
    r[0] = a[1] * b[2] - a[2] * b[1];
    r[1] = a[2] * b[0] - a[0] * b[2];
    r[2] = a[0] * b[1] - a[1] * b[0];
}

/* math_utils.hpp:138-155 areCollinear: SIGNED distance of p3 to the line p1p2, "< tol" (S7) */
int gcr_o_are_collinear(double x1, double y1, double x2, double y2, double x3, double y3, double tol)
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

/* math_utils.hpp:167-221 gaussElimination<3>: up-front row "pivotisation" decided on the
 * un-eliminated matrix, forward elimination without pivot guard, back-substitution (S9). */
void gcr_o_gauss3(double m[3][4], double x[3])
{
    for (int i = 0; i < 3; i++)
        for (int k = i + 1; k < 3; k++)
            if (fabs(m[i][i]) < fabs(m[k][i]))
                for (int j = 0; j <= 3; j++)
                {
                    const double t = m[i][j];
                    m[i][j] = m[k][j];
                    m[k][j] = t;
                }
    for (int i = 0; i < 2; i++)
        for (int k = i + 1; k < 3; k++)
        {
            const double f = m[k][i] / m[i][i];
            for (int j = 0; j <= 3; j++) m[k][j] = m[k][j] - f * m[i][j];
        }
    for (int i = 0; i < 3; i++)
    {
        const int r = 2 - i;
        x[r] = m[r][3];
        for (int c = r + 1; c < 3; c++) x[r] = x[r] - m[r][c] * x[c];
        x[r] = x[r] / m[r][r];
    }
}

typedef struct { double x, y; } pt2;

/* math_utils.hpp:121-124 Point2D::operator< (lexicographic) */
static int pt_less(const pt2* a, const pt2* b)
{
    return a->x < b->x || (a->x == b->x && a->y < b->y);
}

/* math_utils.hpp:229-232 crossProduct(O, P, Q) */
static double cross2(const pt2* o, const pt2* p, const pt2* q)
{
    return (p->x - o->x) * (q->y - o->y) - (p->y - o->y) * (q->x - o->x);
}

/* math_utils.hpp:239-284 computeConvexHull (Andrew monotone chain; sorts `pts` in place).
 * `hull` must hold 2*n points; returns the vertex count. */
static int convex_hull(pt2* pts, int n, pt2* hull)
{
    if (n <= 1)
    {
        for (int i = 0; i < n; i++) hull[i] = pts[i];
        return n;
    }
    /* std::sort with operator<; equal elements are indistinguishable so any sort gives the same array */
    for (int i = 1; i < n; i++)
    {
        pt2 v = pts[i];
        int j = i - 1;
        while (j >= 0 && pt_less(&v, &pts[j]))
        {
            pts[j + 1] = pts[j];
            j--;
        }
        pts[j + 1] = v;
    }
    int k = 0;
    for (int i = 0; i < n; ++i)
    {
        while (k >= 2 && cross2(&hull[k - 2], &hull[k - 1], &pts[i]) <= 0) k--;
        hull[k++] = pts[i];
    }
    const int t = k + 1;
    for (int i = n - 1; i > 0; --i)
    {
        while (k >= t && cross2(&hull[k - 2], &hull[k - 1], &pts[i - 1]) <= 0) k--;
        hull[k++] = pts[i - 1];
    }
    int size = k - 1;
    if (size == 2)
    {
        const int close_x = fabs(hull[0].x - hull[1].x) < 1e-9;
        const int close_y = fabs(hull[0].y - hull[1].y) < 1e-9;
        if (close_x && close_y) size = 1;
    }
    return size;
}

int gcr_o_convex_hull(const double* pts, int n, double* hull_out)
{
    pt2 p[64], h[128];
    if (n > 64) return -1;
    for (int i = 0; i < n; i++) { p[i].x = pts[2 * i]; p[i].y = pts[2 * i + 1]; }
    const int m = convex_hull(p, n, h);
    for (int i = 0; i < m; i++) { hull_out[2 * i] = h[i].x; hull_out[2 * i + 1] = h[i].y; }
    return m;
}

/* math_utils.hpp:291-321 pointInConvexPolygon: boundary counts as inside, <3 vertices -> outside */
static int point_in_convex_polygon(const pt2* q, const pt2* poly, int n)
{
    if (n < 3) return 0;
    int pos = 0, neg = 0;
    for (int i = 0; i < n; i++)
    {
        const double cp = cross2(&poly[i], &poly[(i + 1) % n], q);
        if (cp > 0) pos = 1;
        else if (cp < 0) neg = 1;
        if (pos && neg) return 0;
    }
    return 1;
}

int gcr_o_point_in_convex_polygon(double x, double y, const double* poly, int n)
{
    pt2 p[128];
    if (n > 128) return -1;
    for (int i = 0; i < n; i++) { p[i].x = poly[2 * i]; p[i].y = poly[2 * i + 1]; }
    const pt2 q = {x, y};
    return point_in_convex_polygon(&q, p, n);
}

/* ---------------------------------------------------------------- model maps (model.h) */

/* model.h:128-131 rectifyPoint (homogeneous); :133-138 unrectifyPoint */
static double rectify_w(double h7, double h8, double x, double y, double w) { return -h7 * x - h8 * y + w; }
static double unrectify_w(double h7, double h8, double x, double y, double w) { return h7 * x + h8 * y + w; }

/* model.h:140-154 the (x,y) overloads */
void gcr_o_rectify_point(double h7, double h8, double* x, double* y)
{
    const double w = rectify_w(h7, h8, *x, *y, 1.0);
    *x = *x / w;
    *y = *y / w;
}
void gcr_o_unrectify_point(double h7, double h8, double* x, double* y)
{
    const double w = unrectify_w(h7, h8, *x, *y, 1.0);
    *x = *x / w;
    *y = *y / w;
}

/* model.h:156-165 rectifiedAngle */
double gcr_o_rectified_angle(double h7, double h8, double x, double y, double angle)
{
    const double ct = cos(angle), st = sin(angle);
    const double numer = (-x * st + y * ct) * h7 + st;
    const double denom = (x * st - y * ct) * h8 + ct;
    return gcr_o_clip_angle(atan2(numer, denom));
}

/* model.h:167-174 unrectifiedAngle */
double gcr_o_unrectified_angle(double h7, double h8, double x, double y, double angle)
{
    const double ct = cos(angle), st = sin(angle);
    const double numer = (x * st - y * ct) * h7 + st;
    const double denom = (-x * st + y * ct) * h8 + ct;
    return gcr_o_clip_angle(atan2(numer, denom));
}

/* model.h:194-204 rectifiedScale = ds * pow(-h7 dx - h8 dy + 1, -3) */
double gcr_o_rectified_scale(double h7, double h8, double x, double y, double s)
{
    return s * pow(-h7 * x - h8 * y + 1.0, -3.0);
}

/* model.h:182-186,206-209 unrectifiedScale = uds * pow(h7 udx + h8 udy + 1, -3) */
double gcr_o_unrectified_scale(double h7, double h8, double x, double y, double s)
{
    return s * pow(h7 * x + h8 * y + 1.0, -3.0);
}

/* ---------------------------------------------------------------- solver (two_sift.hpp) */

/* two_sift.hpp:141-215 isValidSample (S7, S8).  sf/of: the two sampled rows (x,y,s)/(x,y,theta). */
int gcr_o_is_valid_sample(const double* s0, const double* s1, const double* o0, const double* o1)
{
    double l1[3], l2[3], vp[3];
    gcr_o_line_from_point_angle(o0[0], o0[1], o0[2], l1);
    gcr_o_line_from_point_angle(o1[0], o1[1], o1[2], l2);
    cross3(l1, l2, vp);
    if (fabs(vp[0]) < 1e-6 && fabs(vp[1]) < 1e-6 && fabs(vp[2]) < 1e-6) return 0;
    if (fabs(vp[2]) < 1e-6) return 1;
    const double z = vp[2];
    vp[0] = vp[0] / z;
    vp[1] = vp[1] / z;
    vp[2] = vp[2] / z;
    if (gcr_o_are_collinear(s0[0], s0[1], s1[0], s1[1], vp[0], vp[1], 1.0)) return 0;
    pt2 pts[4] = {{s0[0], s0[1]}, {s1[0], s1[1]}, {o0[0], o0[1]}, {o1[0], o1[1]}};
    pt2 hull[8];
    const int n = convex_hull(pts, 4, hull);
    const pt2 q = {vp[0], vp[1]};
    if (point_in_convex_polygon(&q, hull, n)) return 0;
    return 1;
}

/* two_sift.hpp:261-344 estimateMinimalModel (S9).  out = (h7, h8, alpha, phi). */
int gcr_o_estimate_minimal(const double* s0, const double* s1, const double* o0, const double* o1,
                           double out[4])
{
    double m[3][4];
    /* two_sift.hpp:217-226 setScaleConstraint: [x, y, s^(1/3), 1] */
    m[0][0] = s0[0]; m[0][1] = s0[1]; m[0][2] = pow(s0[2], 1.0 / 3.0); m[0][3] = 1.0;
    m[1][0] = s1[0]; m[1][1] = s1[1]; m[1][2] = pow(s1[2], 1.0 / 3.0); m[1][3] = 1.0;
    double l1[3], l2[3], vp[3];
    gcr_o_line_from_point_angle(o0[0], o0[1], o0[2], l1);
    gcr_o_line_from_point_angle(o1[0], o1[1], o1[2], l2);
    cross3(l1, l2, vp);
    m[2][0] = vp[0]; m[2][1] = vp[1]; m[2][2] = 0.0; m[2][3] = vp[2];
    double sol[3];
    gcr_o_gauss3(m, sol);
    if (sol[0] != sol[0] || sol[1] != sol[1] || sol[2] != sol[2]) return 0; /* hasNaN */
    const double h7 = sol[0], h8 = sol[1], alpha = sol[2];
    if (alpha < 1e-9) return 0;
    /* model.rectifyPoint(vp): vp.z <- -h7 vp.x - h8 vp.y + vp.z, must vanish */
    vp[2] = rectify_w(h7, h8, vp[0], vp[1], vp[2]);
    if (fabs(vp[2]) > 1e-9) return 0;
    out[0] = h7;
    out[1] = h8;
    out[2] = alpha;
    out[3] = gcr_o_clip_angle(atan2(vp[1], vp[0]));
    return 1;
}

/* two_sift.hpp:45-61 isValidModel (S10) */
int gcr_o_is_valid_model(const double model[4])
{
    const double hmax = fmax(fabs(model[0]), fabs(model[1]));
    return hmax >= 1e-3 ? 0 : 1;
}

/* two_sift.hpp:621-645 scaleResidual (S4); the model's normalising transform is the identity (S6):
 * x <- 1*(x - 0*1), s <- s*1, both exact. */
double gcr_o_scale_residual(double x, double y, double s, const double model[4])
{
    const double rectified = s * pow(-model[0] * x - model[1] * y + 1.0, -3.0);
    if (rectified < 1e-9) return DBL_MAX;
    const double alpha_cube = model[2] * model[2] * model[2];
    return fabs(log(alpha_cube * rectified));
}

/* two_sift.hpp:647-665 orientationResidual (S5) */
double gcr_o_orientation_residual(double x, double y, double t, const double model[4])
{
    const double a = gcr_o_rectified_angle(model[0], model[1], x, y, t);
    return fmin(gcr_o_lines_angles_diff(model[3], a),
                gcr_o_lines_angles_diff(gcr_o_clip_angle(model[3] + GCR_O_PI_2), a));
}

/* ---------------------------------------------------------------- MSAC (MSAC_scoring_function.hpp) */

/* MSAC_scoring_function.hpp:53-130 getScore + score.hpp:38-58 bookkeeping (S1-S3).
 * smask/omask: nullable u32 words, bit j of word w <-> feature 32w+j, set for every feature passing
 * the threshold test (the reference's inlier index lists, which the zero-reset does NOT clear).
 * n_inl: Score counts AFTER the zero-reset rule.  Returns Score::value(). */
double gcr_o_score(const double* scale, long ns, const double* orient, long no, const double model[4],
                   const double thr[2], uint32_t n_inl[2], uint32_t* smask, uint32_t* omask)
{
    const double tsq[2] = {(2.25 * thr[0]) * thr[0], (2.25 * thr[1]) * thr[1]}; /* S1 */
    uint64_t cnt[2] = {0, 0};
    double val[2] = {0.0, 0.0}, sum = 0.0;
    if (smask) memset(smask, 0, sizeof(uint32_t) * (size_t)((ns + 31) / 32));
    if (omask) memset(omask, 0, sizeof(uint32_t) * (size_t)((no + 31) / 32));
    for (long i = 0; i < ns; i++)
    {
        const double r = gcr_o_scale_residual(scale[3 * i], scale[3 * i + 1], scale[3 * i + 2], model);
        const double r2 = r * r;
        if (r2 <= tsq[0])
        {
            cnt[0]++;
            val[0] += -r2;
            sum += -r2;
            if (smask) smask[i >> 5] |= 1u << (i & 31);
        }
    }
    for (long i = 0; i < no; i++)
    {
        const double r = gcr_o_orientation_residual(orient[3 * i], orient[3 * i + 1], orient[3 * i + 2], model);
        const double r2 = r * r;
        if (r2 <= tsq[1])
        {
            cnt[1]++;
            val[1] += -r2;
            sum += -r2;
            if (omask) omask[i >> 5] |= 1u << (i & 31);
        }
    }
    /* :108-127 with sampleSize() = {2,2}; reset_value (score.hpp:52-58) updates the running sum */
    for (int i = 0; i < 2; i++)
    {
        if (cnt[i] < 2)
        {
            n_inl[0] = n_inl[1] = 0;
            return 0.0;
        }
        const double msac = val[i] / tsq[i] + (double)cnt[i];
        sum -= val[i];
        val[i] = msac;
        sum += msac;
    }
    n_inl[0] = (uint32_t)cnt[0];
    n_inl[1] = (uint32_t)cnt[1];
    return sum;
}

/* ---------------------------------------------------------------- hypothesis stage */

/* GCRANSAC.h:304-446 for one replayed minimal sample: isValidSample -> estimateModel -> getScore ->
 * isValidModel.  samples int32[H,4] = (scale0, scale1, orient0, orient1).  Outputs as in
 * oracle/ref_harness.cpp::gcr_ref_replay.  Returns the thread count used. */
int gcr_o_replay(const double* scale, long ns, const double* orient, long no, const int32_t* samples,
                 long H, const double* thr, uint8_t* flags, double* models, double* score,
                 uint32_t* n_inl, uint32_t* masks, int nthreads)
{
    const long ws = (ns + 31) / 32, wo = (no + 31) / 32, wt = ws + wo;
    int used = 1;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
    used = nthreads > 0 ? nthreads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 16)
#endif
    for (long h = 0; h < H; h++)
    {
        const double* s0 = scale + 3 * (long)samples[4 * h + 0];
        const double* s1 = scale + 3 * (long)samples[4 * h + 1];
        const double* o0 = orient + 3 * (long)samples[4 * h + 2];
        const double* o1 = orient + 3 * (long)samples[4 * h + 3];
        uint8_t f = 0;
        double m[4] = {0, 0, 0, 0}, val = 0.0;
        uint32_t cnt[2] = {0, 0};
        uint32_t* sm = masks ? masks + h * wt : NULL;
        uint32_t* om = masks ? masks + h * wt + ws : NULL;
        if (masks) memset(masks + h * wt, 0, sizeof(uint32_t) * (size_t)wt);
        if (gcr_o_is_valid_sample(s0, s1, o0, o1))
        {
            f |= GCR_O_SAMPLE_VALID;
            if (gcr_o_estimate_minimal(s0, s1, o0, o1, m))
            {
                f |= GCR_O_SOLVED;
                val = gcr_o_score(scale, ns, orient, no, m, thr, cnt, sm, om);
                if (gcr_o_is_valid_model(m)) f |= GCR_O_MODEL_VALID;
            }
            else
            {
                m[0] = m[1] = m[2] = m[3] = 0.0;
            }
        }
        flags[h] = f;
        for (int k = 0; k < 4; k++) models[4 * h + k] = m[k];
        score[h] = val;
        n_inl[2 * h] = cnt[0];
        n_inl[2 * h + 1] = cnt[1];
    }
    return used;
}

/* GCRANSAC.h:440-462 + score.hpp:28-31 (S10, S11): walk the hypotheses in order, keep the first one
 * whose score strictly beats the best so far (initially 0) AND whose model is valid.
 * Returns the winning hypothesis index or -1. */
long gcr_o_select_best(const double* score, const uint8_t* flags, long H)
{
    double best = 0.0;
    long idx = -1;
    for (long h = 0; h < H; h++)
    {
        if (!(flags[h] & GCR_O_SOLVED)) continue;
        if (best < score[h] && (flags[h] & GCR_O_MODEL_VALID))
        {
            best = score[h];
            idx = h;
        }
    }
    return idx;
}

/* GCRANSAC.h:738-757 getIterationNumber for sample sizes {2,2} */
uint64_t gcr_o_iteration_number(const uint32_t n_inl[2], long ns, long no, double confidence)
{
    double q = 1.0;
    q *= pow((double)n_inl[0] / (double)ns, 2.0);
    q *= pow((double)n_inl[1] / (double)no, 2.0);
    const double l = log(1 - q);
    if (fabs(l) < DBL_EPSILON) return UINT64_MAX;
    return (uint64_t)ceil(log(1.0 - confidence) / l);
}

/* ---------------------------------------------------------------- sampler restatement
 * The reference draws each class's 2 indices as the head of a full std::shuffle seeded from
 * std::random_device (GCRANSAC.h:53-80): a uniform ordered pair without replacement, not replayable.
 * The B200 engine draws the same distribution from Philox4x32-10 (Salmon et al., SC'11; public
 * algorithm) keyed by (seed), counter = global hypothesis index; this is its CPU twin so the tests
 * can check the device draws bit-for-bit. */
static void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1)
{
    for (int r = 0; r < 10; r++)
    {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
}

/* ordered pair (a,b), a != b, uniform over [0,n): a = floor(u0 n / 2^32), b from the remaining n-1 */
static void draw_pair(uint32_t u0, uint32_t u1, uint32_t n, int32_t* a, int32_t* b)
{
    const uint32_t i0 = (uint32_t)(((uint64_t)u0 * n) >> 32);
    uint32_t i1 = (uint32_t)(((uint64_t)u1 * (n - 1)) >> 32);
    if (i1 >= i0) i1++;
    *a = (int32_t)i0;
    *b = (int32_t)i1;
}

void gcr_o_sample_philox(uint64_t seed, uint64_t first_hyp, long H, long ns, long no, int32_t* samples)
{
    for (long h = 0; h < H; h++)
    {
        const uint64_t g = first_hyp + (uint64_t)h;
        uint32_t c[4] = {(uint32_t)g, (uint32_t)(g >> 32), 0x47435231u /* "GCR1" */, 0u};
        philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
        draw_pair(c[0], c[1], (uint32_t)ns, &samples[4 * h + 0], &samples[4 * h + 1]);
        draw_pair(c[2], c[3], (uint32_t)no, &samples[4 * h + 2], &samples[4 * h + 3]);
    }
}

/* ================================================================ scale-only estimators
 * findRectifyingHomographyScaleOnly / ...Original: one feature class, minimal sample of 3 scale features.
 * `original` != 0 selects estimators/solver_rectifying_homography_three_sift_original.hpp, whose rows and
 * residual differ from three_sift.hpp as cited below.  Model = (h7, h8, alpha, 0). */

/* three_sift.hpp:111-143 areAllPointsCollinear for a 3-sample + :40-43 isValidSample (= !collinear).  The
 * collinearity test is the SIGNED one of math_utils.hpp:138-155 (S7). */
int gcr_o_is_valid_sample_scale3(const double* s0, const double* s1, const double* s2)
{
    return gcr_o_are_collinear(s0[0], s0[1], s1[0], s1[1], s2[0], s2[1], 1.0) ? 0 : 1;
}

/* three_sift.hpp:145-193 estimateMinimalModel: rows [x, y, s^(1/3), 1] through gaussElimination<3>;
 * three_sift_original.hpp:170-173: rows [x, y, -s^(-1/3), -1].  hasNaN reject, alpha < 1e-9 reject. */
int gcr_o_estimate_minimal_scale3(const double* s0, const double* s1, const double* s2, int original, double model[4])
{
    const double* f[3] = {s0, s1, s2};
    double m[3][4], x[3];
    for (int i = 0; i < 3; i++)
    {
        m[i][0] = f[i][0];
        m[i][1] = f[i][1];
        m[i][2] = original ? -pow(f[i][2], -1.0 / 3.0) : pow(f[i][2], 1.0 / 3.0);
        m[i][3] = original ? -1.0 : 1.0;
    }
    gcr_o_gauss3(m, x);
    if (isnan(x[0]) || isnan(x[1]) || isnan(x[2])) return 0;
    if (x[2] < 1e-9) return 0;
    model[0] = x[0];
    model[1] = x[1];
    model[2] = x[2];
    model[3] = 0.0;
    return 1;
}

/* three_sift.hpp:293-317 (== the hybrid scale residual) / three_sift_original.hpp:293-317 */
double gcr_o_scale_residual_variant(double x, double y, double s, const double model[4], int original)
{
    if (!original) return gcr_o_scale_residual(x, y, s, model);
    const double rectified = s * pow(-model[0] * x - model[1] * y + 1.0, -3.0);
    if (rectified < 1e-9) return DBL_MAX;
    const double alpha_cube = model[2] * model[2] * model[2];
    return fabs(log(rectified / alpha_cube));
}

/* MSAC for one class (MSAC_scoring_function.hpp:53-130 with sampleSize() = {3}).  n_inl after the zero-reset. */
double gcr_o_score_scale(const double* scale, long ns, const double model[4], double thr, int original,
                         uint32_t* n_inl, uint32_t* smask)
{
    const double tsq = (2.25 * thr) * thr;
    uint64_t cnt = 0;
    double val = 0.0;
    if (smask) memset(smask, 0, sizeof(uint32_t) * (size_t)((ns + 31) / 32));
    for (long i = 0; i < ns; i++)
    {
        const double r = gcr_o_scale_residual_variant(scale[3 * i], scale[3 * i + 1], scale[3 * i + 2], model, original);
        const double r2 = r * r;
        if (r2 <= tsq)
        {
            cnt++;
            val += -r2;
            if (smask) smask[i >> 5] |= 1u << (i & 31);
        }
    }
    if (cnt < 3)
    {
        *n_inl = 0;
        return 0.0;
    }
    *n_inl = (uint32_t)cnt;
    /* score.hpp:52-58 reset_value: sum <- sum - old + new with sum == old == val */
    const double msac = val / tsq + (double)cnt;
    double sum = val;
    sum -= val;
    sum += msac;
    return sum;
}

/* Hypothesis stage for replayed 3-samples: samples int32[H,4] = (s0, s1, s2, ignored).  Outputs as gcr_o_replay with
 * n_inl[h] = (count, 0) and masks of ceil(ns/32) words; every solved model is valid (solver_engine.h:80-90). */
int gcr_o_replay_scale(const double* scale, long ns, const int32_t* samples, long H, double thr, int original,
                       uint8_t* flags, double* models, double* score, uint32_t* n_inl, uint32_t* masks, int nthreads)
{
    const long ws = (ns + 31) / 32;
    int used = 1;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
    used = nthreads > 0 ? nthreads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 16)
#endif
    for (long h = 0; h < H; h++)
    {
        const double* s0 = scale + 3 * (long)samples[4 * h + 0];
        const double* s1 = scale + 3 * (long)samples[4 * h + 1];
        const double* s2 = scale + 3 * (long)samples[4 * h + 2];
        uint8_t f = 0;
        double m[4] = {0, 0, 0, 0}, val = 0.0;
        uint32_t cnt = 0;
        uint32_t* sm = masks ? masks + h * ws : NULL;
        if (masks) memset(masks + h * ws, 0, sizeof(uint32_t) * (size_t)ws);
        if (gcr_o_is_valid_sample_scale3(s0, s1, s2))
        {
            f |= GCR_O_SAMPLE_VALID;
            if (gcr_o_estimate_minimal_scale3(s0, s1, s2, original, m))
            {
                f |= GCR_O_SOLVED | GCR_O_MODEL_VALID;
                val = gcr_o_score_scale(scale, ns, m, thr, original, &cnt, sm);
            }
            else
            {
                m[0] = m[1] = m[2] = m[3] = 0.0;
            }
        }
        flags[h] = f;
        for (int k = 0; k < 4; k++) models[4 * h + k] = m[k];
        score[h] = val;
        n_inl[2 * h] = cnt;
        n_inl[2 * h + 1] = 0;
    }
    return used;
}

/* CPU twin of the device sampler for scale-only problems: ordered triple without replacement from Philox words
 * 0..2 (pair as draw_pair, third index uniform over the n-2 remaining values, mapped around the two taken). */
void gcr_o_sample_philox_scale3(uint64_t seed, uint64_t first_hyp, long H, long ns, int32_t* samples)
{
    for (long h = 0; h < H; h++)
    {
        const uint64_t g = first_hyp + (uint64_t)h;
        uint32_t c[4] = {(uint32_t)g, (uint32_t)(g >> 32), 0x47435231u, 0u};
        philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
        int32_t a, b;
        draw_pair(c[0], c[1], (uint32_t)ns, &a, &b);
        int32_t t = (int32_t)(((uint64_t)c[2] * (uint32_t)(ns - 2)) >> 32);
        const int32_t lo = a < b ? a : b, hi = a < b ? b : a;
        if (t >= lo) t++;
        if (t >= hi) t++;
        samples[4 * h + 0] = a;
        samples[4 * h + 1] = b;
        samples[4 * h + 2] = t;
        samples[4 * h + 3] = -1;
    }
}
