// Non-minimal model fits on the device: the least-squares step of the local optimisation and of the final refit
// (SURVEY.md 8f rows 1 and 3).  Replaces, for K index-list fits per call,
//   RectifyingHomographyEstimator::estimateModelNonminimal   rectifying_homography_estimator.h:164-227
//   RectifyingHomographyTwoSIFTSolver::estimateNonMinimalModel two_sift.hpp:423-579
//     rows  [x, y, s^(1/3)] . (h7, h8, alpha) = 1  per scale feature                            two_sift.hpp:228-238
//           [vp.x, vp.y, 0] . (h7, h8, alpha) = vp.z  per PAIR of orientation features, vp = l_i x l_j divided by its
//           largest |coordinate| when that exceeds 1                                              two_sift.hpp:240-260
//     solve  colPivHouseholderQr().solve                                                          two_sift.hpp:523-537
//     phi    0.5-degree weighted histogram mode of the rectified orientations                     two_sift.hpp:354-394,539-576
//   and the scale-only forms three_sift.hpp:195-254, three_sift_original.hpp:229-234.
//
// The pair rows are the cost: n(n-1)/2 of them (6.5 million at C3's 3.6k orientation inliers, 156 ms on one host
// core).  Here every row is produced and consumed in registers: the 3 x 3 Gram matrix and the right-hand side are
// accumulated in double-double arithmetic (106-bit significand), so that the normal equations lose nothing against
// the reference's QR on the dense matrix (error of the solution ~ cond^2 * 2^-104 instead of cond * 2^-53), and the
// 3 x 3 system is solved in double-double by a diagonally pivoted elimination (== column-pivoted QR in exact
// arithmetic, same rank decision).  Parity with the reference is at tolerance level (model <= 1e-6 relative), as for
// the host fit it replaces.
#pragma once
#include "gcr_exact.cuh"

namespace gcr
{

// phase stamps of the finishing CTA (tools/fit_stamps.cu only)
#ifdef GCR_FIT_STAMPS
__device__ unsigned long long g_fit_stamps[32]; // [i]: %globaltimer (ns), [16 + i]: clock64 (SM cycles)
__device__ int g_fit_passes = 1;                // the finish is run this many times (the stamps are those of the last pass)
// (branch-free: an `if (threadIdx.x == 0)` around the stores can leave lane 0 split from its warp, and every shuffle that follows then
// takes the compiler's WARPSYNC.COLLECTIVE fallback at ~250 cycles -- the stamps would measure their own side effect)
#define GCR_FIT_STAMP(i)                                                                                               \
    do {                                                                                                               \
        unsigned long long t_, c_ = static_cast<unsigned long long>(clock64());                                        \
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                                         \
        asm volatile("{ .reg .pred p; setp.eq.u32 p, %0, 0; @p st.global.u64 [%1], %2; @p st.global.u64 [%3], %4; }"   \
                     :: "r"(threadIdx.x), "l"(&g_fit_stamps[i]), "l"(t_), "l"(&g_fit_stamps[16 + (i)]), "l"(c_) : "memory"); \
    } while (0)
#else
#define GCR_FIT_STAMP(i) do { } while (0)
#endif

constexpr int kFitThreads = 256;
constexpr int kFitUnitRows = 8192;    // most rows of one fit that one CTA reduces (FitDesc::unit_rows: 1024 ... 8192)
constexpr int kFitUnitsAim = 256;     // a fit is cut into at most this many units if units of kFitUnitRows allow it (the solve sums them with one warp)
constexpr int kFitGram = 9;           // g00 g01 g02 g11 g12 g22 | r0 r1 r2
constexpr int kFitSmemLines = 1024;   // fits with at most this many orientation features build their lines in shared memory (24 KB)
constexpr int kPhiBins = 364;         // round(angle / 0.5 deg) for angle in [0, pi]: 0..360
constexpr double kPhiBinWidth = 0.5 * kPi / 180.0;
constexpr double kPhiFixedPoint = 1099511627776.0; // 2^40: the bin moments are exact integer sums of quantised angles

struct FitDesc
{
    long long s_off, o_off; // offsets of this fit's index lists in the concatenated scale / orientation lists
    int ns, no;             // list lengths
    unsigned int unit_first; // first row unit of this fit (prefix sum over fits of ceil(rows / unit_rows))
    unsigned int ent_off;    // first histogram entry of this fit in the output list
    int valid;               // 0: the reference's solver declines this fit (too few constraints)
    int unit_rows;           // rows per unit of this fit: small fits are cut finer (a 19 000-row refit is 19 CTAs instead of 3)
};

struct FitEntry // one non-empty histogram bin, emitted in order of first appearance in the index list
{
    int bin;
    unsigned int count;
    unsigned long long moment; // sum of llrint(angle * 2^40) over the bin
};

// ---- double-double arithmetic (error-free transformations; needs IEEE fma, no reassociation)
struct dd
{
    double hi, lo;
};
__device__ __forceinline__ dd dd_make(double a) { return dd{a, 0.0}; }
__device__ __forceinline__ dd two_sum(double a, double b)
{
    const double s = a + b, bb = s - a;
    return dd{s, (a - (s - bb)) + (b - bb)};
}
__device__ __forceinline__ dd quick_two_sum(double a, double b) // |a| >= |b|
{
    const double s = a + b;
    return dd{s, b - (s - a)};
}
__device__ __forceinline__ dd two_prod(double a, double b)
{
    const double p = a * b;
    return dd{p, __fma_rn(a, b, -p)};
}
__device__ __forceinline__ dd dd_add(dd a, dd b)
{
    dd s = two_sum(a.hi, b.hi);
    const dd t = two_sum(a.lo, b.lo);
    s.lo += t.hi;
    s = quick_two_sum(s.hi, s.lo);
    s.lo += t.lo;
    return quick_two_sum(s.hi, s.lo);
}
__device__ __forceinline__ dd dd_neg(dd a) { return dd{-a.hi, -a.lo}; }
__device__ __forceinline__ dd dd_sub(dd a, dd b) { return dd_add(a, dd_neg(b)); }
__device__ __forceinline__ dd dd_mul(dd a, dd b)
{
    dd p = two_prod(a.hi, b.hi);
    p.lo += a.hi * b.lo + a.lo * b.hi;
    return quick_two_sum(p.hi, p.lo);
}
// a / b: three quotient digits, each from the exact double-double residual of the ones before.  The digits need b.hi's reciprocal to a few
// ulps only (an error there moves weight into the next digit), so one division serves all three.
__device__ __forceinline__ dd dd_div(dd a, dd b)
{
    const double inv = 1.0 / b.hi;
    const double q1 = a.hi * inv;
    dd r = dd_sub(a, dd_mul(b, dd_make(q1)));
    const double q2 = r.hi * inv;
    r = dd_sub(r, dd_mul(b, dd_make(q2)));
    const double q3 = r.hi * inv;
    return dd_add(quick_two_sum(q1, q2), dd_make(q3));
}
// acc += a * b, exactly rounded product
__device__ __forceinline__ void dd_fma_acc(dd& acc, double a, double b)
{
    const dd p = two_prod(a, b);
    dd s = two_sum(acc.hi, p.hi);
    s.lo += acc.lo + p.lo;
    acc = quick_two_sum(s.hi, s.lo);
}

struct FitBuffers
{
    const double* scale;  // f64[ns,3] rows (x, y, s), as uploaded
    const double* orient; // f64[no,3] rows (x, y, theta)
    const FitDesc* desc;
    const unsigned int* idx_s;
    const unsigned int* idx_o;
    double* lines;   // f64[sum no_k, 3]: (sin, -cos, y cos - x sin) of every listed orientation feature; null: no fit of the call has more
                     // than kFitSmemLines of them and every row unit builds its fit's lines in shared memory (no k_fit_lines launch)
    unsigned int* done; // u32[K], zero between calls: row units of fit k that have stored their sums (the last one finishes the fit)
    double* partial; // f64[units, 2 * kFitGram]: per-unit Gram sums (hi, lo)
    double* models;  // f64[K,4] out: h7, h8, alpha, (phi: filled by the caller from the histogram entries)
    int* ok;         // i32[K] out
    unsigned int* n_ent; // u32[K] out
    FitEntry* ent;       // histogram entries
    // page-locked, device-mapped mirrors of the four outputs (null: the caller copies them back itself): small calls -- a local-optimisation
    // batch, the refit -- are chains of microsecond kernels, and a copy engine transfer at either end costs as much as two of them
    double* h_models;
    int* h_ok;
    unsigned int* h_n_ent;
    FitEntry* h_ent;
    int n_fits;
    unsigned int n_units;
    long long n_lines;
    int problem;
    // bucket-count schedule of the host's std::unordered_map<int, double> (probed once by the engine): the map rehashes to rehash_to[i]
    // buckets while inserting its rehash_at[i]-th key (libstdc++: 13 at 1, 29 at 14, 59 at 30, 127 at 60, 257 at 128, 541 at 258)
    int n_rehash;
    int rehash_at[8], rehash_to[8];
};
constexpr int kPhiMaxBuckets = 1024;

// lineFromPointAndAngle (math_utils.hpp:104-109) of every listed orientation feature
__global__ void __launch_bounds__(kFitThreads) k_fit_lines(FitBuffers F)
{
    pdl_launch_dependents(); // chained launch: the next kernel of the call may start launching now; it waits at its own top
    pdl_wait();
    const long long i = static_cast<long long>(blockIdx.x) * kFitThreads + threadIdx.x;
    if (i >= F.n_lines) return;
    const double* f = F.orient + 3 * static_cast<size_t>(F.idx_o[i]);
    double l[3];
    line_from_point_angle(f[0], f[1], f[2], l);
    F.lines[3 * i] = l[0];
    F.lines[3 * i + 1] = l[1];
    F.lines[3 * i + 2] = l[2];
}

// fit that owns row unit u: the last k with desc[k].unit_first <= u
__device__ __forceinline__ int fit_of_unit(const FitBuffers& F, unsigned int u)
{
    int lo = 0, hi = F.n_fits - 1;
    while (lo < hi)
    {
        const int mid = (lo + hi + 1) >> 1;
        if (F.desc[mid].unit_first <= u) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// Row unit -> partial Gram sums.  Rows of a fit: its scale features first, then its orientation pairs (a, b), a < b, in
// lexicographic order.  A thread walks rows r0 + tid, + kFitThreads, ...; consecutive threads take consecutive b.
// (`L`: the fit's lines, in global memory behind k_fit_lines or in the CTA's shared memory)
__device__ __forceinline__ void fit_rows_unit(const FitBuffers& F, unsigned int u, const FitDesc& D, const double* L)
{
    __shared__ double s_red[kFitThreads / 32][2 * kFitGram];
    const long long n = D.no;
    const long long rows = D.ns + (F.problem == 0 ? n * (n - 1) / 2 : 0);
    const long long r_begin = static_cast<long long>(u - D.unit_first) * D.unit_rows;
    const long long r_end = min(rows, r_begin + D.unit_rows);
    dd g[kFitGram];
#pragma unroll
    for (int i = 0; i < kFitGram; i++) g[i] = dd{0.0, 0.0};

    long long r = r_begin + threadIdx.x;
    // scale rows
    for (; r < min(r_end, static_cast<long long>(D.ns)); r += kFitThreads)
    {
        const double* f = F.scale + 3 * static_cast<size_t>(F.idx_s[D.s_off + r]);
        double a0 = f[0], a1 = f[1], a2, b;
        if (F.problem == 2) { a2 = -pow(f[2], -1.0 / 3.0); b = -1.0; } // three_sift_original.hpp:229-234
        else { a2 = pow(f[2], 1.0 / 3.0); b = 1.0; }
        dd_fma_acc(g[0], a0, a0); dd_fma_acc(g[1], a0, a1); dd_fma_acc(g[2], a0, a2);
        dd_fma_acc(g[3], a1, a1); dd_fma_acc(g[4], a1, a2); dd_fma_acc(g[5], a2, a2);
        dd_fma_acc(g[6], a0, b); dd_fma_acc(g[7], a1, b); dd_fma_acc(g[8], a2, b);
    }
    // pair rows
    if (r < r_end)
    {
        long long p = r - D.ns; // linear pair index; row a starts at S(a) = a (2n - a - 1) / 2 and holds n - 1 - a pairs
        const double tn = static_cast<double>(2 * n - 1);
        long long a = static_cast<long long>(floor(0.5 * (tn - sqrt(fmax(tn * tn - 8.0 * static_cast<double>(p), 0.0)))));
        a = max(0ll, min(a, n - 2));
        while (a + 1 <= n - 2 && (a + 1) * (2 * n - a - 2) / 2 <= p) a++;
        while (a > 0 && a * (2 * n - a - 1) / 2 > p) a--;
        long long o = p - a * (2 * n - a - 1) / 2; // offset inside row a: b = a + 1 + o
        double p0 = L[3 * a], p1 = L[3 * a + 1], p2 = L[3 * a + 2];
        for (; r < r_end; r += kFitThreads)
        {
            const double* q = L + 3 * (a + 1 + o);
            const double q0 = q[0], q1 = q[1], q2 = q[2];
            double v0 = p1 * q2 - p2 * q1, v1 = p2 * q0 - p0 * q2, v2 = p0 * q1 - p1 * q0; // Eigen cross, two_sift.hpp:249
            const double big = fmax(fabs(v0), fmax(fabs(v1), fabs(v2)));
            if (big > 1.0) { v0 /= big; v1 /= big; v2 /= big; }
            dd_fma_acc(g[0], v0, v0); dd_fma_acc(g[1], v0, v1); dd_fma_acc(g[3], v1, v1);
            dd_fma_acc(g[6], v0, v2); dd_fma_acc(g[7], v1, v2);
            o += kFitThreads;
            if (o >= n - 1 - a)
            {
                do { o -= n - 1 - a; a++; } while (o >= n - 1 - a && a < n - 2);
                if (a <= n - 2) { p0 = L[3 * a]; p1 = L[3 * a + 1]; p2 = L[3 * a + 2]; }
            }
        }
    }
    // CTA reduction in a fixed order: lanes by shuffle tree, then warps in index order
    __syncwarp(); // (lanes converged after the row loops: a split warp takes a ~250-cycle WARPSYNC fallback per shuffle, 90 of them here)
#pragma unroll
    for (int i = 0; i < kFitGram; i++)
    {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1)
        {
            dd other;
            other.hi = __shfl_down_sync(0xffffffffu, g[i].hi, off);
            other.lo = __shfl_down_sync(0xffffffffu, g[i].lo, off);
            g[i] = dd_add(g[i], other);
        }
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0)
    {
#pragma unroll
        for (int i = 0; i < kFitGram; i++)
        {
            s_red[warp][2 * i] = g[i].hi;
            s_red[warp][2 * i + 1] = g[i].lo;
        }
    }
    __syncthreads();
    if (threadIdx.x < kFitGram)
    {
        dd acc = dd{s_red[0][2 * threadIdx.x], s_red[0][2 * threadIdx.x + 1]};
        for (int w = 1; w < kFitThreads / 32; w++) acc = dd_add(acc, dd{s_red[w][2 * threadIdx.x], s_red[w][2 * threadIdx.x + 1]});
        F.partial[static_cast<size_t>(u) * 2 * kFitGram + 2 * threadIdx.x] = acc.hi;
        F.partial[static_cast<size_t>(u) * 2 * kFitGram + 2 * threadIdx.x + 1] = acc.lo;
        __threadfence(); // the sums are read by the CTA that stores the fit's last unit (k_fit_rows)
    }
}

// Solution of the 3 x 3 normal equations in double-double, with diagonal pivoting (the pivots are the squares of the
// diagonal of a column-pivoted QR of the row matrix, so the rank decision |R_kk| > 3 eps |R_00| of the reference's
// colPivHouseholderQr becomes pivot_k > 9 eps^2 pivot_0); free variables of a rank-deficient system are zero.
// By a whole warp: lane l < 12 holds element (l / 4, l % 4) of the augmented matrix [N | r] (the other lanes shadow lane 0), so an
// elimination step is one division, one product and one difference deep whatever the number of elements, and the code is a rolled loop of a
// few hundred instructions: 3 300 + 2 800 cycles for elimination and back-substitution (tools/fit_stamps.cu).  (One lane walking the same
// elimination as straight-line code was 2 900 dependent FP64 instructions at 8 cycles each.)
__device__ __noinline__ dd dd_div_call(dd a, dd b) { return dd_div(a, b); }
__device__ __forceinline__ dd shfl_dd(dd v, int src)
{
    __syncwarp(); // (see solve_normal3_warp: never reach a shuffle with the warp split)
    return dd{__shfl_sync(0xffffffffu, v.hi, src), __shfl_sync(0xffffffffu, v.lo, src)};
}
// `g`: the Gram sums (shared memory).  Every lane returns the same verdict and solution.
__device__ inline bool solve_normal3_warp(const dd* g, int lane, double x[3])
{
    const int e = lane < 12 ? lane : 0, i = e >> 2, j = e & 3;
    // [N | r] = [g0 g1 g2 | g6; g1 g3 g4 | g7; g2 g4 g5 | g8]
    const int t = j == 3 ? 6 + i : (i == j ? (i == 0 ? 0 : (i == 1 ? 3 : 5)) : (i + j == 1 ? 1 : (i + j == 2 ? 2 : 4)));
    dd v = g[t];
    // The lanes must be CONVERGED here (they are: the caller's __syncthreads precedes, and nothing below branches on the lane): a warp
    // that reaches a shuffle split -- e.g. after an `if (lane == 0)` block the compiler chose not to reconverge -- takes the fallback behind
    // BRA.DIV, one WARPSYNC.COLLECTIVE per shuffle at ~250 cycles each (seen with an earlier, branching form of the phase stamps:
    // 6 000 cycles per elimination step instead of 1 100).
    __syncwarp();
    int perm = 0 | (1 << 2) | (2 << 4); // perm[k] = (perm >> 2k) & 3
    int rank = 0;
    double pivot0 = 0.0;
    const double eps = 2.220446049250313e-16;
#pragma unroll 1
    for (int k = 0; k < 3; k++)
    {
#ifdef GCR_FIT_STAMPS
        {
            const unsigned long long am_ = __activemask();
            asm volatile("{ .reg .pred p; setp.eq.u32 p, %0, 0; @p st.global.u64 [%1], %2; }" :: "r"(lane), "l"(&g_fit_stamps[13 + (k > 0 ? 1 : 0)]), "l"(am_) : "memory");
        }
#endif
        __syncwarp();
        const double d0 = __shfl_sync(0xffffffffu, v.hi, 0), d1 = __shfl_sync(0xffffffffu, v.hi, 5), d2 = __shfl_sync(0xffffffffu, v.hi, 10);
        int best = k;
        double d = k == 0 ? d0 : (k == 1 ? d1 : d2); // the largest remaining diagonal entry, the first of equals
        if (k < 1 && d1 > d) { best = 1; d = d1; }
        if (k < 2 && d2 > d) { best = 2; d = d2; }
        if (best != k) // (uniform) rows k <-> best with their right-hand sides, columns k <-> best
        {
            const int si = i == k ? best : (i == best ? k : i);
            const int sj = j == 3 ? 3 : (j == k ? best : (j == best ? k : j));
            v = shfl_dd(v, 4 * si + sj);
            const int pk = (perm >> (2 * k)) & 3, pb = (perm >> (2 * best)) & 3;
            perm = (perm & ~((3 << (2 * k)) | (3 << (2 * best)))) | (pb << (2 * k)) | (pk << (2 * best));
        }
        if (k == 0) pivot0 = d;
        if (!(d > pivot0 * 9.0 * eps * eps) || !(d > 0.0)) break;
        rank++;
        const dd l = dd_div_call(shfl_dd(v, 4 * i + k), shfl_dd(v, 5 * k));
        const dd nv = dd_sub(v, dd_mul(l, shfl_dd(v, 4 * k + j)));
        if (i > k && j >= k) v = nv;
    }
    GCR_FIT_STAMP(12);
    dd y0{0.0, 0.0}, y1{0.0, 0.0}, y2{0.0, 0.0}; // (every lane, the same values)
    if (rank > 2) y2 = dd_div_call(shfl_dd(v, 11), shfl_dd(v, 10));
    if (rank > 1)
    {
        dd acc = shfl_dd(v, 7);
        if (rank > 2) acc = dd_sub(acc, dd_mul(shfl_dd(v, 6), y2));
        y1 = dd_div_call(acc, shfl_dd(v, 5));
    }
    if (rank > 0)
    {
        dd acc = shfl_dd(v, 3);
        if (rank > 1) acc = dd_sub(acc, dd_mul(shfl_dd(v, 1), y1));
        if (rank > 2) acc = dd_sub(acc, dd_mul(shfl_dd(v, 2), y2));
        y0 = dd_div_call(acc, shfl_dd(v, 0));
    }
    const double s0 = y0.hi + y0.lo, s1 = y1.hi + y1.lo, s2 = y2.hi + y2.lo;
    const int q0 = perm & 3, q1 = (perm >> 2) & 3; // x[perm[k]] = y[k]
#pragma unroll
    for (int c = 0; c < 3; c++) x[c] = q0 == c ? s0 : (q1 == c ? s1 : s2);
    return isfinite(x[0]) && isfinite(x[1]) && isfinite(x[2]);
}

// The units of a fit summed in a fixed order by the CTA, then the solve by warp 0.  Thread (group, component) = (tid / 9, tid % 9) sums the
// units u = group (mod kFitSumGroups) of one Gram component one after the other; thread `component` then sums the groups in index order.
// (One warp, a lane per 32 units and a 9 x 5-step shuffle butterfly in double-double, was 6 200 cycles for a single unit; this is 1 500.)
constexpr int kFitSumGroups = kFitThreads / kFitGram; // 28
__device__ __forceinline__ void fit_solve_block(const FitBuffers& F, int k)
{
    __shared__ dd s_part[kFitSumGroups][kFitGram];
    __shared__ dd s_sum[kFitGram];
    const FitDesc D = F.desc[k];
    const unsigned int u_end = k + 1 < F.n_fits ? F.desc[k + 1].unit_first : F.n_units;
    const unsigned int n_units = u_end - D.unit_first;
    const int comp = threadIdx.x % kFitGram, grp = threadIdx.x / kFitGram;
    if (grp < kFitSumGroups && static_cast<unsigned int>(grp) < n_units)
    {
        dd acc{0.0, 0.0};
        for (unsigned int u = D.unit_first + grp; u < u_end; u += kFitSumGroups)
        {
            // (L2 loads: other CTAs of this launch wrote these, and a unit's 144 bytes share cache lines with its neighbours')
            const double* P = F.partial + static_cast<size_t>(u) * 2 * kFitGram + 2 * comp;
            acc = dd_add(acc, dd{__ldcg(P), __ldcg(P + 1)});
        }
        s_part[grp][comp] = acc;
    }
    __syncthreads();
    GCR_FIT_STAMP(9);
    if (threadIdx.x < kFitGram)
    {
        dd acc = s_part[0][threadIdx.x];
        const int n_grp = static_cast<int>(min(n_units, static_cast<unsigned int>(kFitSumGroups)));
        for (int gi = 1; gi < n_grp; gi++) acc = dd_add(acc, s_part[gi][threadIdx.x]);
        s_sum[threadIdx.x] = acc;
    }
    __syncthreads();
    GCR_FIT_STAMP(10);
    if (threadIdx.x < 32)
    {
        double x[3] = {0.0, 0.0, 0.0};
        const bool solved = solve_normal3_warp(s_sum, threadIdx.x, x);
        if (threadIdx.x == 0)
        {
            bool ok = D.valid != 0 && solved;
            if (D.valid == 0) x[0] = x[1] = x[2] = 0.0;
            if (ok && x[2] < 1e-9) ok = false; // two_sift.hpp:533-537, three_sift.hpp:243-247
            F.models[4 * k] = x[0];
            F.models[4 * k + 1] = x[1];
            F.models[4 * k + 2] = x[2];
            F.models[4 * k + 3] = 0.0;
            F.ok[k] = ok ? 1 : 0;
            F.n_ent[k] = 0u;
        }
    }
    GCR_FIT_STAMP(11);
}

// Histogram of the rectified orientations of one fit (two_sift.hpp:539-576: angles folded into [0, pi], bins of
// 0.5 degree, unit weights).  The heaviest bin is chosen by the caller, which needs the bins in order of first
// appearance to resolve equal counts exactly as the reference's unordered_map does; emitted here in that order.
__device__ __forceinline__ void fit_phi_block(const FitBuffers& F, int k)
{
    static_assert(kPhiBins < 1024, "sort keys hold a position in 10 bits");
    __shared__ unsigned int s_first[kPhiBins], s_count[kPhiBins], s_sortkey[kPhiBins], s_bfirst[kPhiMaxBuckets];
    __shared__ unsigned long long s_moment[kPhiBins], s_best;
    __shared__ short s_key[kPhiBins], s_node[kPhiBins], s_ord[2][kPhiBins];
    __shared__ unsigned int s_total;
    const FitDesc D = F.desc[k];
    for (int b = threadIdx.x; b < kPhiBins; b += kFitThreads)
    {
        s_count[b] = 0u;
        s_first[b] = 0xffffffffu;
        s_moment[b] = 0ull;
    }
    if (threadIdx.x == 0)
    {
        s_total = 0u;
        s_best = 0ull;
    }
    __syncthreads();
    Model4 m{F.models[4 * k], F.models[4 * k + 1], F.models[4 * k + 2], 0.0};
    for (int a = threadIdx.x; a < D.no; a += kFitThreads)
    {
        const double* f = F.orient + 3 * static_cast<size_t>(F.idx_o[D.o_off + a]);
        double st, ct;
        sincos(f[2], &st, &ct);
        const double numer = (-f[0] * st + f[1] * ct) * m.h7 + st; // model.h:156-165
        const double denom = (f[0] * st - f[1] * ct) * m.h8 + ct;
        double t = clip_angle(atan2(numer, denom));
        if (t > kPi) t -= kPi; // theta and theta + pi are the same line
        int bin = static_cast<int>(round(t / kPhiBinWidth));
        if (!(bin >= 0)) bin = 0; // NaN angle
        if (bin >= kPhiBins) bin = kPhiBins - 1;
        atomicAdd(&s_count[bin], 1u);
        atomicMin(&s_first[bin], static_cast<unsigned int>(a));
        atomicAdd(&s_moment[bin], static_cast<unsigned long long>(llrint(t * kPhiFixedPoint)));
    }
    __syncthreads();
    GCR_FIT_STAMP(5);
    // the non-empty bins, in any order, with their first positions side by side: the rank of a bin among them (its place in the
    // order of first appearance) is a count over that list, not over all 364 bins
    unsigned int n_bins = 0;
    for (int b = threadIdx.x; b < kPhiBins; b += kFitThreads)
        if (s_count[b] != 0u)
        {
            const unsigned int at = atomicAdd(&s_total, 1u);
            s_node[at] = static_cast<short>(b);
            s_sortkey[at] = s_first[b];
        }
    __syncthreads();
    const unsigned int n_list = s_total;
    for (unsigned int q = threadIdx.x; q < n_list; q += kFitThreads)
    {
        const int b = s_node[q];
        const unsigned int mine = s_sortkey[q];
        unsigned int rank = 0;
        for (unsigned int c = 0; c < n_list; c++) rank += s_sortkey[c] < mine ? 1u : 0u;
        FitEntry e;
        e.bin = b;
        e.count = s_count[b];
        e.moment = s_moment[b];
        F.ent[D.ent_off + rank] = e;
        s_key[rank] = static_cast<short>(b);
        n_bins++;
    }
    if (n_bins) atomicAdd(&F.n_ent[k], n_bins);
    // phi on the device too (so that the fitted models can be scored without a round trip: gcr_fit_score_models): the mean angle of the
    // heaviest bin, and among equally heavy bins the one the REFERENCE meets first (two_sift.hpp:354-394: it walks an
    // std::unordered_map<int, double> filled in order of first appearance and keeps the first strictly heavier entry).  Small index lists
    // (14 orientations of a local-optimisation trial) have equally heavy bins as a rule, so the map's iteration order is restated here.
    // libstdc++ keeps one node list: a new key whose bucket is empty goes to the FRONT of the list, otherwise to the front of its
    // bucket's run of nodes; a rehash re-links the nodes, walked in list order, by the same two rules (hashtable.h:
    // _M_insert_bucket_begin, _M_rehash_aux).  So between two rehashes the list is a function of the sequence Q = (list at the rehash)
    // ++ (keys inserted since): runs by bucket, the run created LAST first, and inside a run the node met last first -- that is Q sorted by
    // (position in Q of the bucket's first node, own position in Q), both descending.  One such sort per bucket count the host's own
    // map goes through (FitBuffers::rehash_*, at most six), each a rank count over at most 364 keys by the whole CTA.  (Walking the
    // list with one thread, as the library does, is a chain of dependent shared-memory accesses: 6.7 us for 14 keys and 48 us for 150
    // against 2.6 and 6 us this way, tools/fit_stamps.cu.)  The caller still
    // replays the real map on the entries above and re-scores where the two disagree, so a library with another policy costs time,
    // not correctness.
    __syncthreads(); // (s_key complete)
    GCR_FIT_STAMP(6);
    const int n = static_cast<int>(n_list);
    int cur = 0;
    for (int r = -1; r < F.n_rehash; r++) // (block-uniform) r = -1: the single bucket of an empty map, if keys are inserted under it
    {
        const int first_new = r < 0 ? 0 : F.rehash_at[r] - 1; // nodes [0, first_new) are in the list, in s_ord[cur] order
        if (first_new >= n) break;
        const int nb = r < 0 ? 1 : min(F.rehash_to[r], kPhiMaxBuckets);
        const int m = min(n, r + 1 < F.n_rehash ? F.rehash_at[r + 1] - 1 : n);
        if (m <= first_new) continue;
        for (int b = threadIdx.x; b < nb; b += kFitThreads) s_bfirst[b] = 0xffffffffu;
        __syncthreads();
        for (int e = threadIdx.x; e < m; e += kFitThreads)
        {
            const short node = e < first_new ? s_ord[cur][e] : static_cast<short>(e);
            s_node[e] = node;
            atomicMin(&s_bfirst[s_key[node] % nb], static_cast<unsigned int>(e));
        }
        __syncthreads();
        for (int e = threadIdx.x; e < m; e += kFitThreads) s_sortkey[e] = (s_bfirst[s_key[s_node[e]] % nb] << 10) | static_cast<unsigned int>(e);
        __syncthreads();
        for (int e = threadIdx.x; e < m; e += kFitThreads)
        {
            const unsigned int mine = s_sortkey[e];
            int pos = 0;
            for (int c = 0; c < m; c++) pos += s_sortkey[c] > mine ? 1 : 0;
            s_ord[cur ^ 1][pos] = s_node[e];
        }
        __syncthreads();
        cur ^= 1;
    }
    // the first strictly heavier entry of the walk == the heaviest bin, the earliest in list order among equals
    for (int pos = threadIdx.x; pos < n; pos += kFitThreads)
        atomicMax(&s_best, (static_cast<unsigned long long>(s_count[s_key[s_ord[cur][pos]]]) << 10) | static_cast<unsigned long long>(1023 - pos));
    __syncthreads();
    if (threadIdx.x == 0 && n > 0)
    {
        const int mode = s_key[s_ord[cur][1023 - static_cast<int>(s_best & 1023ull)]];
        F.models[4 * k + 3] = static_cast<double>(s_moment[mode]) / kPhiFixedPoint / static_cast<double>(s_best >> 10);
    }
}

// Candidate slots of the scoring pass that follows a fit in gcr_fit_score_models (what k_set_models does for caller-supplied models).
struct FitSlots
{
    Model4* model; // null: the fits are not scored in this call
    unsigned int* hyp;
    unsigned long long* qsum;
    unsigned int* rawcnt;
};

// What follows a fit's row sums, by one CTA: the solve (warp 0), the histogram mode of the rectified orientations (hybrid problem,
// accepted fits), the fitted model's candidate slot with cleared accumulators, the page-locked mirrors of the outputs.
#ifdef GCR_FIT_STAMPS
__device__ __noinline__ void fit_finish_block(const FitBuffers& F, const FitSlots& S, int k) // (one copy of the code for both passes)
#else
__device__ __forceinline__ void fit_finish_block(const FitBuffers& F, const FitSlots& S, int k)
#endif
{
    fit_solve_block(F, k);
    __syncthreads(); // the model and its verdict, written by thread 0 to global memory, are read by the whole CTA below
    GCR_FIT_STAMP(4);
    if (F.problem == 0 && F.ok[k]) fit_phi_block(F, k); // (block-uniform)
    GCR_FIT_STAMP(7);
    if (S.model != nullptr && threadIdx.x == 0)
    {
        S.model[k] = Model4{F.models[4 * k], F.models[4 * k + 1], F.models[4 * k + 2], F.models[4 * k + 3]};
        S.hyp[k] = static_cast<unsigned int>(k);
        S.qsum[2 * k] = S.qsum[2 * k + 1] = 0ull;
        S.rawcnt[2 * k] = S.rawcnt[2 * k + 1] = 0u;
    }
    if (F.h_models != nullptr)
    {
        // (entries and their count: written before the barriers of fit_phi_block; the model and its verdict: by this thread 0)
        const unsigned int n_ent = __ldcg(&F.n_ent[k]); // (accumulated by atomics of this CTA)
        const unsigned int e0 = F.desc[k].ent_off;
        for (unsigned int i = threadIdx.x; i < n_ent; i += kFitThreads) F.h_ent[e0 + i] = F.ent[e0 + i];
        if (threadIdx.x == 0)
        {
#pragma unroll
            for (int c = 0; c < 4; c++) F.h_models[4 * k + c] = F.models[4 * k + c];
            F.h_ok[k] = F.ok[k];
            F.h_n_ent[k] = n_ent;
        }
    }
}

// One launch per call of gcr_fit_models / gcr_fit_score_models for everything up to the fitted models: a CTA per row unit; the CTA that
// stores the LAST unit of a fit (a counter per fit, left at zero again) finishes that fit -- units summed in index order whichever CTA
// that is, so the result does not depend on the schedule.  Before: lines | rows | solve | phi | candidate slots as five dependent
// launches of a few microseconds each, which is what the local optimisation and the refit of a small scene consisted of.
__global__ void __launch_bounds__(kFitThreads) k_fit_rows(FitBuffers F, FitSlots S)
{
    pdl_launch_dependents(); // chained launch: the next kernel of the call may start launching now; it waits at its own top
    pdl_wait();
    __shared__ double s_lines[3 * kFitSmemLines];
    __shared__ int s_last;
    GCR_FIT_STAMP(0);
    const unsigned int u = blockIdx.x;
    const int k = fit_of_unit(F, u);
    const FitDesc D = F.desc[k];
    const double* L = F.lines != nullptr ? F.lines + 3 * D.o_off : s_lines;
    const long long r_end = static_cast<long long>(u - D.unit_first + 1) * D.unit_rows;
    if (F.lines == nullptr && F.problem == 0 && r_end > D.ns) // (block-uniform) this unit holds pair rows
    {
        for (int a = threadIdx.x; a < D.no; a += kFitThreads)
        {
            const double* f = F.orient + 3 * static_cast<size_t>(F.idx_o[D.o_off + a]);
            line_from_point_angle(f[0], f[1], f[2], s_lines + 3 * a);
        }
        __syncthreads();
    }
    GCR_FIT_STAMP(1);
    fit_rows_unit(F, u, D, L);
    __syncthreads(); // (the writers fenced their sums)
    GCR_FIT_STAMP(2);
    if (threadIdx.x == 0)
    {
        const unsigned int u_end = k + 1 < F.n_fits ? F.desc[k + 1].unit_first : F.n_units;
        s_last = atomicAdd(&F.done[k], 1u) + 1u == u_end - D.unit_first ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (threadIdx.x == 0) F.done[k] = 0u;
    GCR_FIT_STAMP(3);
#ifdef GCR_FIT_STAMPS
    for (int pass = 1; pass < g_fit_passes; pass++)
    {
        fit_finish_block(F, S, k);
        __syncthreads();
        GCR_FIT_STAMP(3);
    }
#endif
    fit_finish_block(F, S, k);
    GCR_FIT_STAMP(8);
}

} // namespace gcr
