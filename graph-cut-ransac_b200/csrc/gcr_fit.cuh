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

constexpr int kFitThreads = 256;
constexpr int kFitUnitRows = 8192;    // rows of one fit that one CTA reduces
constexpr int kFitGram = 9;           // g00 g01 g02 g11 g12 g22 | r0 r1 r2
constexpr int kPhiBins = 364;         // round(angle / 0.5 deg) for angle in [0, pi]: 0..360
constexpr double kPhiBinWidth = 0.5 * kPi / 180.0;
constexpr double kPhiFixedPoint = 1099511627776.0; // 2^40: the bin moments are exact integer sums of quantised angles

struct FitDesc
{
    long long s_off, o_off; // offsets of this fit's index lists in the concatenated scale / orientation lists
    int ns, no;             // list lengths
    unsigned int unit_first; // first row unit of this fit (prefix sum over fits of ceil(rows / kFitUnitRows))
    unsigned int ent_off;    // first histogram entry of this fit in the output list
    int valid;               // 0: the reference's solver declines this fit (too few constraints)
    int pad_;
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
__device__ __forceinline__ dd dd_div(dd a, dd b)
{
    const double q1 = a.hi / b.hi;
    dd r = dd_sub(a, dd_mul(b, dd_make(q1)));
    const double q2 = r.hi / b.hi;
    r = dd_sub(r, dd_mul(b, dd_make(q2)));
    const double q3 = r.hi / b.hi;
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
    double* lines;   // f64[sum no_k, 3]: (sin, -cos, y cos - x sin) of every listed orientation feature
    double* partial; // f64[units, 2 * kFitGram]: per-unit Gram sums (hi, lo)
    double* models;  // f64[K,4] out: h7, h8, alpha, (phi: filled by the caller from the histogram entries)
    int* ok;         // i32[K] out
    unsigned int* n_ent; // u32[K] out
    FitEntry* ent;       // histogram entries
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
__global__ void __launch_bounds__(kFitThreads) k_fit_rows(FitBuffers F)
{
    pdl_launch_dependents(); // chained launch: the next kernel of the call may start launching now; it waits at its own top
    pdl_wait();
    __shared__ double s_red[kFitThreads / 32][2 * kFitGram];
    const unsigned int u = blockIdx.x;
    const int k = fit_of_unit(F, u);
    const FitDesc D = F.desc[k];
    const long long n = D.no;
    const long long rows = D.ns + (F.problem == 0 ? n * (n - 1) / 2 : 0);
    const long long r_begin = static_cast<long long>(u - D.unit_first) * kFitUnitRows;
    const long long r_end = min(rows, r_begin + kFitUnitRows);
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
        const double* L = F.lines + 3 * D.o_off;
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
    }
}

// Solution of the 3 x 3 normal equations in double-double, with diagonal pivoting (the pivots are the squares of the
// diagonal of a column-pivoted QR of the row matrix, so the rank decision |R_kk| > 3 eps |R_00| of the reference's
// colPivHouseholderQr becomes pivot_k > 9 eps^2 pivot_0); free variables of a rank-deficient system are zero.
__device__ inline bool solve_normal3(const dd g[kFitGram], double x[3])
{
    dd N[3][4];
    N[0][0] = g[0]; N[0][1] = g[1]; N[0][2] = g[2];
    N[1][0] = g[1]; N[1][1] = g[3]; N[1][2] = g[4];
    N[2][0] = g[2]; N[2][1] = g[4]; N[2][2] = g[5];
    N[0][3] = g[6]; N[1][3] = g[7]; N[2][3] = g[8];
    int perm[3] = {0, 1, 2};
    int rank = 0;
    double pivot0 = 0.0;
    const double eps = 2.220446049250313e-16;
    for (int k = 0; k < 3; k++)
    {
        int best = k;
        for (int j = k + 1; j < 3; j++)
            if (N[j][j].hi > N[best][best].hi) best = j;
        if (best != k)
        {
            for (int j = 0; j < 4; j++) { const dd t = N[k][j]; N[k][j] = N[best][j]; N[best][j] = t; }
            for (int i = 0; i < 3; i++) { const dd t = N[i][k]; N[i][k] = N[i][best]; N[i][best] = t; }
            const int t = perm[k]; perm[k] = perm[best]; perm[best] = t;
        }
        const double d = N[k][k].hi;
        if (k == 0) pivot0 = d;
        if (!(d > pivot0 * 9.0 * eps * eps) || !(d > 0.0)) break;
        rank++;
        for (int i = k + 1; i < 3; i++)
        {
            const dd l = dd_div(N[i][k], N[k][k]);
            for (int j = k; j < 4; j++) N[i][j] = dd_sub(N[i][j], dd_mul(l, N[k][j]));
        }
    }
    dd y[3] = {dd{0, 0}, dd{0, 0}, dd{0, 0}};
    for (int i = rank - 1; i >= 0; i--)
    {
        dd acc = N[i][3];
        for (int j = i + 1; j < rank; j++) acc = dd_sub(acc, dd_mul(N[i][j], y[j]));
        y[i] = dd_div(acc, N[i][i]);
    }
    for (int k = 0; k < 3; k++) x[perm[k]] = y[k].hi + y[k].lo;
    return isfinite(x[0]) && isfinite(x[1]) && isfinite(x[2]);
}

// One warp per fit: units of the fit summed in a fixed order, then the solve.
__global__ void __launch_bounds__(32) k_fit_solve(FitBuffers F)
{
    pdl_launch_dependents(); // chained launch: the next kernel of the call may start launching now; it waits at its own top
    pdl_wait();
    const int k = blockIdx.x, lane = threadIdx.x;
    const FitDesc D = F.desc[k];
    const unsigned int u_end = k + 1 < F.n_fits ? F.desc[k + 1].unit_first : F.n_units;
    dd g[kFitGram];
#pragma unroll
    for (int i = 0; i < kFitGram; i++) g[i] = dd{0.0, 0.0};
    for (unsigned int u = D.unit_first + lane; u < u_end; u += 32)
    {
        const double* P = F.partial + static_cast<size_t>(u) * 2 * kFitGram;
#pragma unroll
        for (int i = 0; i < kFitGram; i++) g[i] = dd_add(g[i], dd{P[2 * i], P[2 * i + 1]});
    }
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
    if (lane == 0)
    {
        double x[3] = {0.0, 0.0, 0.0};
        bool ok = D.valid != 0 && solve_normal3(g, x);
        if (ok && x[2] < 1e-9) ok = false; // two_sift.hpp:533-537, three_sift.hpp:243-247
        F.models[4 * k] = x[0];
        F.models[4 * k + 1] = x[1];
        F.models[4 * k + 2] = x[2];
        F.models[4 * k + 3] = 0.0;
        F.ok[k] = ok ? 1 : 0;
        F.n_ent[k] = 0u;
    }
}

// Histogram of the rectified orientations of one fit (two_sift.hpp:539-576: angles folded into [0, pi], bins of
// 0.5 degree, unit weights).  The heaviest bin is chosen by the caller, which needs the bins in order of first
// appearance to resolve equal counts exactly as the reference's unordered_map does; emitted here in that order.
__global__ void __launch_bounds__(kFitThreads) k_fit_phi(FitBuffers F)
{
    pdl_launch_dependents(); // chained launch: the next kernel of the call may start launching now; it waits at its own top
    pdl_wait();
    __shared__ unsigned int s_count[kPhiBins], s_first[kPhiBins];
    __shared__ unsigned long long s_moment[kPhiBins];
    __shared__ short s_key[kPhiBins], s_next[kPhiBins + 1], s_bucket[kPhiMaxBuckets];
    __shared__ unsigned int s_total;
    const int k = blockIdx.x;
    if (!F.ok[k]) return;
    const FitDesc D = F.desc[k];
    for (int b = threadIdx.x; b < kPhiBins; b += kFitThreads)
    {
        s_count[b] = 0u;
        s_first[b] = 0xffffffffu;
        s_moment[b] = 0ull;
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
    unsigned int n_bins = 0;
    for (int b = threadIdx.x; b < kPhiBins; b += kFitThreads)
    {
        if (s_count[b] == 0u) continue;
        unsigned int rank = 0;
        for (int c = 0; c < kPhiBins; c++) rank += (s_first[c] < s_first[b]) ? 1u : 0u;
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
    // (14 orientations of a local-optimisation trial) have equally heavy bins as a rule, so the map's iteration order is restated here:
    // libstdc++'s node list -- a new key whose bucket is empty goes to the FRONT of the list, otherwise behind its bucket's first node;
    // a rehash re-links the nodes in list order by the same rule (hashtable.h: _M_insert_bucket_begin, _M_rehash_aux) -- with the bucket
    // counts the host's own map goes through (FitBuffers::rehash_*).  The caller still replays the real map on the entries above and
    // re-scores where the two disagree, so a library with another policy costs time, not correctness.
    if (threadIdx.x == 0) s_total = 0u;
    __syncthreads();
    if (n_bins) atomicAdd(&s_total, n_bins);
    __syncthreads();
    if (threadIdx.x == 0)
    {
        constexpr short kNone = -2, kBB = kPhiBins; // no node / the list's before-begin sentinel
        const int n = static_cast<int>(s_total);
        int nb = 1, ri = 0;
        s_next[kBB] = -1;
        s_bucket[0] = kNone;
        for (int i = 0; i < n; i++)
        {
            if (ri < F.n_rehash && i + 1 == F.rehash_at[ri])
            {
                nb = min(F.rehash_to[ri++], kPhiMaxBuckets);
                for (int b = 0; b < nb; b++) s_bucket[b] = kNone;
                short p = s_next[kBB];
                s_next[kBB] = -1;
                int bbegin_bkt = 0;
                while (p >= 0)
                {
                    const short nx = s_next[p];
                    const int b = s_key[p] % nb;
                    if (s_bucket[b] == kNone)
                    {
                        s_next[p] = s_next[kBB];
                        s_next[kBB] = p;
                        s_bucket[b] = kBB;
                        if (s_next[p] >= 0) s_bucket[bbegin_bkt] = p;
                        bbegin_bkt = b;
                    }
                    else
                    {
                        s_next[p] = s_next[s_bucket[b]];
                        s_next[s_bucket[b]] = p;
                    }
                    p = nx;
                }
            }
            const int b = s_key[i] % nb;
            if (s_bucket[b] != kNone)
            {
                s_next[i] = s_next[s_bucket[b]];
                s_next[s_bucket[b]] = static_cast<short>(i);
            }
            else
            {
                s_next[i] = s_next[kBB];
                s_next[kBB] = static_cast<short>(i);
                if (s_next[i] >= 0) s_bucket[s_key[s_next[i]] % nb] = static_cast<short>(i);
                s_bucket[b] = kBB;
            }
        }
        unsigned int best = 0u;
        int mode = -1;
        for (short p = s_next[kBB]; p >= 0; p = s_next[p])
            if (s_count[s_key[p]] > best)
            {
                best = s_count[s_key[p]];
                mode = s_key[p];
            }
        if (mode >= 0) F.models[4 * k + 3] = static_cast<double>(s_moment[mode]) / kPhiFixedPoint / static_cast<double>(best);
    }
}

} // namespace gcr
