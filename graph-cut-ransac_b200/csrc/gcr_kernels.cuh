// Device kernels of the B200 hybrid-RANSAC hypothesis engine (sm_100a).
//
//   k_prep_features   per-feature invariants -> FP32 chunked structure-of-arrays (+ scene maxima, irregular list)
//   k_sample_solve    Philox / replayed minimal sample -> degeneracy test -> FP64 minimal solver -> FP32
//                     hypothesis parameters with their proven error bands -> compaction of solved hypotheses
//   k_score_fast      THE hot kernel: multi-class MSAC over every feature of every class, FP32, features
//                     streamed through shared memory by TMA bulk copies (cp.async.bulk + mbarrier ring),
//                     one thread <-> HB hypotheses held in registers, guard-band pairs appended to a worklist
//   k_fixup           FP64 reference-order re-evaluation of guard-band pairs and of irregular features
//   k_exact_items     FP64 reference-order scoring of whole hypotheses (out-of-range ones, near-best candidates,
//                     caller-supplied models), one CTA per hypothesis, ballot-packed masks
//   k_finalize        MSAC value per hypothesis (S2/S3) + arg-max key
//   k_collect         near-best candidates for exact re-scoring
//   k_select          final selection rule (S10/S11) + winner record
//
// Reference semantics implemented are cited at each function.
#pragma once
#include "gcr_kernels_common.cuh"
#include "gcr_score.cuh"
#include <cooperative_groups.h>

namespace gcr
{

// ------------------------------------------------------------------------------------------------
// K0: feature preparation.  Hoists what the reference recomputes per (model, feature):
//   scale       (x, y, s)     -> (x/q, y/q, 1/q), q = cbrt(s): then  s (1-h7x-h8y)^-3 = v^-3 with
//                                v = 1/q - h7 x/q - h8 y/q          (two_sift.hpp:621-645, model.h:194-204)
//   orientation (x, y, theta) -> (k = x sin - y cos, sin, cos): numer = sin - k h7, denom = cos + k h8
//                                                                  (model.h:156-165)
// Features whose prepared values are non-finite or absurdly large are "irregular": they are kept out of
// the FP32 arrays (sentinel = never an inlier) and always evaluated in FP64 (k_fixup).
__global__ void k_prep_features(const double* __restrict__ scale, int ns, const double* __restrict__ orient, int no,
                                float* __restrict__ fs, float* __restrict__ fo, int ncs, int nco,
                                double* __restrict__ fsd, double* __restrict__ fod, int ncsd, int ncod, Scene* scene,
                                unsigned int* irr_s, unsigned int* irr_o, double* __restrict__ aux_s, double* __restrict__ aux_o,
                                int problem, double* __restrict__ scale_copy, double* __restrict__ orient_copy)
{
    constexpr int kCD = kChunk / 2; // features per FP64 chunk (same byte size as an FP32 chunk)
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    pdl_launch_dependents(); // a following k_sample_solve may fetch its samples (replay: over PCIe) while this kernel runs
    const float inf = __int_as_float(0x7f800000);
    const double dinf = __longlong_as_double(0x7ff0000000000000ll), dnan = __longlong_as_double(0x7ff8000000000000ll);
    if (i < ncs * kChunk) // (whole warps: the bound is a multiple of the block size)
    {
        float a = 0.f, b = 0.f, c = inf; // pad: v = +inf -> never an inlier, never in the band
        double ad = 0.0, bd = 0.0, cd = dinf;
        double xa = 0.0, ya = 0.0, ia = 0.0;
        bool ok = false;
        if (i < ns)
        {
            const double x = scale[3 * i], y = scale[3 * i + 1], s = scale[3 * i + 2];
            if (scale_copy) // the rows were read in place from pinned host memory: keep the device copy the other kernels use
            {
                scale_copy[3 * i] = x;
                scale_copy[3 * i + 1] = y;
                scale_copy[3 * i + 2] = s;
            }
            // the minimal solvers' third column (two_sift.hpp:288-299, three_sift.hpp:153-160, three_sift_original.hpp):
            // a per-FEATURE value, evaluated here once with the very function the solver would call per hypothesis
            aux_s[i] = problem == 2 ? pow(s, -1.0 / 3.0) : pow(s, 1.0 / 3.0);
            const double q = cbrt(s);
            xa = x / q; ya = y / q; ia = 1.0 / q;
            ok = isfinite(xa) && isfinite(ya) && isfinite(ia) && fabs(xa) < 1e12 && fabs(ya) < 1e12 && fabs(ia) < 1e12 && fabs(ia) > 1e-12;
        }
        // Features leave the FP32 / FP64 planes in aligned PAIRS (2j, 2j + 1): an irregular feature takes its partner along, and so does the
        // padding behind the last feature of an odd count.  The winner-only scoring loop shares one reciprocal between the two features of a
        // pair (score_chunk_w): a NaN sentinel must never meet a real feature there.
        const bool pair_ok = ok && __shfl_xor_sync(0xffffffffu, ok, 1);
        if (i < ns)
        {
            if (pair_ok)
            {
                a = static_cast<float>(xa);
                b = static_cast<float>(ya);
                c = static_cast<float>(ia);
                ad = xa; bd = ya; cd = ia;
            }
            else
            {
                const unsigned int slot = atomicAdd(&scene->n_irr_s, 1u);
                if (slot < kIrrCap) irr_s[slot] = i; else scene->irr_overflow = 1;
            }
        }
        {
            // scene maxima: one atomic per warp and quantity (the bit pattern of a non-negative float orders like the float).  Measured
            // neutral even at 7 M features (the hardware already merges a warp's same-address atomics): the kernel's time is its FP64
            // pow / cbrt / sincos.
            const bool real = i < ns && pair_ok;
            const int m0 = __reduce_max_sync(0xffffffffu, real ? __float_as_int(fabsf(a)) : 0);
            const int m1 = __reduce_max_sync(0xffffffffu, real ? __float_as_int(fabsf(b)) : 0);
            const int m2 = __reduce_max_sync(0xffffffffu, real ? __float_as_int(fabsf(c)) : 0);
            if ((threadIdx.x & 31) == 0 && (m0 | m1 | m2) != 0)
            {
                atomicMax(reinterpret_cast<int*>(&scene->xq_max), m0);
                atomicMax(reinterpret_cast<int*>(&scene->yq_max), m1);
                atomicMax(reinterpret_cast<int*>(&scene->iq_max), m2);
            }
        }
        float* base = fs + (i / kChunk) * kChunkFloats + (i % kChunk);
        base[0] = a;
        base[kChunk] = b;
        base[2 * kChunk] = c;
        if (i < ncsd * kCD)
        {
            double* bd_ = fsd + static_cast<size_t>(i / kCD) * (3 * kCD) + (i % kCD);
            bd_[0] = ad;
            bd_[kCD] = bd;
            bd_[2 * kCD] = cd;
        }
    }
    if (i < nco * kChunk)
    {
        float a = 0.f, b = 0.f, c = 0.f;
        double ad = 0.0, bd = dnan, cd = 0.0; // pad: NaN sine
        bool pad = true, ok = false;
        double k = 0.0, st = 0.0, ct = 0.0;
        if (i < no)
        {
            const double x = orient[3 * i], y = orient[3 * i + 1], t = orient[3 * i + 2];
            if (orient_copy)
            {
                orient_copy[3 * i] = x;
                orient_copy[3 * i + 1] = y;
                orient_copy[3 * i + 2] = t;
            }
            sincos(t, &st, &ct);
            aux_o[2 * i] = st; // lineFromPointAndAngle's sine / cosine (math_utils.hpp:104-109), per feature
            aux_o[2 * i + 1] = ct;
            k = x * st - y * ct;
            ok = isfinite(k) && isfinite(st) && isfinite(ct) && fabs(k) < 1e12;
        }
        const bool pair_ok = ok && __shfl_xor_sync(0xffffffffu, ok, 1); // aligned pairs, as for the scale class
        if (i < no)
        {
            if (pair_ok)
            {
                a = static_cast<float>(k);
                b = static_cast<float>(st);
                c = static_cast<float>(ct);
                ad = k; bd = st; cd = ct;
                pad = false;
            }
            else
            {
                const unsigned int slot = atomicAdd(&scene->n_irr_o, 1u);
                if (slot < kIrrCap) irr_o[slot] = i; else scene->irr_overflow = 1;
            }
        }
        {
            const int mk = __reduce_max_sync(0xffffffffu, pad ? 0 : __float_as_int(fabsf(a))); // (one atomic per warp, as for the scale class)
            if ((threadIdx.x & 31) == 0 && mk != 0) atomicMax(reinterpret_cast<int*>(&scene->k_max), mk);
        }
        if (pad)
        {
            // NaN sine: every comparison in the score kernel is false -> certain outlier, never in the band
            a = 0.f;
            b = __int_as_float(0x7fc00000);
            c = 0.f;
        }
        float* base = fo + (i / kChunk) * kChunkFloats + (i % kChunk);
        base[0] = a;
        base[kChunk] = b;
        base[2 * kChunk] = c;
        if (i < ncod * kCD)
        {
            double* bd_ = fod + static_cast<size_t>(i / kCD) * (3 * kCD) + (i % kCD);
            bd_[0] = ad;
            bd_[kCD] = bd;
            bd_[2 * kCD] = cd;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K1+K2: sampler + degeneracy test + minimal solver + FP32 parameter derivation + compaction.

// Philox4x32-10 (Salmon et al., SC'11).  Counter = (global hypothesis index, "GCR1", 0), key = seed.
// Replaces get_random_subset (GCRANSAC.h:53-80): a uniform ORDERED pair without replacement per class.
__device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1)
{
#pragma unroll
    for (int r = 0; r < 10; r++)
    {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
        const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
        c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
}
__device__ __forceinline__ void draw_pair(uint32_t u0, uint32_t u1, uint32_t n, int& a, int& b)
{
    const uint32_t i0 = __umulhi(u0, n);
    uint32_t i1 = __umulhi(u1, n - 1);
    if (i1 >= i0) i1++;
    a = static_cast<int>(i0);
    b = static_cast<int>(i1);
}

// Error bands of the FP32 path for one model: |delta (A - lg2 v)| <= drho / 2 and |delta t| <= dprime / max + t_band0 in the
// FP32 tiles (v, numer, denom from FP32 features and FP32 h7, h8), and the same with the three cancelling FMAs done in FP64 and rounded
// once (precise tiles, and the mid-precision refinement of the contenders in k_tail).
struct PathBands
{
    double A;                 // log2 of the (effective) alpha
    double drho32, dprime32;  // FP32 tiles
    double drho64, dprime64;  // FP64 front end
    bool ok32, ok64, common_ok;
    bool ok32_bound; // the FP32 bands are VALID (finite, v stays positive over the inlier range) though possibly wide
};
__device__ __forceinline__ PathBands path_bands(const Model4& m, const Scene& sc, const Thresholds& T, bool all_exact)
{
    PathBands r;
    const double ah7 = fabs(m.h7), ah8 = fabs(m.h8);
    const double mag = ah7 * sc.xq_max + ah8 * sc.yq_max + sc.iq_max; // bound of the terms of v
    // scale-only "original": residual |log(rectified / alpha^3)| = 3 ln2 |-lg2 alpha - lg2 v|: alpha enters reciprocally
    const double alpha_eff = T.problem == 2 ? 1.0 / m.alpha : m.alpha;
    r.A = log2(alpha_eff);
    const double Tp = T.Tp;
    const double v_lo = alpha_eff * T.exp2_neg_Tp, v_hi = alpha_eff * T.exp2_pos_Tp; // hypothesis-independent factors from the host
    const double kh = sc.k_max * (ah7 + ah8);
    const double eps_lg = 2.384185791015625e-7 + 3.0 * kU * (fabs(r.A) + Tp + 1.0); // MUFU.LG2 + roundings of A, d
    const double dv = 8.0 * kU * mag;
    r.drho32 = 2.0 * (dv / ((v_lo - dv) * kLn2) + eps_lg);
    r.dprime32 = 6.0 * kU * (2.0 + 3.0 * kh) * 1.0001;
    r.ok32_bound = isfinite(dv) && (v_lo > 4.0 * dv) && isfinite(r.drho32) && isfinite(r.dprime32);
    r.ok32 = r.ok32_bound && (r.drho32 < 1e-3 * Tp) && (r.dprime32 < 1e-3 * T.tanT);
    const double kU64 = 1.1102230246251565e-16;
    const double dv64 = 8.0 * kU64 * mag;
    r.drho64 = 2.0 * (dv64 / ((v_lo - dv64) * kLn2) + 1.5 * kU / kLn2 + eps_lg);
    r.dprime64 = 6.0 * kU64 * (2.0 + 3.0 * kh) * 1.0001;
    r.ok64 = isfinite(dv64) && (v_lo > 4.0 * dv64) && (r.drho64 < 1e-3 * Tp);
    r.common_ok = !all_exact && !T.orient_all_exact && !sc.irr_overflow && isfinite(r.A) && (v_hi * 1.01 < 1000.0) && (kh < 1e6) &&
                  (ah7 < 1e6) && (ah8 < 1e6);
    return r;
}

__global__ void __launch_bounds__(128) k_sample_solve(Buffers B, Thresholds T, int all_exact)
{
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned lane = threadIdx.x & 31;
    if (h == 0) B.phase_ns[6] = global_ns(); // diagnostics: start of the stage as seen by the first CTA
    bool solved = false, hard = false, sample_valid = false, sens_s = false, sens_o = false;
    bool scored = false; // goes to one of the three scoring lists (winner-only runs leave out the models isValidModel rejects)
    const bool winner = !B.per_hyp;
    Model4 m = {0.0, 0.0, 0.0, 0.0};
    float4 p0 = make_float4(0.f, 0.f, 0.f, 0.f), p1 = make_float4(0.f, 1.f, 0.f, 0.f);
    int s0 = 0, s1 = 0, o0 = 0, o1 = 0; // scale-only problems: (s0, s1, o0) index the scale class, o1 is unused (-1)
    const bool scale_only = T.problem != 0;
    if (h < B.n_hyp)
    {
        if (B.have_samples)
        {
            if (B.samples_u16)
            {
                const ushort4 s = reinterpret_cast<const ushort4*>(B.samples)[h];
                s0 = s.x; s1 = s.y; o0 = s.z; o1 = s.w == 0xffffu ? -1 : static_cast<int>(s.w);
            }
            else
            {
                const int4 s = reinterpret_cast<const int4*>(B.samples)[h];
                s0 = s.x; s1 = s.y; o0 = s.z; o1 = s.w;
            }
        }
        else
        {
            const unsigned long long g = B.first_hyp + static_cast<unsigned long long>(h);
            uint32_t c[4] = {static_cast<uint32_t>(g), static_cast<uint32_t>(g >> 32), 0x47435231u, 0u};
            philox4x32_10(c, static_cast<uint32_t>(B.seed), static_cast<uint32_t>(B.seed >> 32));
            draw_pair(c[0], c[1], static_cast<uint32_t>(B.ns), s0, s1);
            if (scale_only)
            {
                // third distinct index: uniform over the n - 2 remaining values, mapped around the two taken ones
                int t = static_cast<int>(__umulhi(c[2], static_cast<uint32_t>(B.ns) - 2u));
                const int lo = min(s0, s1), hi = max(s0, s1);
                if (t >= lo) t++;
                if (t >= hi) t++;
                o0 = t;
                o1 = -1;
            }
            else
                draw_pair(c[2], c[3], static_cast<uint32_t>(B.no), o0, o1);
            if (B.samples) reinterpret_cast<int4*>(B.samples)[h] = make_int4(s0, s1, o0, o1);
        }
    }
    // Every thread: wait for the kernel before this one (k_prep_features in the host-buffer path: the features and their per-feature
    // columns), and only THEN let the scoring kernel start: its prologue stages the feature planes that k_prep_features writes, and a
    // dependent's griddepcontrol.wait covers this kernel only, so the trigger must not fire while k_prep_features may still run.
    pdl_wait();
    pdl_launch_dependents();
    if (h < B.n_hyp)
    {
        const unsigned n_third = static_cast<unsigned>(scale_only ? B.ns : B.no);
        const bool in_range = static_cast<unsigned>(s0) < static_cast<unsigned>(B.ns) &&
                              static_cast<unsigned>(s1) < static_cast<unsigned>(B.ns) &&
                              static_cast<unsigned>(o0) < n_third && (scale_only || static_cast<unsigned>(o1) < n_third);
        if (!in_range)
        {
            atomicAdd(&B.ctr->n_bad_sample, 1u);
        }
        else if (scale_only)
        {
            double f0[3], f1[3], f2[3];
#pragma unroll
            for (int j = 0; j < 3; j++)
            {
                f0[j] = B.scale[3 * s0 + j];
                f1[j] = B.scale[3 * s1 + j];
                f2[j] = B.scale[3 * o0 + j];
            }
            const double pw[3] = {B.aux_s[s0], B.aux_s[s1], B.aux_s[o0]};
            sample_valid = !are_collinear(f0[0], f0[1], f1[0], f1[1], f2[0], f2[1], 1.0); // three_sift.hpp:111-143
            if (sample_valid) solved = estimate_minimal_scale3(f0, f1, f2, pw, T.problem == 2, m);
        }
        else
        {
            double fs0[3], fs1[3], fo0[3], fo1[3], vp[3];
#pragma unroll
            for (int j = 0; j < 3; j++)
            {
                fs0[j] = B.scale[3 * s0 + j];
                fs1[j] = B.scale[3 * s1 + j];
                fo0[j] = B.orient[3 * o0 + j];
                fo1[j] = B.orient[3 * o1 + j];
            }
            const double sc0[2] = {B.aux_o[2 * o0], B.aux_o[2 * o0 + 1]}, sc1[2] = {B.aux_o[2 * o1], B.aux_o[2 * o1 + 1]};
            const double pw0 = B.aux_s[s0], pw1 = B.aux_s[s1];
            sample_valid = is_valid_sample(fs0, fs1, fo0, fo1, sc0, sc1, vp); // GCRANSAC.h:315
            if (sample_valid) solved = estimate_minimal(fs0, fs1, pw0, pw1, vp, m); // GCRANSAC.h:326
        }
        unsigned char f = 0;
        if (sample_valid) f |= 1;
        if (solved)
        {
            f |= 2;
            // consulted at selection time, GCRANSAC.h:440-446; the scale-only solvers keep SolverEngine's default (always valid)
            const bool model_valid = T.problem != 0 || is_valid_model(m);
            if (model_valid) f |= 4;
            // The reference scores every model and consults isValidModel only when it would become the best one (GCRANSAC.h:440-446):
            // a rejected model can never be selected, so a winner-only run does not score it (its slot reports flags without bit 4)
            scored = model_valid || !winner;
            // ---- scoring-path parameters and their proven error bands (DESIGN.md "error budget")
            const PathBands pb = path_bands(m, *B.scene, T, all_exact != 0);
            const double A = pb.A, Tp = T.Tp;
            double drho = pb.drho32, dprime = pb.dprime32;
            // The 1e-3 caps keep the FP32 residuals of the inliers within the 1e-5 score bar; hypotheses that
            // miss them in FP32 (tiny alpha: cancellation in v; large |h|: angle noise) go to the precise tiles,
            // the rest (out of any proven range) to the FP64 reference-order kernel.
            // (A winner-only run needs no narrow bands, only valid ones: a wide band widens that hypothesis' error bound -- such models, tiny
            //  alpha or large |h|, score poorly and do not become contenders -- and the precise tiles, twice the cost per evaluation plus an
            //  F2F conversion on the XU pipe, shrink from ~4 % of the hypotheses to the few whose FP32 v could change sign.  Same-box A/B:
            //  C3 kernel 5.558 -> 5.384 ms, 125k-slot shard 0.724 -> 0.693 ms, C5 0.214 -> 0.208 ms, 11 instead of 1 contenders at C3.)
            if (pb.common_ok && (winner ? pb.ok32_bound : pb.ok32)) { }
            else if (pb.common_ok && pb.ok64)
            {
                sens_s = true;
                drho = pb.drho64;
                dprime = pb.dprime64;
            }
            else hard = true;
            if (hard && scored) f |= 8;
            double sphi, cphi;
            sincos(m.phi, &sphi, &cphi);
            if (winner)
            {
                // score_chunk_w: dq = A/Tp - lg2(v)/Tp; the class values carry an error of at most err_ab.x + err_ab.y * (sum of 1/mx over the orientation evaluations) + 8e-6 value
                // (DESIGN.md 5b): |delta dq| <= |delta (A - lg2 v)| / Tp + roundings of A/Tp and of the FMA (drho is twice its own bound of
                // |delta (A - lg2 v)|; that safety factor is left out here, the 8u of dv already doubles the first-order 4u); a feature near or
                // inside the threshold moves its term 1 - dq^2 by <= 2.1 |delta dq| + 2u.  Orientation: the term 1 - r^2/tsq moves by
                // <= 2.1 (delta t) / sqrt(tsq), delta t <= dprime / mx + t_band0, + 2.5e-6 (polynomial fit 5e-7, Horner and FFMA.SAT roundings,
                // the two extra roundings of the shared reciprocal)
                const double ddq = 0.5 * drho / Tp + 2.0 * kU * (fabs(A) / Tp + 2.0);
                // (a term 1 - x^2, |x| <= 1, moves by at most min(1, (2 + d) d) when x moves by d: exact for the per-hypothesis scale part; the
                //  orientation part is a sum over features of a + b / mx_f, for which 2.5 (a + b / mx_f) >= min(1, 2 x + x^2) at EVERY x >= 0)
                const double e_s = fmin(1.0, (2.0 + ddq) * ddq) + 2.0 * kU;
                const double e_o0 = 2.5 * static_cast<double>(T.t_band0) / static_cast<double>(T.sqrt_tsq1) + 2.5e-6;
                const double e_o1 = 2.5 * dprime / static_cast<double>(T.sqrt_tsq1);
                const double no_eff = T.problem == 0 ? static_cast<double>(B.no) : 0.0;
                B.err_ab[h] = make_float2(__double2float_ru(static_cast<double>(B.ns) * e_s + no_eff * e_o0), __double2float_ru(1.001 * e_o1));
                p0 = make_float4(static_cast<float>(-m.h7), static_cast<float>(-m.h8), static_cast<float>(A / Tp), 0.f);
                p1 = make_float4(0.f, static_cast<float>(cphi), static_cast<float>(sphi), 0.f);
            }
            else
            {
                // reporting modes: error of a CERTAIN inlier's term r^2/tsq (guard-band evaluations are re-done in FP64): scale
                // delta (d/Tp)^2 <= 1.05 drho / Tp (|delta d| <= drho / 2, |d| <= Tp) + 4u; orientation 2.1 (delta t) / sqrt(tsq) with
                // delta t <= dprime / max + t_band0, + 1.2e-6 for the polynomial (finalize_phase sums them over the counted inliers)
                const double dd_rep = 0.5 * drho / Tp;
                B.err_ab[h] = make_float2(__double2float_ru(fmin(1.0, (2.0 + dd_rep) * dd_rep) + 4.0 * kU),
                                          __double2float_ru(2.5 * dprime / static_cast<double>(T.sqrt_tsq1)));
                p0 = make_float4(static_cast<float>(-m.h7), static_cast<float>(-m.h8), static_cast<float>(A),
                                 __double2float_rd(Tp - drho));
                p1 = make_float4(__double2float_ru(Tp + drho), static_cast<float>(cphi), static_cast<float>(sphi),
                                 __double2float_ru(dprime));
            }
        }
        B.flags[h] = f;
        B.models[h] = m;
        if (!winner) B.score[h] = 0.0; // (its low word collects a float sum in the scoring kernels, see flush_hyps)
        B.sums[2 * h] = 0.0;
        B.sums[2 * h + 1] = 0.0;
        B.cnts[2 * h] = 0u;
        B.cnts[2 * h + 1] = 0u;
    }
    // block-aggregated compaction of solved hypotheses into the fast / precise / exact lists: warp ballots, a prefix over
    // the block's warps in shared memory, ONE returning atomic per list and block (per-warp atomics on the same three
    // counters serialise in L2 and were half of this kernel's time)
    __shared__ unsigned int s_cnt[5][4];  // [list][warp]: fast, sens, hard, sample_valid, solved
    __shared__ unsigned int s_base[3];
    const unsigned warp = threadIdx.x >> 5;
    const bool sens = sens_s || sens_o;
    const unsigned m_valid = __ballot_sync(0xffffffffu, sample_valid);
    const unsigned m_fast = __ballot_sync(0xffffffffu, scored && !hard && !sens);
    const unsigned m_sens = __ballot_sync(0xffffffffu, scored && !hard && sens);
    const unsigned m_hard = __ballot_sync(0xffffffffu, scored && hard);
    const unsigned m_solved = __ballot_sync(0xffffffffu, solved);
    if (lane == 0)
    {
        s_cnt[0][warp] = __popc(m_fast);
        s_cnt[1][warp] = __popc(m_sens);
        s_cnt[2][warp] = __popc(m_hard);
        s_cnt[3][warp] = __popc(m_valid);
        s_cnt[4][warp] = __popc(m_solved);
    }
    __syncthreads();
    if (threadIdx.x < 3)
    {
        const unsigned int tot = s_cnt[threadIdx.x][0] + s_cnt[threadIdx.x][1] + s_cnt[threadIdx.x][2] + s_cnt[threadIdx.x][3];
        unsigned int* ctr = threadIdx.x == 0 ? &B.ctr->n_fast : (threadIdx.x == 1 ? &B.ctr->n_sens : &B.ctr->n_hard);
        s_base[threadIdx.x] = tot ? atomicAdd(ctr, tot) : 0u;
    }
    else if (threadIdx.x == 32)
    {
        const unsigned int nv = s_cnt[3][0] + s_cnt[3][1] + s_cnt[3][2] + s_cnt[3][3];
        const unsigned int ns = s_cnt[4][0] + s_cnt[4][1] + s_cnt[4][2] + s_cnt[4][3];
        if (nv) atomicAdd(&B.ctr->n_sample_valid, nv);
        if (ns) atomicAdd(&B.ctr->n_solved, ns);
    }
    __syncthreads();
    unsigned base_fast = s_base[0], base_sens = s_base[1], base_hard = s_base[2];
    for (unsigned w = 0; w < warp; w++)
    {
        base_fast += s_cnt[0][w];
        base_sens += s_cnt[1][w];
        base_hard += s_cnt[2][w];
    }
    const unsigned below = (1u << lane) - 1u;
    if (scored && !hard)
    {
        // normal hypotheses fill the compact arrays from the front, score-sensitive ones from the back, so
        // that the rare-path traffic of the latter stays in their own warps
        const unsigned c = sens ? static_cast<unsigned>(B.n_hyp) - 1u - (base_sens + __popc(m_sens & below))
                                : base_fast + __popc(m_fast & below);
        B.pf0[c] = p0;
        B.pf1[c] = p1;
        B.fast_hyp[c] = h;
    }
    if (scored && hard) B.hard_hyp[base_hard + __popc(m_hard & below)] = h;
}

// ------------------------------------------------------------------------------------------------
// K3x: FP64 re-evaluation of guard-band pairs and of irregular features, patched in with atomics.
__device__ __forceinline__ void fixup_phase(const Buffers& B, const Thresholds& T)
{
    const unsigned int n_wl = min(__ldcg(&B.ctr->n_worklist), B.worklist_cap);
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int t0 = blockIdx.x * blockDim.x + threadIdx.x;
    for (unsigned int i = t0; i < n_wl; i += stride)
    {
        const uint2 w = B.worklist[i];
        const unsigned int h = w.x, cls = w.y >> 31, f = w.y & 0x7fffffffu;
        const Model4 m = B.models[h];
        double r;
        if (cls == 0) r = scale_residual(B.scale[3 * f], B.scale[3 * f + 1], B.scale[3 * f + 2], m, T.problem == 2);
        else r = orientation_residual(B.orient[3 * f], B.orient[3 * f + 1], B.orient[3 * f + 2], m);
        const double r2 = r * r;
        if (r2 <= T.tsq[cls]) // S1
        {
            atomicAdd(&B.sums[2 * h + cls], r2);
            atomicAdd(&B.cnts[2 * h + cls], 1u);
            if (B.masks) atomicOr(&B.masks[static_cast<size_t>(h) * B.wmask + (cls ? B.ws : 0) + (f >> 5)], 1u << (f & 31));
        }
    }
    // irregular features x every fast-path hypothesis
    const unsigned int n_is = min(B.scene->n_irr_s, static_cast<unsigned>(kIrrCap));
    const unsigned int n_io = min(B.scene->n_irr_o, static_cast<unsigned>(kIrrCap));
    const unsigned int n_irr = n_is + n_io;
    if (n_irr == 0) return;
    const unsigned int n_fast = B.ctr->n_fast, n_both = n_fast + B.ctr->n_sens;
    const unsigned long long total = static_cast<unsigned long long>(n_irr) * n_both;
    for (unsigned long long i = t0; i < total; i += stride)
    {
        const unsigned int c = static_cast<unsigned int>(i / n_irr), k = static_cast<unsigned int>(i % n_irr);
        const unsigned int h = B.fast_hyp[c < n_fast ? c : static_cast<unsigned>(B.n_hyp) - 1u - (c - n_fast)];
        const unsigned int cls = k >= n_is;
        const unsigned int f = cls ? B.irr_o[k - n_is] : B.irr_s[k];
        const Model4 m = B.models[h];
        double r;
        if (cls == 0) r = scale_residual(B.scale[3 * f], B.scale[3 * f + 1], B.scale[3 * f + 2], m, T.problem == 2);
        else r = orientation_residual(B.orient[3 * f], B.orient[3 * f + 1], B.orient[3 * f + 2], m);
        const double r2 = r * r;
        if (r2 <= T.tsq[cls])
        {
            if (!B.per_hyp) // winner-only run: the accumulators are class values sum max(0, 1 - r^2 / tsq)
            {
                if (T.tsq[cls] > 0.0) atomicAdd(&B.sums[2 * h + cls], 1.0 - r2 / T.tsq[cls]);
                continue;
            }
            atomicAdd(&B.sums[2 * h + cls], r2);
            atomicAdd(&B.cnts[2 * h + cls], 1u);
            if (B.masks) atomicOr(&B.masks[static_cast<size_t>(h) * B.wmask + (cls ? B.ws : 0) + (f >> 5)], 1u << (f & 31));
        }
    }
}

// ------------------------------------------------------------------------------------------------
// FP64 reference-order MSAC of whole hypotheses: MSACScoringFunction::getScore
// (MSAC_scoring_function.hpp:53-130).  Work unit = (hypothesis, block of 256 consecutive features of one class);
// a warp covers 32 consecutive features so that __ballot_sync yields the packed mask word directly.
// Partial sums are accumulated ORDER-INDEPENDENTLY as 64-bit fixed point (units of tsq * 2^-44 per class), so the
// result does not depend on how units are scheduled: equal models get bit-equal scores (S11 ties stay ties).
//   mode 0: the hard list (accumulators: the hypothesis' own sums/cnts slots, reinterpreted as u64)
//   mode 1: the candidate list; mode 2: `n_direct` caller-supplied models in candidate slots [0, n_direct)
constexpr int kExactThreads = 256;
constexpr double kFix = 17592186044416.0; // 2^44

__device__ __forceinline__ double fix_to_double(unsigned long long q, double tsq) { return static_cast<double>(q) * (tsq / kFix); }

// S2/S3: per-class value = -sum r^2 / tsq + n; any class below its minimal sample size zeroes everything
__device__ __forceinline__ double msac_value(const Thresholds& T, double s0, double s1, unsigned int c0, unsigned int c1,
                                             unsigned int& n0, unsigned int& n1)
{
    if (c0 < T.min_inl[0] || c1 < T.min_inl[1])
    {
        n0 = n1 = 0u;
        return 0.0;
    }
    n0 = c0;
    n1 = c1;
    const double v0 = -s0 / T.tsq[0] + static_cast<double>(c0);
    if (T.problem != 0) return v0; // one class
    return v0 + (-s1 / T.tsq[1] + static_cast<double>(c1));
}

__device__ __forceinline__ void exact_items_phase(const Buffers& B, const Thresholds& T, int mode, unsigned int n_direct,
                                                  unsigned int hard_first = 0u, unsigned int hard_count = 0xffffffffu)
{
    __shared__ unsigned long long red[kExactThreads / 32];
    __shared__ unsigned int redc[kExactThreads / 32];
    const unsigned int n_items = mode == 0 ? (hard_count != 0xffffffffu ? hard_count : __ldcg(&B.ctr->n_hard)) : (mode == 1 ? min(__ldcg(&B.ctr->n_candidates) - n_direct, static_cast<unsigned>(kCandCap)) : n_direct);
    const unsigned int ups = (B.ns + kExactThreads - 1) / kExactThreads, upo = (B.no + kExactThreads - 1) / kExactThreads;
    const unsigned int upi = ups + upo;
    const unsigned long long n_units = static_cast<unsigned long long>(n_items) * upi;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // mode 0 runs in the tail's first phase behind the fix-up pairs, which fill the CTAs from the front: units from the back,
    // so that no CTA has both chains (each one FP64 residual long, ~3.5 us) to run one after the other
    const unsigned int first_cta = mode == 0 ? gridDim.x - 1u - blockIdx.x : blockIdx.x;
    for (unsigned long long unit = first_cta; unit < n_units; unit += gridDim.x)
    {
        const unsigned int it = static_cast<unsigned int>(unit / upi), u = static_cast<unsigned int>(unit % upi);
        const int cls = u >= ups;
        const int f = static_cast<int>((cls ? u - ups : u) * kExactThreads) + tid;
        const int n = cls ? B.no : B.ns;
        const double* F = cls ? B.orient : B.scale;
        const unsigned int h = mode == 0 ? B.hard_hyp[hard_first + it] : 0u;
        const Model4 m = mode == 0 ? B.models[h] : B.cand_model[it];
        unsigned int* mask_out = mode == 0 ? (B.masks ? B.masks + static_cast<size_t>(h) * B.wmask : nullptr)
                                           : B.cand_masks + static_cast<size_t>(it) * B.wmask;
        bool in = false;
        unsigned long long q = 0ull;
        if (f < n)
        {
            const double r = cls ? orientation_residual(F[3 * f], F[3 * f + 1], F[3 * f + 2], m)
                                 : scale_residual(F[3 * f], F[3 * f + 1], F[3 * f + 2], m, T.problem == 2);
            const double r2 = r * r;
            in = T.strict ? r2 < T.tsq[cls] : r2 <= T.tsq[cls]; // S1 (strict: one-class graph-cut relabelling)
            if (in) q = static_cast<unsigned long long>(r2 / T.tsq[cls] * kFix + 0.5);
        }
        const unsigned int word = __ballot_sync(0xffffffffu, in);
        if (mask_out && lane == 0 && (f - lane) < n) mask_out[(cls ? B.ws : 0) + ((f - lane) >> 5)] = word;
        unsigned int k = __popc(word);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
        if (lane == 0)
        {
            red[warp] = q;
            redc[warp] = k;
        }
        __syncthreads();
        if (tid == 0)
        {
            unsigned long long tq = 0ull;
            unsigned int tk = 0u;
            for (int w = 0; w < kExactThreads / 32; w++)
            {
                tq += red[w];
                tk += redc[w];
            }
            unsigned long long* qs = mode == 0 ? reinterpret_cast<unsigned long long*>(B.sums) + 2 * static_cast<size_t>(h)
                                               : B.cand_qsum + 2 * static_cast<size_t>(it);
            unsigned int* cs = mode == 0 ? B.cnts + 2 * static_cast<size_t>(h) : B.cand_rawcnt + 2 * static_cast<size_t>(it);
            if (tk)
            {
                atomicAdd(qs + cls, tq);
                atomicAdd(cs + cls, tk);
            }
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kExactThreads) k_exact_items(Buffers B, Thresholds T, int mode, unsigned int n_direct = 1u)
{
    pdl_launch_dependents(); // chained launch: the next kernel of the call may start launching now; it waits at its own top
    pdl_wait();
    exact_items_phase(B, T, mode, n_direct);
}

// candidate accumulators -> MSAC value + counts after the zero-reset rule
__device__ __forceinline__ void cand_value(const Buffers& B, const Thresholds& T, unsigned int i)
{
    unsigned int n0, n1;
    B.cand_score[i] = msac_value(T, fix_to_double(B.cand_qsum[2 * i], T.tsq[0]), fix_to_double(B.cand_qsum[2 * i + 1], T.tsq[1]),
                                 B.cand_rawcnt[2 * i], B.cand_rawcnt[2 * i + 1], n0, n1);
    B.cand_cnt[2 * i] = n0;
    B.cand_cnt[2 * i + 1] = n1;
}
__global__ void k_cand_value(Buffers B, Thresholds T, unsigned int n)
{
    pdl_launch_dependents(); // chained launch: the next kernel of the call may start launching now; it waits at its own top
    pdl_wait();
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) cand_value(B, T, i);
}

// gcr_score_models: values of the n caller models, written together with their counts and (optionally) packed masks
// straight into page-locked host memory (device-mapped): the caller synchronises and reads, no device-to-host copies
__global__ void k_score_out(Buffers B, Thresholds T, unsigned int n, double* __restrict__ h_score, unsigned int* __restrict__ h_cnt,
                            unsigned int* __restrict__ h_masks)
{
    pdl_launch_dependents(); // chained launch: the next kernel of the call may start launching now; it waits at its own top
    pdl_wait();
    const unsigned int t = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    for (unsigned int i = t; i < n; i += stride)
    {
        cand_value(B, T, i);
        h_score[i] = B.cand_score[i];
        h_cnt[2 * i] = B.cand_cnt[2 * i];
        h_cnt[2 * i + 1] = B.cand_cnt[2 * i + 1];
    }
    if (h_masks)
    {
        const size_t words = static_cast<size_t>(n) * B.wmask;
        for (size_t w = t; w < words; w += stride) h_masks[w] = B.cand_masks[w];
    }
}

// ------------------------------------------------------------------------------------------------
// K4-K6 as phases of ONE cooperative kernel (k_tail): at small sizes (C2) six dependent launches of 4-7 us each cost
// as much as the scoring kernel; grid-wide barriers between the phases cost ~1-2 us.
//
// finalize: per-hypothesis MSAC value + arg-max key over model-valid hypotheses.
// MSAC value of a scored hypothesis from its accumulators (flag 8: fixed-point accumulators of the FP64 kernel)
__device__ __forceinline__ double hyp_value(const Buffers& B, const Thresholds& T, unsigned int h, unsigned char f,
                                            unsigned int& n0, unsigned int& n1)
{
    double s0 = B.sums[2 * h], s1 = B.sums[2 * h + 1];
    if (f & 8)
    {
        s0 = fix_to_double(reinterpret_cast<const unsigned long long*>(B.sums)[2 * h], T.tsq[0]);
        s1 = fix_to_double(reinterpret_cast<const unsigned long long*>(B.sums)[2 * h + 1], T.tsq[1]);
    }
    return msac_value(T, s0, s1, B.cnts[2 * h], B.cnts[2 * h + 1], n0, n1);
}

// i-th scored hypothesis: FP32 tiles (front of fast_hyp), precise tiles (back), FP64 list
__device__ __forceinline__ unsigned int scored_hyp(const Buffers& B, unsigned int i, unsigned int n_fast, unsigned int n_sens)
{
    if (i < n_fast) return B.fast_hyp[i];
    if (i < n_fast + n_sens) return B.fast_hyp[static_cast<unsigned int>(B.n_hyp) - 1u - (i - n_fast)];
    return B.hard_hyp[i - n_fast - n_sens];
}

// Small batches walk the slots directly (one dependent load less than through the compact lists, and the phase is
// latency-bound there); big ones walk the compact lists (29 % of the slots at C2's validity rate).
__device__ __forceinline__ bool tail_direct(const Buffers& B)
{
    return B.per_hyp || static_cast<unsigned int>(B.n_hyp) <= 2u * gridDim.x * blockDim.x;
}

__device__ __forceinline__ void finalize_phase(const Buffers& B, const Thresholds& T)
{
    // reporting modes: value and inlier counts of EVERY slot are outputs; the selection works on proven bounds: counts are exact
    // (guard-band pairs were decided in FP64), the FP32 sum of r^2/tsq over the n0 + n1 certain inliers is off by at most
    //   err = n0 e_s + e_o1 (sum of 1/max over them) + n1 (2.1 t_band0 / sqrt(tsq) + 1.2e-6) + 8e-6 (n0 + n1)
    // (k_sample_solve's per-hypothesis coefficients err_ab; the last term: FP32 accumulation inside a chunk), kept in err_ab.x for collect_phase
    unsigned long long key = 0ull;
    const double e_o0 = 2.5 * static_cast<double>(T.t_band0) / fmax(static_cast<double>(T.sqrt_tsq1), 1e-300) + 1.2e-6;
    for (int base = blockIdx.x * blockDim.x; base < B.n_hyp; base += gridDim.x * blockDim.x)
    {
        const int h = base + threadIdx.x;
        if (h >= B.n_hyp) continue;
        const unsigned char f = B.flags[h];
        unsigned int n0 = 0, n1 = 0;
        double v = 0.0, err = 0.0;
        if (f & 2)
        {
            const float risum = *reinterpret_cast<const float*>(B.score + h);
            const unsigned int c0 = B.cnts[2 * h], c1 = B.cnts[2 * h + 1];
            v = hyp_value(B, T, h, f, n0, n1);
            if (!(f & 8))
            {
                const float2 ab = B.err_ab[h];
                err = static_cast<double>(c0) * ab.x + static_cast<double>(ab.y) * risum + static_cast<double>(c1) * e_o0 +
                      8e-6 * static_cast<double>(c0 + c1) + 1e-12;
            }
            B.err_ab[h].x = __double2float_ru(err);
        }
        B.score[h] = v;
        B.n_inl[2 * h] = n0;
        B.n_inl[2 * h + 1] = n1;
        if ((f & 4) && v > 0.0) // S10 + S11 (initial best is 0, strict '<'); counts are exact here, so v > 0 is the reference's own test
        {
            const unsigned long long k = ordered_bits(v - err);
            key = k > key ? k : key;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other > key ? other : key;
    }
    // one atomic per CTA (2 368 warp leaders on one address serialise in L2 for ~1 us)
    __shared__ unsigned long long s_key[kExactThreads / 32];
    if ((threadIdx.x & 31) == 0) s_key[threadIdx.x >> 5] = key;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        for (unsigned int w = 1; w < blockDim.x / 32; w++) key = s_key[w] > key ? s_key[w] : key;
        if (key) atomicMax(&B.ctr->best_key, key);
    }
}

// collect: hypotheses whose FP32-path score is within the path's error of the best one are re-scored in FP64
// so that the selection rule sees reference-order scores.
// Candidate slots are handed out from cumulative counters (slot = ticket - cand_base of this round).  A hypothesis that
// finds the 1024 slots of the round taken is marked "deferred" (flag bit 16) and collected by a later round.
__device__ __forceinline__ void collect_candidate(const Buffers& B, unsigned int h, unsigned int base, unsigned char f)
{
    const unsigned int slot = atomicAdd(&B.ctr->n_candidates, 1u) - base;
    if (slot < kCandCap)
    {
        B.cand_hyp[slot] = h;
        B.cand_model[slot] = B.models[h];
        B.cand_qsum[2 * slot] = B.cand_qsum[2 * slot + 1] = 0ull;
        B.cand_rawcnt[2 * slot] = B.cand_rawcnt[2 * slot + 1] = 0u;
        if (f & 16) B.flags[h] = f & ~16;
    }
    else
    {
        atomicAdd(&B.ctr->n_cand_overflow, 1u);
        if (!(f & 16)) B.flags[h] = f | 16;
    }
}

// round 0: every model-valid hypothesis inside the window; later rounds: the deferred ones
__device__ __forceinline__ void collect_phase(const Buffers& B, const Thresholds& T, int round, unsigned int base)
{
    const unsigned long long bk = __ldcg(&B.ctr->best_key);
    if (bk == 0ull) return;
    const double best_lb = from_ordered_bits(bk); // best LOWER bound (finalize_phase)
    for (int h = blockIdx.x * blockDim.x + threadIdx.x; h < B.n_hyp; h += gridDim.x * blockDim.x)
    {
        const unsigned char f = B.flags[h];
        if ((f & 6) != 6) continue; // solved and model-valid
        if (round > 0)
        {
            if (f & 16) collect_candidate(B, h, base, f);
            continue;
        }
        const double v = B.score[h];
        if (!(v > 0.0) || v + static_cast<double>(B.err_ab[h].x) < best_lb) continue; // upper bound below the best lower bound: cannot win
        collect_candidate(B, h, base, f);
    }
}

// ---- winner-only runs (Buffers::per_hyp == 0): a three-level cascade.
//  1. The scoring kernels leave per class the truncated-quadratic value S_i = sum_f max(0, 1 - r^2/tsq_i) in sums[2h + i] and the largest
//     1/mx of the orientation evaluations in cnts[2h + 1]; the FP32 evaluation error of S_0 + S_1 is bounded per hypothesis by
//     err = err_ab.x + err_ab.y * (sum of 1/mx, bounded chunk by chunk) + 8e-6 value (k_sample_solve).  That bound is a sum over ALL features (which of them
//     sit near the threshold is not known), i.e. loose: ~1e-3 of the value for 70 000 features.
//  2. CONTENDERS -- hypotheses whose upper bound reaches the best lower bound -- are re-scored by refine_phase in mid precision with exact
//     inlier decisions (FP64 front end, FP32 MUFU, a guard band of a few ulps decided by the FP64 reference-order residuals): exact counts,
//     hence the zero-reset rule S3, and a value whose error is bounded over the INLIERS only (~1e-6 of the value).
//  3. The few contenders whose refined upper bound reaches the best refined lower bound are scored in FP64 reference order (candidates),
//     and the selection rule (GCRANSAC.h:440-446, strict '<', earliest wins) sees the reference's own values.
// Value and loose error bound of a scored, model-valid hypothesis (level 1).
__device__ __forceinline__ double hyp_value_w(const Buffers& B, const Thresholds& T, unsigned int h, unsigned char f, double& err)
{
    err = 0.0;
    if (f & 8) // scored by the FP64 reference-order kernel: exact, zero-reset rule applied
    {
        unsigned int n0, n1;
        return hyp_value(B, T, h, f, n0, n1);
    }
    const double s0 = B.sums[2 * h], s1 = T.problem == 0 ? B.sums[2 * h + 1] : 0.0;
    const float2 ab = B.err_ab[h];
    const double v = s0 + s1;
    err = static_cast<double>(ab.x) + static_cast<double>(ab.y) * static_cast<double>(__uint_as_float(B.cnts[2 * h + 1])) + 8e-6 * v + 1e-9;
    return v;
}

// Calls fn(h, flags) for every scored, model-valid hypothesis of the run, grid-strided (directly over the slots for small batches,
// through the compact lists for big ones, as finalize_phase does).
template <typename F>
__device__ __forceinline__ void for_each_scored_valid(const Buffers& B, F fn)
{
    if (tail_direct(B))
    {
        for (int h = blockIdx.x * blockDim.x + threadIdx.x; h < B.n_hyp; h += gridDim.x * blockDim.x)
        {
            const unsigned char f = B.flags[h];
            if ((f & 6) == 6) fn(static_cast<unsigned int>(h), f);
        }
        return;
    }
    const unsigned int n_fast = __ldcg(&B.ctr->n_fast), n_sens = __ldcg(&B.ctr->n_sens), n_hard = __ldcg(&B.ctr->n_hard);
    const unsigned int total = n_fast + n_sens + n_hard;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x)
    {
        const unsigned int h = scored_hyp(B, i, n_fast, n_sens);
        const unsigned char f = B.flags[h];
        if (f & 4) fn(h, f);
    }
}

// block-wide maximum of an ordered key, one atomic per CTA
__device__ __forceinline__ void publish_key(unsigned long long* dst, unsigned long long key)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other > key ? other : key;
    }
    __shared__ unsigned long long s_keyw[kExactThreads / 32];
    if ((threadIdx.x & 31) == 0) s_keyw[threadIdx.x >> 5] = key;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        for (unsigned int w = 1; w < blockDim.x / 32; w++) key = s_keyw[w] > key ? s_keyw[w] : key;
        if (key) atomicMax(dst, key);
    }
    __syncthreads();
}

// level 1: best LOWER bound over the hypotheses that may have a positive score (S11: a model needs one)
// What a thread met in finalize_phase_w, kept in registers across the barrier: (hypothesis, upper bound) of up to kUbCache hypotheses, so that
// the first contender collection decides from registers instead of walking the lists and re-loading five arrays per hypothesis again.
constexpr int kUbCache = 4;
struct UbCache
{
    unsigned int h[kUbCache];
    double ub[kUbCache];
    int n;         // entries held
    bool overflow; // this thread met more than kUbCache hypotheses: it walks its share of the lists again
};
__device__ __forceinline__ void finalize_phase_w(const Buffers& B, const Thresholds& T, UbCache& cache)
{
    unsigned long long key = 0ull;
    cache.n = 0;
    cache.overflow = false;
    for_each_scored_valid(B, [&](unsigned int h, unsigned char f) {
        double err;
        const double v = hyp_value_w(B, T, h, f, err);
        if (v + err > 0.0)
        {
            const unsigned long long k = ordered_bits(v - err);
            key = k > key ? k : key;
            if (cache.n < kUbCache)
            {
#pragma unroll
                for (int i = 0; i < kUbCache; i++) // (static indices: the arrays stay in registers)
                    if (i == cache.n)
                    {
                        cache.h[i] = h;
                        cache.ub[i] = v + err;
                    }
                cache.n++;
            }
            else
                cache.overflow = true;
        }
    });
    publish_key(&B.ctr->best_key, key);
}

// level 2, collection: every not yet refined hypothesis whose upper bound reaches `floor_v` takes a contender slot; the collecting thread
// derives the refinement's parameters (FP64 model, FP32 band limits of the FP64-front arithmetic).  Slots beyond kContCap: the overflow flag
// sends the stragglers straight to level 3.
constexpr unsigned int kDirectCand = 16;
__device__ __forceinline__ void contender_collect_phase(const Buffers& B, const Thresholds& T, double floor_v, bool speculate,
                                                        const UbCache* cache = nullptr)
{
    auto consider = [&](unsigned int h, unsigned char f) {
        if (f & 32) return; // already a contender
        double err;
        const double v = hyp_value_w(B, T, h, f, err);
        if (!(v + err > 0.0) || v + err < floor_v) return;
        const unsigned int slot = atomicAdd(&B.ts->n_cont, 1u);
        if (slot >= B.cont_cap)
        {
            B.ts->cont_overflow = 1u;
            return;
        }
        B.flags[h] = f | 32;
        // class values that prove the minimal inlier counts by themselves (every term is <= 1): this contender cannot lose its score to the
        // zero-reset rule, its lower bound is a valid one
        if (!(f & 8))
        {
            const double s0 = B.sums[2 * h], s1 = B.sums[2 * h + 1];
            if (!(s0 - err > static_cast<double>(T.min_inl[0]) - 1.0) || (T.problem == 0 && !(s1 - err > static_cast<double>(T.min_inl[1]) - 1.0)))
                B.ts->unproven = 1u;
        }
        ContPar P;
        const Model4 m = B.models[h];
        P.hyp = h;
        P.exact = (f & 8) ? 1u : 0u;
        P.h7 = m.h7;
        P.h8 = m.h8;
        sincos(m.phi, &P.sphi, &P.cphi);
        const PathBands pb = path_bands(m, *B.scene, T, false);
        const double Tp = T.Tp;
        P.A = static_cast<float>(pb.A);
        P.Tin = __double2float_rd(Tp - pb.drho64);
        P.Tout = __double2float_ru(Tp + pb.drho64);
        P.dpr = __double2float_ru(pb.dprime64);
        // error of a CERTAIN inlier's term r^2/tsq in this arithmetic (band evaluations are decided and valued in FP64):
        //   scale        |delta d| <= drho64 / 2  ->  delta (d/Tp)^2 <= 1.05 drho64 / Tp (+ the FP32 rounding of d^2's factor)
        //   orientation  |delta t| <= 2 band0r (evaluations whose dprime64 / mx exceeds band0r are sent to FP64)  ->  2.1 * 2 band0r / sqrt(tsq),
        //                + 1.2e-6 for the polynomial fit (5e-7) and its FP32 evaluation
        const double band0r = 8.0 * kU * static_cast<double>(T.tanT) + 1e-12;
        const double dd_ref = 0.5 * pb.drho64 / Tp;
        P.e_s = __double2float_ru(fmin(1.0, (2.0 + dd_ref) * dd_ref) + 4.0 * kU);
        P.e_o = __double2float_ru(5.0 * band0r / static_cast<double>(T.sqrt_tsq1) + 1.2e-6);
        P.ri_cap = __double2float_rd(band0r / fmax(pb.dprime64, 1e-300));
        P.band0r = __double2float_ru(band0r);
        B.cont_par[slot] = P;
        B.cont_q[2 * slot] = B.cont_q[2 * slot + 1] = 0ull;
        B.cont_n[2 * slot] = B.cont_n[2 * slot + 1] = 0u;
        // first round: the first few contenders also take candidate slots at once; when they turn out to be all there is (and none can
        // lose its score to the zero-reset rule) the FP64 level starts without a collection pass and its barrier
        if (speculate && slot < kDirectCand) collect_candidate(B, h, 0u, f | 32);
    };
    if (cache && !cache->overflow)
    {
        // first round: this thread's hypotheses and their upper bounds are still in its registers (finalize_phase_w)
#pragma unroll
        for (int i = 0; i < kUbCache; i++)
            if (i < cache->n && cache->ub[i] >= floor_v) consider(cache->h[i], B.flags[cache->h[i]]);
        return;
    }
    for_each_scored_valid(B, consider);
}

// level 2, evaluation: unit = (contender, 256 consecutive features of one class), thread <-> feature.  The three cancelling FMAs
// (v = 1/q - h7 x/q - h8 y/q; numer, denom) and the rotation by -phi in FP64 from the FP64 planes, rounded once; MUFU.LG2 / MUFU.RCP
// and the threshold tests in FP32 against band limits that bound what is left (2^-22 of lg2 + roundings; a few ulps of t): certainly
// in -> r^2 from the FP32 value, certainly out -> nothing, in between (a few 1e-5 of the evaluations) and irregular features -> the
// FP64 reference-order residual decides and gives r^2.  Sums of r^2/tsq in 64-bit fixed point (order-independent), exact counts.
constexpr int kRefineFW = 64;  // features per warp and unit
constexpr int kRefineIlp = 4;  // independent evaluations in flight per thread
__device__ __forceinline__ void refine_phase(const Buffers& B, const Thresholds& T, unsigned int c_first, unsigned int c_end)
{
    constexpr int kCD = kChunk / 2;
    constexpr int kFB = kRefineFW * (kExactThreads / 32); // features per unit (one CTA)
    if (c_end <= c_first) return;
    // The mapping of the scoring kernels: one LANE per contender (its parameters and sums in registers, no reduction), the features read as
    // warp-uniform loads of the FP64 planes; unit = (32 contenders) x (kFB features of one class), each warp of the CTA a kRefineFW slice.
    const unsigned int n_groups = (c_end - c_first + 31) / 32;
    const unsigned int bs = (B.ns + kFB - 1) / kFB, bo = (B.no + kFB - 1) / kFB, fbk = bs + bo;
    const unsigned int n_units = n_groups * fbk;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const double unit_s = static_cast<double>(T.scale_unit);
    __shared__ double s_pl[3][kFB]; // the unit's block of the three FP64 planes (12 KB), staged once per unit
    for (unsigned int unit = blockIdx.x; unit < n_units; unit += gridDim.x)
    {
        const unsigned int g = unit / fbk, u = unit % fbk;
        const int cls = u >= bs;
        const int n = cls ? B.no : B.ns;
        const int blk0 = static_cast<int>((cls ? u - bs : u) * kFB);
        const double* planes = cls ? B.fod : B.fsd;
        __syncthreads(); // the previous unit's readers are done
        for (int i = threadIdx.x; i < kFB; i += blockDim.x)
        {
            const int f = min(blk0 + i, n - 1); // (slots past the end repeat the last feature and are not counted)
            const double* pl = planes + static_cast<size_t>(f / kCD) * (3 * kCD) + (f % kCD);
            s_pl[0][i] = pl[0];
            s_pl[1][i] = pl[kCD];
            s_pl[2][i] = pl[2 * kCD];
        }
        __syncthreads();
        const int f_begin = blk0 + warp * kRefineFW;
        const int f_end = min(n, f_begin + kRefineFW);
        const unsigned int c = c_first + g * 32 + lane;
        ContPar P;
        P.exact = 1u;
        if (c < c_end) P = B.cont_par[c];
        const bool active = !P.exact; // (a contender the FP64 kernel scored already needs no refinement)
        unsigned long long q = 0ull;
        unsigned int cnt = 0u;
        for (int fb = f_begin; fb < f_end; fb += kRefineIlp)
        {
            float x0[kRefineIlp], x1[kRefineIlp]; // scale: d; orientation: t, 1/mx.  The arithmetic of all features first (independent chains)
            bool bad[kRefineIlp];                 // irregular feature (sentinel plane value): FP64 decides
#pragma unroll
            for (int j = 0; j < kRefineIlp; j++)
            {
                const int i = fb + j - blk0; // (< kFB: kRefineFW is a multiple of kRefineIlp)
                const double pa = s_pl[0][i], pb = s_pl[1][i], pc = s_pl[2][i];
                if (cls == 0)
                {
                    const float v = static_cast<float>(fma(-P.h8, pb, fma(-P.h7, pa, pc)));
                    x0[j] = P.A - lg2_approx(v);
                    x1[j] = 0.f;
                    bad[j] = !(pc < 1e300);
                }
                else
                {
                    const double nn = fma(pa, -P.h7, pb), dd = fma(pa, P.h8, pc);
                    const double c1 = fma(nn, P.sphi, dd * P.cphi), s1 = fma(nn, P.cphi, -(dd * P.sphi));
                    const float a1 = static_cast<float>(fabs(c1)), a2 = static_cast<float>(fabs(s1));
                    const float mn = fminf(a1, a2), mx = fmaxf(fmaxf(a1, a2), 1e-30f);
                    x1[j] = rcp_approx(mx);
                    x0[j] = mn * x1[j];
                    bad[j] = pb != pb;
                }
            }
#pragma unroll
            for (int j = 0; j < kRefineIlp; j++)
            {
                const int f = fb + j;
                if (f >= f_end || !active) continue;
                bool in, fallback;
                double r2;
                if (cls == 0)
                {
                    const float d = x0[j];
                    in = fabsf(d) <= P.Tin;
                    fallback = (!in && fabsf(d) <= P.Tout) || bad[j]; // guard band, or an irregular feature
                    r2 = static_cast<double>(d) * static_cast<double>(d) * unit_s;
                }
                else
                {
                    const float t = x0[j], ri = x1[j];
                    const float band = fmaf(P.dpr, ri, P.band0r);
                    const bool wide = !(ri <= P.ri_cap); // the FP64 part of the band is no longer negligible: decide in FP64
                    in = !wide && t <= T.tanT - band;
                    fallback = (!in && t <= T.tanT + band) || bad[j];
                    const float z = t * t;
                    r2 = static_cast<double>(z * fmaf(fmaf(fmaf(T.pc[3], z, T.pc[2]), z, T.pc[1]), z, T.pc[0]));
                }
                if (fallback)
                {
                    const Model4 m = B.models[P.hyp];
                    const double* F = cls ? B.orient : B.scale;
                    const double r = cls ? orientation_residual(F[3 * f], F[3 * f + 1], F[3 * f + 2], m)
                                         : scale_residual(F[3 * f], F[3 * f + 1], F[3 * f + 2], m, T.problem == 2);
                    r2 = r * r;
                    in = r2 <= T.tsq[cls]; // S1
                }
                if (in)
                {
                    q += static_cast<unsigned long long>(fmin(r2 / T.tsq[cls], 1.0) * kFix + 0.5);
                    cnt++;
                }
            }
        }
        if (active && cnt) // fixed point: the order of the additions does not matter
        {
            atomicAdd(&B.cont_q[2 * c + cls], q);
            atomicAdd(&B.cont_n[2 * c + cls], cnt);
        }
    }
}

// The same evaluation for a FEW contenders (<= kRefineFewMax): thread <-> feature, the unit's features held in registers while the contenders
// pass over them, one warp reduction + one pair of atomics per warp and contender.  The lane-per-contender form above executes all 32 lanes'
// worth of arithmetic per feature whatever the number of contenders (21-25 us at C3 for 3-12 of them); this one costs ~0.7 us per contender.
constexpr unsigned int kRefineFewMax = 24;
__device__ __forceinline__ void refine_phase_few(const Buffers& B, const Thresholds& T, unsigned int c_first, unsigned int c_end)
{
    constexpr int kCD = kChunk / 2;
    if (c_end <= c_first) return;
    const unsigned int fbs = (B.ns + kExactThreads - 1) / kExactThreads, fbo = (B.no + kExactThreads - 1) / kExactThreads, n_units = fbs + fbo;
    const int tid = threadIdx.x, lane = tid & 31;
    const double unit_s = static_cast<double>(T.scale_unit);
    for (unsigned int u = blockIdx.x; u < n_units; u += gridDim.x)
    {
        const int cls = u >= fbs;
        const int f = static_cast<int>((cls ? u - fbs : u) * kExactThreads) + tid;
        const int n = cls ? B.no : B.ns;
        double pa = 0.0, pb = 0.0, pc = 0.0;
        bool bad = false; // irregular feature (sentinel plane value): FP64 decides
        if (f < n)
        {
            const double* pl = (cls ? B.fod : B.fsd) + static_cast<size_t>(f / kCD) * (3 * kCD) + (f % kCD);
            pa = pl[0];
            pb = pl[kCD];
            pc = pl[2 * kCD];
            bad = cls ? (pb != pb) : !(pc < 1e300);
        }
        for (unsigned int c = c_first; c < c_end; c++)
        {
            const ContPar P = B.cont_par[c];
            if (P.exact) continue; // (uniform) already scored in FP64
            bool in = false, fallback = false;
            double r2 = 0.0;
            if (f < n)
            {
                if (cls == 0)
                {
                    const float v = static_cast<float>(fma(-P.h8, pb, fma(-P.h7, pa, pc)));
                    const float d = P.A - lg2_approx(v);
                    in = fabsf(d) <= P.Tin;
                    fallback = (!in && fabsf(d) <= P.Tout) || bad;
                    r2 = static_cast<double>(d) * static_cast<double>(d) * unit_s;
                }
                else
                {
                    const double nn = fma(pa, -P.h7, pb), dd = fma(pa, P.h8, pc);
                    const double c1 = fma(nn, P.sphi, dd * P.cphi), s1 = fma(nn, P.cphi, -(dd * P.sphi));
                    const float a1 = static_cast<float>(fabs(c1)), a2 = static_cast<float>(fabs(s1));
                    const float mn = fminf(a1, a2), mx = fmaxf(fmaxf(a1, a2), 1e-30f);
                    const float ri = rcp_approx(mx);
                    const float t = mn * ri;
                    const float band = fmaf(P.dpr, ri, P.band0r);
                    const bool wide = !(ri <= P.ri_cap);
                    in = !wide && t <= T.tanT - band;
                    fallback = (!in && t <= T.tanT + band) || bad;
                    const float z = t * t;
                    r2 = static_cast<double>(z * fmaf(fmaf(fmaf(T.pc[3], z, T.pc[2]), z, T.pc[1]), z, T.pc[0]));
                }
                if (fallback)
                {
                    const Model4 m = B.models[P.hyp];
                    const double* F = cls ? B.orient : B.scale;
                    const double r = cls ? orientation_residual(F[3 * f], F[3 * f + 1], F[3 * f + 2], m)
                                         : scale_residual(F[3 * f], F[3 * f + 1], F[3 * f + 2], m, T.problem == 2);
                    r2 = r * r;
                    in = r2 <= T.tsq[cls]; // S1
                }
            }
            unsigned long long q = in ? static_cast<unsigned long long>(fmin(r2 / T.tsq[cls], 1.0) * kFix + 0.5) : 0ull;
            const unsigned int cnt = __popc(__ballot_sync(0xffffffffu, in));
            if (cnt == 0u) continue; // (warp-uniform)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
            if (lane == 0)
            {
                atomicAdd(&B.cont_q[2 * c + cls], q);
                atomicAdd(&B.cont_n[2 * c + cls], cnt);
            }
        }
    }
}

// level 2, values: refined value (exact counts: zero-reset rule S3 applied) and its error bound per contender; best refined lower bound
__device__ __forceinline__ void contender_value_phase(const Buffers& B, const Thresholds& T, unsigned int c_first, unsigned int c_end)
{
    unsigned long long key = 0ull;
    for (unsigned int c = c_first + blockIdx.x * blockDim.x + threadIdx.x; c < c_end; c += gridDim.x * blockDim.x)
    {
        const ContPar P = B.cont_par[c];
        double v, err = 0.0;
        unsigned int n0, n1;
        if (P.exact) v = hyp_value(B, T, P.hyp, B.flags[P.hyp], n0, n1);
        else
        {
            v = msac_value(T, fix_to_double(B.cont_q[2 * c], T.tsq[0]), fix_to_double(B.cont_q[2 * c + 1], T.tsq[1]), B.cont_n[2 * c],
                           B.cont_n[2 * c + 1], n0, n1);
            err = static_cast<double>(n0) * P.e_s + static_cast<double>(n1) * P.e_o + 1e-12 * static_cast<double>(n0 + n1) + 1e-12;
        }
        B.cont_val[2 * c] = v;
        B.cont_val[2 * c + 1] = err;
        if (v > 0.0) // (a reset or inlier-free contender has no score)
        {
            const unsigned long long k = ordered_bits(v - err);
            key = k > key ? k : key;
        }
    }
    publish_key(&B.ts->lb_ref_key, key);
}

// level 2, values, FEW contenders (all of them fit one CTA's threads): every CTA computes every contender's refined value and the best refined
// lower bound for itself (a few loads per thread and one block reduction: cheaper than a pass of the grid and its barrier); CTA 0 also records
// them.  Thread c keeps contender c's value and bound for the collection that follows (collect_phase_w, few form).
__device__ __forceinline__ double contender_value_local(const Buffers& B, const Thresholds& T, unsigned int n_cont, double& v_c, double& err_c)
{
    __shared__ unsigned long long s_lbk[kExactThreads / 32];
    unsigned long long key = 0ull;
    v_c = 0.0;
    err_c = 0.0;
    const unsigned int c = threadIdx.x;
    if (c < n_cont)
    {
        const ContPar P = B.cont_par[c];
        unsigned int n0, n1;
        if (P.exact) v_c = hyp_value(B, T, P.hyp, B.flags[P.hyp], n0, n1);
        else
        {
            v_c = msac_value(T, fix_to_double(__ldcg(&B.cont_q[2 * c]), T.tsq[0]), fix_to_double(__ldcg(&B.cont_q[2 * c + 1]), T.tsq[1]),
                             __ldcg(&B.cont_n[2 * c]), __ldcg(&B.cont_n[2 * c + 1]), n0, n1);
            err_c = static_cast<double>(n0) * P.e_s + static_cast<double>(n1) * P.e_o + 1e-12 * static_cast<double>(n0 + n1) + 1e-12;
        }
        if (blockIdx.x == 0)
        {
            B.cont_val[2 * c] = v_c;
            B.cont_val[2 * c + 1] = err_c;
        }
        if (v_c > 0.0) key = ordered_bits(v_c - err_c);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other > key ? other : key;
    }
    __syncthreads(); // (s_lbk may still be read from an earlier round)
    if ((threadIdx.x & 31) == 0) s_lbk[threadIdx.x >> 5] = key;
    __syncthreads();
    for (unsigned int w = 0; w < blockDim.x / 32; w++) key = s_lbk[w] > key ? s_lbk[w] : key;
    if (blockIdx.x == 0 && threadIdx.x == 0 && key) atomicMax(&B.ts->lb_ref_key, key); // (for a later, crowded round's collection)
    return key ? from_ordered_bits(key) : -1.0;
}

// level 3, collection: contenders whose refined upper bound reaches the best refined lower bound (the reference's winner, and everything
// tied with it, is among them: its true value is >= every other true value >= every lower bound); after a contender overflow also every
// hypothesis that found no slot, by its loose bound.  Later rounds: the deferred ones (more than kCandCap candidates: massive ties).
__device__ __forceinline__ void collect_phase_w(const Buffers& B, const Thresholds& T, int round, unsigned int base, unsigned int n_cont,
                                                double floor_loose, bool direct, bool few, double v_c, double err_c, double lb_local)
{
    if (few && round == 0)
    {
        // thread c of CTA 0 holds contender c's refined value and bound (contender_value_local); no contender overflow in this form
        if (blockIdx.x == 0 && threadIdx.x < n_cont && v_c > 0.0 && v_c + err_c >= lb_local)
        {
            const unsigned int h = B.cont_par[threadIdx.x].hyp;
            collect_candidate(B, h, base, B.flags[h]);
        }
        return;
    }
    if (round > 0)
    {
        for (int h = blockIdx.x * blockDim.x + threadIdx.x; h < B.n_hyp; h += gridDim.x * blockDim.x)
        {
            const unsigned char f = B.flags[h];
            if ((f & 6) == 6 && (f & 16)) collect_candidate(B, h, base, f);
        }
        return;
    }
    const unsigned long long lk = __ldcg(&B.ts->lb_ref_key);
    const double best_lb = lk ? from_ordered_bits(lk) : 0.0;
    for (unsigned int c = blockIdx.x * blockDim.x + threadIdx.x; c < n_cont; c += gridDim.x * blockDim.x)
    {
        const double v = B.cont_val[2 * c], err = B.cont_val[2 * c + 1];
        if (direct || (v > 0.0 && v + err >= best_lb))
        {
            const unsigned int h = B.cont_par[c].hyp;
            collect_candidate(B, h, base, B.flags[h]);
        }
    }
    if (__ldcg(&B.ts->cont_overflow))
        for_each_scored_valid(B, [&](unsigned int h, unsigned char f) {
            if (f & 32) return;
            double err;
            const double v = hyp_value_w(B, T, h, f, err);
            if (v + err > 0.0 && v + err >= fmin(floor_loose, best_lb)) collect_candidate(B, h, base, f);
        });
}

struct Record // mirrors gcr_record
{
    double score;
    long long index;
    double model[4];
    unsigned int n_inl[2];
    unsigned long long reserved;
};

// ---- winner exchange across GPUs, executed by warp 0 of the selecting CTA (the "collective" of the path, fused into
// the kernel that produces the record): lane p stores this rank's record into peer p's buffer over NVLink
// (st.release.sys after the payload), then lane q waits for rank q's record in the local buffer (ld.acquire.sys).
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ bool record_better(double sa, long long ia, double sb, long long ib) // a beats b (S11 across ranks)
{
    if (!(sa > 0.0) || ia < 0) return false;
    if (!(sb > 0.0) || ib < 0) return true;
    return sa > sb || (sa == sb && ia < ib);
}

// Called by all 32 lanes of warp 0 with the local record in `w` (8 u64 words, same in every lane).  Returns the global
// winner's words and its rank; on a timeout (a peer never arrived: 4 s) keeps the local record and flags the run.
__device__ __forceinline__ void exchange_records(const Buffers& B, unsigned long long (&w)[8], unsigned int& winner_rank)
{
    const int lane = threadIdx.x & 31;
    const int world = B.xchg_world, me = B.xchg_rank;
    const unsigned long long seq = B.xchg_seq;
    const size_t par = static_cast<size_t>(seq & 1ull) * kMaxPeers * kXchgSlotWords;
    if (lane < world)
    {
        unsigned long long* dst = B.xchg_peer[lane] + par + static_cast<size_t>(me) * kXchgSlotWords;
#pragma unroll
        for (int k = 0; k < 8; k++) dst[k] = w[k];
        st_release_sys(dst + 8, seq);
    }
    bool timeout = false;
    unsigned long long mine[8];
#pragma unroll
    for (int k = 0; k < 8; k++) mine[k] = w[k];
    if (lane < world)
    {
        const unsigned long long* src = B.xchg_peer[me] + par + static_cast<size_t>(lane) * kXchgSlotWords;
        const unsigned long long t0 = global_ns();
        while (ld_acquire_sys(src + 8) < seq)
        {
            if (global_ns() - t0 > 4000000000ull) { timeout = true; break; }
            __nanosleep(64);
        }
        if (!timeout)
        {
#pragma unroll
            for (int k = 0; k < 8; k++) mine[k] = ld_relaxed_sys(src + k);
        }
    }
    timeout = __any_sync(0xffffffffu, timeout);
    // lane q holds rank q's record: butterfly arg-best over (score, index)
    double sc = lane < world ? __longlong_as_double(static_cast<long long>(mine[0])) : 0.0;
    long long ix = lane < world ? static_cast<long long>(mine[1]) : -1;
    unsigned int rk = lane;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
        const double os = __shfl_xor_sync(0xffffffffu, sc, o);
        const long long oi = __shfl_xor_sync(0xffffffffu, ix, o);
        const unsigned int orank = __shfl_xor_sync(0xffffffffu, rk, o);
        if (record_better(os, oi, sc, ix)) { sc = os; ix = oi; rk = orank; }
    }
    // every lane now agrees on rk (ties cannot occur: global indices are unique); fetch that lane's words
    if (!(sc > 0.0) || ix < 0) rk = static_cast<unsigned int>(me);
#pragma unroll
    for (int k = 0; k < 8; k++) w[k] = __shfl_sync(0xffffffffu, mine[k], static_cast<int>(rk));
    winner_rank = rk;
    if (timeout)
    {
        if (lane == 0) B.ctr->xchg_timeout = 1u;
        winner_rank = static_cast<unsigned int>(me);
#pragma unroll
        for (int k = 0; k < 8; k++) w[k] = __shfl_sync(0xffffffffu, mine[k], me); // lane `me` read its own slot (or kept w)
    }
}

// select (one block): final selection (GCRANSAC.h:440-462): highest exact score, earliest hypothesis on ties; writes
// the winner record (also into the caller's buffer, e.g. an NCCL send buffer) and the winner's packed masks; then
// snapshots the run counters into the result block for the host and clears them for the next run.
// One round of the selection: best of this round's candidates (score, then lowest hypothesis index: S11), merged with
// the running best of earlier rounds (kept in the result block).  `final`: no deferred candidates remain.
__device__ __forceinline__ void select_phase(const Buffers& B, const Thresholds& T, int round, bool final, unsigned int cand_base)
{
    __shared__ double s_val[kExactThreads / 32];
    __shared__ unsigned int s_h[kExactThreads / 32], s_i[kExactThreads / 32];
    __shared__ unsigned int s_hyp, s_slot;
    __shared__ unsigned long long s_rec[8];
    const unsigned int n = min(__ldcg(&B.ctr->n_candidates) - cand_base, static_cast<unsigned>(kCandCap));
    // every thread: value of its candidates and its own best (score, then lowest hypothesis index)
    double best = 0.0;
    unsigned int bh = 0xffffffffu, bs = 0;
    for (unsigned int i = threadIdx.x; i < n; i += blockDim.x)
    {
        cand_value(B, T, i);
        const double v = B.cand_score[i];
        const unsigned int h = B.cand_hyp[i];
        if (v > best || (v == best && v > 0.0 && h < bh)) { best = v; bh = h; bs = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o);
        const unsigned int oh = __shfl_xor_sync(0xffffffffu, bh, o), oi = __shfl_xor_sync(0xffffffffu, bs, o);
        if (ov > best || (ov == best && ov > 0.0 && oh < bh)) { best = ov; bh = oh; bs = oi; }
    }
    if ((threadIdx.x & 31) == 0)
    {
        s_val[threadIdx.x >> 5] = best;
        s_h[threadIdx.x >> 5] = bh;
        s_i[threadIdx.x >> 5] = bs;
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        for (unsigned int w = 1; w < blockDim.x / 32; w++)
            if (s_val[w] > best || (s_val[w] == best && best > 0.0 && s_h[w] < bh)) { best = s_val[w]; bh = s_h[w]; bs = s_i[w]; }
        Record r;
        r.score = 0.0;
        r.index = -1;
        r.model[0] = r.model[1] = r.model[2] = r.model[3] = 0.0;
        r.n_inl[0] = r.n_inl[1] = 0;
        r.reserved = 0;
        bool keep_running = false;
        if (round > 0)
        {
            // running best of the earlier rounds (its masks are already in the result block)
            const Record run = *reinterpret_cast<const Record*>(B.record);
            const long long cand_index = bh != 0xffffffffu ? static_cast<long long>(B.first_hyp + bh) : -1;
            if (run.index >= 0 && !(bh != 0xffffffffu && (best > run.score || (best == run.score && cand_index < run.index))))
            {
                r = run;
                keep_running = true;
            }
        }
        if (!keep_running && bh != 0xffffffffu)
        {
            const Model4 m = B.cand_model[bs];
            r.score = best;
            r.index = static_cast<long long>(B.first_hyp + bh);
            r.model[0] = m.h7; r.model[1] = m.h8; r.model[2] = m.alpha; r.model[3] = m.phi;
            r.n_inl[0] = B.cand_cnt[2 * bs];
            r.n_inl[1] = B.cand_cnt[2 * bs + 1];
        }
        s_hyp = keep_running ? 0xfffffffeu : bh; // 0xfffffffe: keep the masks of the running best
        s_slot = bs;
        *reinterpret_cast<Record*>(s_rec) = r;
    }
    __syncthreads();
    if (threadIdx.x < 32)
    {
        unsigned long long w[8];
#pragma unroll
        for (int k = 0; k < 8; k++) w[k] = s_rec[k];
        unsigned int winner_rank = B.xchg_world > 1 ? static_cast<unsigned int>(B.xchg_rank) : 0u;
        if (final && B.xchg_world > 1) exchange_records(B, w, winner_rank); // the GLOBAL winner replaces the local record
        if (threadIdx.x == 0)
        {
            if (final) w[7] = winner_rank; // gcr_record.reserved: rank that owns the winning hypothesis
            unsigned long long* out = reinterpret_cast<unsigned long long*>(B.record);
            unsigned long long* usr = reinterpret_cast<unsigned long long*>(B.record_user);
#pragma unroll
            for (int k = 0; k < 8; k++)
            {
                out[k] = w[k];
                if (final && usr) usr[k] = w[k];
            }
            if (final)
            {
                B.ctr->n_refined = min(B.ts->n_cont, B.cont_cap);
                *B.ctr_snap = B.ctr[0];
                Counters z;
                memset(&z, 0, sizeof(z));
                B.ctr[0] = z;
                TailState tz;
                memset(&tz, 0, sizeof(tz));
                B.ts[0] = tz;
            }
        }
    }
    __syncthreads();
    // The winner's packed masks -> the result block on the device and its mirror in page-locked host memory (the caller needs no
    // device-to-host copy after the run, only the stream synchronisation: [record 16 words][counters 16][phase stamps 16][winner masks]).
    // Eight words per thread are requested before any is stored: a load-store pair per iteration is one L2 round trip each, which at
    // C3's 2 189 mask words was 17 trips in a row through this one CTA (select 12.4 us against 7 us at C2's 79 words).
    const bool keep = s_hyp == 0xfffffffeu, have = s_hyp != 0xffffffffu;
    const bool mirror = final && B.host_block != nullptr;
    const unsigned int* src_masks = keep ? B.best_masks : B.cand_masks + static_cast<size_t>(s_slot) * B.wmask;
    if (!keep || mirror)
        for (int base = 0; base < B.wmask; base += 8 * static_cast<int>(blockDim.x))
        {
            unsigned int v[8];
#pragma unroll
            for (int q = 0; q < 8; q++)
            {
                const int w = base + q * static_cast<int>(blockDim.x) + static_cast<int>(threadIdx.x);
                v[q] = (have && w < B.wmask) ? __ldcg(src_masks + w) : 0u;
            }
#pragma unroll
            for (int q = 0; q < 8; q++)
            {
                const int w = base + q * static_cast<int>(blockDim.x) + static_cast<int>(threadIdx.x);
                if (w < B.wmask)
                {
                    if (!keep) B.best_masks[w] = v[q];
                    if (mirror) B.host_block[kResultHeaderWords + w] = v[q];
                }
            }
        }
    if (mirror)
    {
        const unsigned int* src = reinterpret_cast<const unsigned int*>(B.record); // (written by thread 0 before the barrier above)
        for (int w = threadIdx.x; w < 32; w += blockDim.x) B.host_block[w] = __ldcg(src + w);
    }
}

// The fused tail: fixup + whole-hypothesis FP64 items | finalize | collect | FP64 re-scoring of the candidates | select.
// Cooperative launch (all CTAs co-resident), kExactThreads threads per CTA.
__global__ void __launch_bounds__(kExactThreads) k_tail(const __grid_constant__ Buffers B, const __grid_constant__ Thresholds T)
{
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
#ifdef GCR_NO_TAIL_STAMPS
    const bool stamp = false; // (A/B build: what the phase stamps themselves cost)
#else
    const bool stamp = blockIdx.x == 0 && threadIdx.x == 0; // phase boundaries as seen by CTA 0 (diagnostics, ns)
#endif
    pdl_wait(); // results of the scoring kernel
    if (stamp) B.phase_ns[0] = global_ns();
    const bool winner = !B.per_hyp;
    // (every CTA reads the same counters: they were final when the scoring kernel ended)
    if (__ldcg(&B.ctr->n_worklist) | __ldcg(&B.ctr->n_hard) | B.scene->n_irr_s | B.scene->n_irr_o)
    {
        fixup_phase(B, T);
        exact_items_phase(B, T, 0, 0u);
        grid.sync();
    }
    if (stamp) B.phase_ns[1] = global_ns();
    unsigned int n_cont = 0u, spec_cands = 0u;
    double floor_loose = 0.0, v_c = 0.0, err_c = 0.0, lb_local = -1.0;
    bool direct = false, few = false;
    if (winner)
    {
        UbCache cache;
        finalize_phase_w(B, T, cache);
        grid.sync();
#ifdef GCR_TAIL_DEBUG
        if (stamp) B.phase_ns[2] = global_ns();
#endif
        // Contender rounds: one, unless the refinement lowers the bar (the loosely best hypothesis lost its score to the zero-reset rule, or
        // its refined lower bound falls below its loose one) or the slots run out.  Every CTA reads the same counters after each barrier.
        const unsigned long long k0 = __ldcg(&B.ctr->best_key);
        if (k0)
        {
            floor_loose = from_ordered_bits(k0) + B.dbg_floor_bump; // (tests raise the first bar to force a second contender round)
            for (;;)
            {
                contender_collect_phase(B, T, floor_loose, n_cont == 0u, n_cont == 0u ? &cache : nullptr);
                grid.sync();
#ifdef GCR_TAIL_DEBUG
                if (stamp) B.phase_ns[3] = global_ns();
#endif
                const unsigned int n_now = min(__ldcg(&B.ts->n_cont), B.cont_cap);
                const bool overflow = __ldcg(&B.ts->cont_overflow) != 0u;
                if (n_cont == 0u) spec_cands = min(__ldcg(&B.ts->n_cont), kDirectCand);
                // (... and whose FP64 scoring is at most one wave of the grid: one candidate of a 70 000-feature scene, sixteen of a 2 500-feature one)
                if (n_cont == 0u && n_now >= 1u && n_now <= kDirectCand && !overflow && !__ldcg(&B.ts->unproven) &&
                    static_cast<unsigned long long>(n_now) * static_cast<unsigned int>(B.ns + B.no) <= static_cast<unsigned long long>(gridDim.x) * blockDim.x)
                {
                    // a handful of contenders whose scores cannot be reset (small scenes: the loose bound is already tight): their FP64
                    // scoring is cheaper than the refinement round and its two barriers
                    n_cont = n_now;
                    direct = true;
                    break;
                }
                few = n_now - n_cont <= kRefineFewMax && n_now <= blockDim.x && !overflow;
                if (few) refine_phase_few(B, T, n_cont, n_now);
                else refine_phase(B, T, n_cont, n_now);
                grid.sync();
#ifdef GCR_TAIL_DEBUG
                if (stamp) B.phase_ns[4] = global_ns();
#endif
                double lb_ref;
                if (few) lb_ref = lb_local = contender_value_local(B, T, n_now, v_c, err_c); // (no barrier: every CTA has the same numbers)
                else
                {
                    contender_value_phase(B, T, n_cont, n_now);
                    grid.sync();
                    const unsigned long long lk = __ldcg(&B.ts->lb_ref_key);
                    lb_ref = lk ? from_ordered_bits(lk) : -1.0; // no refined contender has a score yet: everything positive stays in play
                }
                n_cont = n_now;
                if (overflow || lb_ref >= floor_loose) break;
                floor_loose = lb_ref;
            }
        }
    }
    else
    {
        finalize_phase(B, T);
        grid.sync();
    }
#ifndef GCR_TAIL_DEBUG
    if (stamp) B.phase_ns[2] = global_ns();
#endif
    // Candidate rounds: one, unless more than kCandCap hypotheses sit inside the window (massive ties).  The candidate
    // counters are cumulative; every CTA tracks the bases of the current round in registers (the counters are stable
    // between the barrier after a collect phase and the next collect phase, so all CTAs read the same values).
    unsigned int cand_base = direct ? 0u : spec_cands, ovf_base = 0u; // (speculative slots that were not used: a spent round)
    for (int round = 0;; round++)
    {
        if (!(direct && round == 0)) // direct: the contenders took their candidate slots when they were collected
        {
            if (winner) collect_phase_w(B, T, round, cand_base, n_cont, floor_loose, direct, few, v_c, err_c, lb_local);
            else collect_phase(B, T, round, cand_base);
            grid.sync();
        }
#ifndef GCR_TAIL_DEBUG
        if (stamp && round == 0) B.phase_ns[3] = global_ns();
#endif
        exact_items_phase(B, T, 1, cand_base);
        grid.sync();
#ifndef GCR_TAIL_DEBUG
        if (stamp && round == 0) B.phase_ns[4] = global_ns();
#endif
        // '>' (not '!='): in the last round the selecting CTA zeroes the counters while slower CTAs may still be reading
        const unsigned int cand_now = __ldcg(&B.ctr->n_candidates), ovf_now = __ldcg(&B.ctr->n_cand_overflow);
        const bool more = ovf_now > ovf_base;
        if (blockIdx.x == 0) select_phase(B, T, round, !more, cand_base);
        if (!more) break;
        cand_base = cand_now;
        ovf_base = ovf_now;
        grid.sync(); // the running best is in the result block before the next round's selection reads it
    }
    if (stamp)
    {
        B.phase_ns[5] = global_ns();
        if (B.host_block)
            for (int k = 0; k < 8; k++) reinterpret_cast<unsigned long long*>(B.host_block + 32)[k] = B.phase_ns[k];
    }
}

// ------------------------------------------------------------------------------------------------
// Per-hypothesis reporting modes only (GCR_RUN_KEEP_PER_HYP / GCR_RUN_EMIT_MASKS): the MSAC value n - sum(r^2)/tsq
// cancels when the inliers sit near the threshold, which amplifies the FP32 path's ~4e-6 n absolute error beyond the
// 1e-5 relative bar for low-score hypotheses (never contenders).  Those with value < 0.6 n are re-scored by the FP64
// reference-order kernel after the tail: collect -> k_exact_items(mode 0) -> refinalize.
__global__ void k_refine_collect(Buffers B, int refine_all)
{
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= B.n_hyp) return;
    const unsigned char f = B.flags[h];
    if (!(f & 2) || (f & 8)) return;
    const unsigned int n = B.n_inl[2 * h] + B.n_inl[2 * h + 1];
    if (n == 0u || !(refine_all || B.score[h] < 0.6 * static_cast<double>(n))) return;
    B.hard_hyp[atomicAdd(&B.ctr->n_hard, 1u)] = h;
    reinterpret_cast<unsigned long long*>(B.sums)[2 * h] = 0ull; // fixed-point accumulators of k_exact_items
    reinterpret_cast<unsigned long long*>(B.sums)[2 * h + 1] = 0ull;
    B.cnts[2 * h] = B.cnts[2 * h + 1] = 0u;
    B.flags[h] = f | 8;
}

__global__ void k_refine_finalize(Buffers B, Thresholds T)
{
    const unsigned int n = B.ctr->n_hard;
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        const unsigned int h = B.hard_hyp[i];
        unsigned int n0, n1;
        B.score[h] = msac_value(T, fix_to_double(reinterpret_cast<const unsigned long long*>(B.sums)[2 * h], T.tsq[0]),
                                fix_to_double(reinterpret_cast<const unsigned long long*>(B.sums)[2 * h + 1], T.tsq[1]),
                                B.cnts[2 * h], B.cnts[2 * h + 1], n0, n1);
        B.n_inl[2 * h] = n0;
        B.n_inl[2 * h + 1] = n1;
    }
}

__global__ void k_refine_done(Buffers B) { B.ctr->n_hard = 0u; }

// caller-supplied models -> candidate slots [0, n) (gcr_score_model / gcr_score_models)
__global__ void k_set_models(Buffers B, const Model4* __restrict__ m, unsigned int n)
{
    pdl_launch_dependents(); // chained launch: the next kernel of the call may start launching now; it waits at its own top
    pdl_wait();
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    B.cand_model[i] = m[i];
    B.cand_hyp[i] = i;
    B.cand_qsum[2 * i] = B.cand_qsum[2 * i + 1] = 0ull;
    B.cand_rawcnt[2 * i] = B.cand_rawcnt[2 * i + 1] = 0u;
}

// helper-math probe (gcr_debug_eval)
__global__ void k_debug_eval(int which, const double* in, int n, Model4 m, double* out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double a = in[3 * i], b = in[3 * i + 1], c = in[3 * i + 2];
    double r = 0.0;
    switch (which)
    {
    case 0: r = clip_angle(a); break;
    case 1: r = min_angle_diff(a, b); break;
    case 2: r = lines_angles_diff(a, b); break;
    case 3: r = scale_residual(a, b, c, m); break;
    default: r = orientation_residual(a, b, c, m); break;
    }
    out[i] = r;
}

} // namespace gcr
