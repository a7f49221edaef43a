// K3: the multi-class MSAC scoring kernel (the hot kernel of the engine).
//
// Work mapping: one thread <-> HB hypotheses held in registers; the features of both classes stream
// through a shared-memory ring filled by 1-D TMA bulk copies (cp.async.bulk + mbarrier full/empty pairs)
// and are read as broadcast 128-bit loads.  No reductions: every thread owns its hypotheses' partial sums.
// A work item is (tile of kTileHyps hypotheses) x (range of feature chunks); items are strided over a
// persistent grid, the split is chosen on the device from the number of solved hypotheses.
//
// Two tile flavours share the code (template PRECISE):
//   FP32 tiles     v, numer, denom from FP32 SoA features: the bulk of the hypotheses;
//   precise tiles  the two FMAs that suffer cancellation (v = 1/q - h7 x/q - h8 y/q for tiny alpha,
//                  numer/denom for large |h|) are done in FP64 from FP64 SoA features and rounded once,
//                  everything after is FP32.  They hold the hypotheses whose FP32 band would be wide.
// Both produce, per (hypothesis, feature): a certain inlier / certain outlier / "inside the proven error
// band" verdict; the last kind is appended to a worklist and decided in FP64 by k_fixup.
#pragma once
#include "gcr_kernels_common.cuh"

namespace gcr
{

// Diagnostic timeline (builds with -DGCR_TRACE only; tools/trace_score.py): per CTA (ring kernel) or per warp (resident
// kernel) a row of kTraceCols %globaltimer stamps: [0] = SM id, [1] = kernel entry, [2] = plan ready, then per work item
// (claimed, first features ready, done), 0-terminated.
#ifdef GCR_TRACE
constexpr int kTraceRows = 8192, kTraceCols = 64;
__device__ unsigned long long g_trace[kTraceRows * kTraceCols];
struct Tracer
{
    unsigned long long* row;
    int n;
    __device__ __forceinline__ Tracer(int r, bool on) : row(on && r < kTraceRows ? g_trace + static_cast<size_t>(r) * kTraceCols : nullptr), n(0)
    {
        if (row)
        {
            unsigned int smid;
            asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
            row[n++] = smid;
        }
    }
    __device__ __forceinline__ void stamp()
    {
        if (row && n < kTraceCols - 1)
        {
            row[n++] = global_ns();
            row[n] = 0;
        }
    }
};
#define GCR_TRACER(name, r, on) Tracer name(r, on)
#define GCR_STAMP(name) name.stamp()
#else
#define GCR_TRACER(name, r, on)
#define GCR_STAMP(name)
#endif

constexpr unsigned int kNoHyp = 0xffffffffu;
struct HypRegs
{
    float nh7, nh8, A, Tin, Tout, cphi, sphi, dpr;
    double h7d, h8d; // precise tiles only
    unsigned int hyp; // resident kernel: kNoHyp marks an inactive lane (padding of the last tile) instead of `active`: one long-lived
                      // predicate more made ptxas keep the band-check predicates in a GPR there (12 LOP3 per 16 scale evaluations)
    bool active;      // ring kernel
    double sum_s, sum_o;
    unsigned int cnt_s, cnt_o;
    float rimax; // winner-only runs: sum over the orientation chunks of (largest 1 / max(|c'|, |s'|) of the chunk's evaluations) x (features of a
                 // chunk), an upper bound of the sum of 1 / max over all orientation evaluations (scales the class value's error bound)
};

struct Verdict
{
    bool in, unc;
};

// rectified scale = v^-3, residual r = 3 ln2 |A - lg2 v|  (S4, two_sift.hpp:621-645).  v <= 0 (beyond the
// horizon: DBL_MAX residual in the reference) gives NaN / inf and fails both tests.
__device__ __forceinline__ Verdict scale_verdict_d(float d, const HypRegs& H)
{
    Verdict e;
    e.in = fabsf(d) <= H.Tin;
    e.unc = (fabsf(d) <= H.Tout) && !e.in;
    return e;
}
__device__ __forceinline__ Verdict scale_verdict(float v, const HypRegs& H, float& d)
{
    d = H.A - lg2_approx(v);
    return scale_verdict_d(d, H);
}

// (denom, numer) of rectifiedAngle (model.h:156-165) rotated by -phi; the residual (S5, two_sift.hpp:647-665)
// is the angle of that vector to the nearest multiple of pi/2 = atan(min/max) of its absolute components;
// r^2 = z p(z), z = tan^2, by an even polynomial.
__device__ __forceinline__ Verdict orient_verdict_cs(float c1, float s1, const HypRegs& H, const Thresholds& T, float& z, float& p, float& ri);
__device__ __forceinline__ Verdict orient_verdict(float n, float d, const HypRegs& H, const Thresholds& T, float& z,
                                                  float& p, float& ri)
{
    const float c1 = fmaf(n, H.sphi, d * H.cphi);
    const float s1 = fmaf(n, H.cphi, -(d * H.sphi));
    return orient_verdict_cs(c1, s1, H, T, z, p, ri);
}
// ri: 1 / max(|c'|, |s'|), the factor of the hypothesis' FP64-derived band dprime in this evaluation's error of t
__device__ __forceinline__ Verdict orient_verdict_cs(float c1, float s1, const HypRegs& H, const Thresholds& T, float& z, float& p, float& ri)
{
    const float a1 = fabsf(c1), a2 = fabsf(s1);
    // the 1e-30 floor keeps an exact 0/0 (numer = denom = 0, where the reference evaluates atan2 of round-off)
    // finite: t = 0 with an enormous band -> worklist
    const float mn = fminf(a1, a2), mx = fmaxf(fmaxf(a1, a2), 1e-30f);
    ri = rcp_approx(mx);
    const float t = mn * ri;
    Verdict e;
    e.in = t <= fmaf(-H.dpr, ri, T.t_lo0);               // tanT - band0 - dpr / max
    e.unc = (t <= fmaf(H.dpr, ri, T.t_hi0)) && !e.in;   // tanT + band0 + dpr / max
    z = t * t;
    // p = Q(z) of the monic form (Thresholds::pm): r^2 = pc_lead * z * Q(z); the lead factor is applied once per chunk
    if (T.poly_deg == 2) p = fmaf(z + T.pm[0], z, T.pm[1]); // uniform branch (kernel parameter)
    else p = fmaf(fmaf(z + T.pm[0], z, T.pm[1]), z, T.pm[2]);
    return e;
}

struct __align__(16) ScoreSmem
{
    float stage[kStages][kChunkFloats];
    uint64_t full[kStages];
    uint64_t empty[kStages];
};

__device__ __forceinline__ void worklist_push(const Buffers& B, unsigned int hyp, unsigned int feat_and_class)
{
    const unsigned int slot = atomicAdd(&B.ctr->n_worklist, 1u);
    if (slot < B.worklist_cap) B.worklist[slot] = make_uint2(hyp, feat_and_class);
    else atomicAdd(&B.ctr->n_worklist_overflow, 1u);
}

// One chunk of one class for the HB hypotheses of this thread.  CLS 0 = scale, 1 = orientation.
// FP32 chunk: 512 features, groups of 4 (float4 per plane).  Precise chunk: 256 features, groups of 2 (double2).
// RANGE: only the groups [g0r, g1r) of the chunk (the resident kernel's segments; both multiples of kScoreUnroll).
template <bool MASKS, bool PRECISE, int CLS, bool RANGE = false, int HB = 1>
__device__ __forceinline__ void score_chunk(const Buffers& B, const Thresholds& T, const float* stage, int fbase,
                                            HypRegs (&H)[HB], int g0r = 0, int g1r = 0)
{
    static_assert(kScoreUnroll % kVoteGroups == 0, "the vote window must divide the unroll factor");
    constexpr int G = PRECISE ? 2 : 4;             // features per group
    constexpr int NF = PRECISE ? kChunk / 2 : kChunk; // features per chunk
    float acc[HB], cntf[HB]; // per-chunk FP32 partials; counts are exact integers (< 2^24)
    float rsum[HB];
    unsigned int mword[HB];
#pragma unroll
    for (int j = 0; j < HB; j++)
    {
        acc[j] = cntf[j] = rsum[j] = 0.f;
        mword[j] = 0u;
    }
    bool unc_j[HB]; // hypothesis j had an evaluation of the last kVoteGroups groups inside the band (per thread)
#pragma unroll
    for (int j = 0; j < HB; j++) unc_j[j] = false;
    const int g0 = RANGE ? g0r : 0, g1 = RANGE ? g1r : NF / G;
    for (int gb = g0; gb < g1; gb += kScoreUnroll)
#pragma unroll
    for (int u = 0; u < kScoreUnroll; u++)
    {
        const int g = gb + u;
        float fa[G], fb[G], fc[G];   // FP32 planes
        double da[G], db[G], dc[G];  // FP64 planes (precise tiles)
        if constexpr (PRECISE)
        {
            const double2 a = reinterpret_cast<const double2*>(stage)[g];
            const double2 b = reinterpret_cast<const double2*>(stage + 2 * NF)[g];
            const double2 c = reinterpret_cast<const double2*>(stage + 4 * NF)[g];
            da[0] = a.x; da[1] = a.y; db[0] = b.x; db[1] = b.y; dc[0] = c.x; dc[1] = c.y;
        }
        else
        {
            const float4 a = reinterpret_cast<const float4*>(stage)[g];
            const float4 b = reinterpret_cast<const float4*>(stage + NF)[g];
            const float4 c = reinterpret_cast<const float4*>(stage + 2 * NF)[g];
            fa[0] = a.x; fa[1] = a.y; fa[2] = a.z; fa[3] = a.w;
            fb[0] = b.x; fb[1] = b.y; fb[2] = b.z; fb[3] = b.w;
            fc[0] = c.x; fc[1] = c.y; fc[2] = c.z; fc[3] = c.w;
        }
        // verdict of hypothesis j on one feature given its three plane values (FP32 or FP64 flavour)
        auto verdict = [&](int j, float a32, float b32, float c32, double a64, double b64, double c64, float& x0,
                           float& x1, float& x2) -> Verdict {
            if constexpr (CLS == 0)
            {
                float v;
                if constexpr (PRECISE) v = static_cast<float>(fma(-H[j].h8d, b64, fma(-H[j].h7d, a64, c64)));
                else v = fmaf(H[j].nh8, b32, fmaf(H[j].nh7, a32, c32));
                x1 = x2 = 0.f;
                return scale_verdict(v, H[j], x0);
            }
            else
            {
                float n, d;
                if constexpr (PRECISE)
                {
                    n = static_cast<float>(fma(a64, -H[j].h7d, b64)); // sin - k h7
                    d = static_cast<float>(fma(a64, H[j].h8d, c64));  // cos + k h8
                }
                else
                {
                    n = fmaf(a32, H[j].nh7, b32);
                    d = fmaf(-a32, H[j].nh8, c32);
                }
                return orient_verdict(n, d, H[j], T, x0, x1, x2);
            }
        };
#pragma unroll
        for (int j = 0; j < HB; j++)
        {
            // FP32 tiles, packed form (GCR_F2): the FMAs that do not depend on a predicate are issued as f32x2 pairs
            // (FFMA2 / FMUL2: two IEEE operations per issue slot, bit-identical to the scalar form)
            float pk0[G], pk1[G], pk2[G];
            Verdict pv[G];
            if constexpr (!PRECISE && kPacked<CLS>)
            {
#pragma unroll
                for (int e = 0; e < G; e += 2)
                {
                    const float2 a2 = make_float2(fa[e], fa[e + 1]), b2 = make_float2(fb[e], fb[e + 1]), c2 = make_float2(fc[e], fc[e + 1]);
                    if constexpr (CLS == 0)
                    {
                        const float2 v = __ffma2_rn(make_float2(H[j].nh8, H[j].nh8), b2,
                                                    __ffma2_rn(make_float2(H[j].nh7, H[j].nh7), a2, c2));
                        const float2 L = make_float2(lg2_approx(v.x), lg2_approx(v.y));
                        const float2 d = __ffma2_rn(L, make_float2(-1.f, -1.f), make_float2(H[j].A, H[j].A)); // A - L, one rounding
                        pk0[e] = d.x; pk0[e + 1] = d.y;
                        pk1[e] = pk1[e + 1] = 0.f;
                        pk2[e] = pk2[e + 1] = 0.f;
                        pv[e] = scale_verdict_d(d.x, H[j]);
                        pv[e + 1] = scale_verdict_d(d.y, H[j]);
                    }
                    else
                    {
                        // same operations, in the same order, as orient_verdict / orient_verdict_cs, two features per instruction
                        const float2 n = __ffma2_rn(a2, make_float2(H[j].nh7, H[j].nh7), b2);
                        const float2 d = __ffma2_rn(a2, make_float2(-H[j].nh8, -H[j].nh8), c2);
                        const float2 c1 = __ffma2_rn(n, make_float2(H[j].sphi, H[j].sphi), __fmul2_rn(d, make_float2(H[j].cphi, H[j].cphi)));
                        const float2 s1 = __ffma2_rn(n, make_float2(H[j].cphi, H[j].cphi), __fmul2_rn(d, make_float2(-H[j].sphi, -H[j].sphi)));
                        if constexpr (GCR_F2 >= 3)
                        {
                            const float a1x = fabsf(c1.x), a2x = fabsf(s1.x), a1y = fabsf(c1.y), a2y = fabsf(s1.y);
                            const float2 mn = make_float2(fminf(a1x, a2x), fminf(a1y, a2y));
                            const float2 ri = make_float2(rcp_approx(fmaxf(fmaxf(a1x, a2x), 1e-30f)), rcp_approx(fmaxf(fmaxf(a1y, a2y), 1e-30f)));
                            const float2 t = __fmul2_rn(mn, ri);
                            const float2 lo = __ffma2_rn(make_float2(-H[j].dpr, -H[j].dpr), ri, make_float2(T.t_lo0, T.t_lo0));
                            const float2 hi = __ffma2_rn(make_float2(H[j].dpr, H[j].dpr), ri, make_float2(T.t_hi0, T.t_hi0));
                            const float2 z = __fmul2_rn(t, t);
                            float2 pl; // Q(z) of the monic form, as in orient_verdict_cs
                            const float2 z0 = make_float2(z.x + T.pm[0], z.y + T.pm[0]);
                            if (T.poly_deg == 2)
                                pl = __ffma2_rn(z0, z, make_float2(T.pm[1], T.pm[1]));
                            else
                                pl = __ffma2_rn(__ffma2_rn(z0, z, make_float2(T.pm[1], T.pm[1])), z, make_float2(T.pm[2], T.pm[2]));
                            pk0[e] = z.x; pk0[e + 1] = z.y;
                            pk1[e] = pl.x; pk1[e + 1] = pl.y;
                            pk2[e] = ri.x; pk2[e + 1] = ri.y;
                            pv[e].in = t.x <= lo.x;
                            pv[e].unc = (t.x <= hi.x) && !pv[e].in;
                            pv[e + 1].in = t.y <= lo.y;
                            pv[e + 1].unc = (t.y <= hi.y) && !pv[e + 1].in;
                        }
                        else
                        {
                            // the rest of the evaluation stays scalar (see GCR_F2 in gcr_kernels_common.cuh)
                            pv[e] = orient_verdict_cs(c1.x, s1.x, H[j], T, pk0[e], pk1[e], pk2[e]);
                            pv[e + 1] = orient_verdict_cs(c1.y, s1.y, H[j], T, pk0[e + 1], pk1[e + 1], pk2[e + 1]);
                        }
                    }
                }
            }
#pragma unroll
            for (int e = 0; e < G; e++)
            {
                float x0, x1, x2;
                Verdict r;
                if constexpr (!PRECISE && kPacked<CLS>)
                {
                    x0 = pk0[e];
                    x1 = pk1[e];
                    x2 = pk2[e];
                    r = pv[e];
                }
                else
                    r = PRECISE ? verdict(j, 0.f, 0.f, 0.f, da[e], db[e], dc[e], x0, x1, x2)
                                : verdict(j, fa[e], fb[e], fc[e], 0.0, 0.0, 0.0, x0, x1, x2);
                if (r.in)
                {
                    acc[j] = CLS == 0 ? fmaf(x0, x0, acc[j]) : fmaf(x0, x1, acc[j]);
                    if (CLS == 1) rsum[j] += x2; // sum of 1 / max over the certain inliers: scales the value's error bound (finalize_phase)
                    cntf[j] += 1.0f;
                    if (MASKS) mword[j] |= 1u << ((g * G + e) & 31);
                }
                unc_j[j] |= r.unc;
            }
        }
        // the band check is voted once per kVoteGroups groups (the vote + branch cost ~5 issue slots)
        if (u % kVoteGroups == kVoteGroups - 1)
        {
            bool any_unc = false;
#pragma unroll
            for (int j = 0; j < HB; j++) any_unc |= unc_j[j];
            if (__any_sync(0xffffffffu, any_unc))
            {
                // rare: re-derive which pairs of these groups are inside the band (same instructions, same inputs,
                // re-read from shared memory so that this path stays a small rolled loop out of the hot stream)
                const int f0 = (g - (kVoteGroups - 1)) * G;
#pragma unroll
                for (int j = 0; j < HB; j++)
                {
                    if (!unc_j[j]) continue; // usually one hypothesis of one lane
#pragma unroll 1
                    for (int f = f0; f < f0 + kVoteGroups * G; f++)
                    {
                        float x0, x1, x2;
                        Verdict r;
                        if constexpr (PRECISE)
                        {
                            const double* sd = reinterpret_cast<const double*>(stage);
                            r = verdict(j, 0.f, 0.f, 0.f, sd[f], sd[NF + f], sd[2 * NF + f], x0, x1, x2);
                        }
                        else
                            r = verdict(j, stage[f], stage[NF + f], stage[2 * NF + f], 0.0, 0.0, 0.0, x0, x1, x2);
                        if (r.unc && (RANGE ? H[j].hyp != kNoHyp : H[j].active))
                            worklist_push(B, H[j].hyp, (CLS ? 0x80000000u : 0u) | static_cast<unsigned>(fbase + f));
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < HB; j++) unc_j[j] = false;
        }
        if (MASKS && ((g + 1) * G) % 32 == 0)
        {
#pragma unroll
            for (int j = 0; j < HB; j++)
            {
                if ((RANGE ? H[j].hyp != kNoHyp : H[j].active) && mword[j])
                    atomicOr(&B.masks[static_cast<size_t>(H[j].hyp) * B.wmask + (CLS ? B.ws : 0) + ((fbase + g * G) >> 5)],
                             mword[j]);
                mword[j] = 0u;
            }
        }
    }
#pragma unroll
    for (int j = 0; j < HB; j++)
    {
        if (CLS == 0)
        {
            H[j].sum_s += static_cast<double>(acc[j]);
            H[j].cnt_s += static_cast<unsigned int>(cntf[j]);
        }
        else
        {
            H[j].sum_o += static_cast<double>(acc[j]) * static_cast<double>(T.pc_lead);
            H[j].cnt_o += static_cast<unsigned int>(cntf[j]);
            H[j].rimax += rsum[j];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Winner-only runs (no per-hypothesis outputs asked for): the class value of MSAC is evaluated as the truncated quadratic
//     value_i = n_i - sum_inl r^2 / tsq_i = sum_f max(0, 1 - r_f^2 / tsq_i)          (MSAC_scoring_function.hpp:22-51,93-125)
// which is CONTINUOUS at the threshold: no inlier decision is taken here, so there is no guard band, no worklist and no count; an
// FP32 evaluation error moves the value by at most 2.1 (delta r) / sqrt(tsq) per feature near or inside the threshold (bound per
// hypothesis in k_sample_solve + the largest 1 / max met here).  The selection then re-scores every hypothesis whose
// upper bound reaches the best lower bound in FP64 reference order (counts, masks and the reference's value come from there), so the
// selected model is the reference's.  Per evaluation: scale 2 FFMA + MUFU.LG2 + FFMA + FFMA.SAT + 1/2 FADD2, orientation 6 FFMA/FMUL
// + FMNMX + FMNMX3.NAN + MUFU.RCP + 2 FMUL + 2-3 FFMA + FFMA.SAT + 1/2 FADD2 + 1/2 FMNMX3 (the FP32 tiles' scale FMAs as FFMA2 pairs).
__device__ __forceinline__ float fma_sat(float a, float b, float c) // saturating FMA: clamps to [0, 1], NaN -> 0 (SASS FFMA.SAT)
{
    float r;
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ float fmax_nan(float a, float b) // NaN-propagating maximum (SASS FMNMX[3].NAN)
{
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}

// (GCR_RCP2=1: ONE reciprocal per pair of orientation evaluations, 1/a = b rcp(ab).  It paid while the XU pipe was the tightest unit
//  (C3 5.60 -> 5.575 ms); with the pair arithmetic packed (GCR_WF2) the FMA pipe is, and the 1.5 extra multiplies cost more than half a
//  MUFU: 5.117 -> 4.976 ms with two reciprocals.)
#ifndef GCR_RCP2
#define GCR_RCP2 0
#endif
// Packed f32x2 arithmetic in the winner-only orientation evaluation: 0 none, 1 numer / denom / rotation of a PAIR of features as
// FFMA2 / FMUL2, 2 + t, z and the polynomial.  Same IEEE results, one issue slot for two features: 17.5 -> 12 issue slots per orientation
// evaluation; C3 kernel 5.371 / 5.181 / 5.109 ms for levels 0 / 1 / 2 (same box, interleaved), C2 28.7 -> 26.6 us.  (Round 1 found the pair
// forms useless in the reporting kernel, whose FMA pipe was the tighter unit; this loop is issue-bound.)
#ifndef GCR_WF2
#define GCR_WF2 2
#endif
template <bool PRECISE, int CLS, bool RANGE = false, int HB = 1>
__device__ __forceinline__ void score_chunk_w(const Thresholds& T, const float* stage, HypRegs (&H)[HB], int g0r = 0, int g1r = 0)
{
    constexpr int G = PRECISE ? 2 : 4;               // features per group
    constexpr int NF = PRECISE ? kChunk / 2 : kChunk; // features per chunk
    float2 acc[HB];                                  // two FP32 partial sums per hypothesis (even / odd features), flushed per chunk
    float rim[HB];                                   // largest 1 / max of this chunk's orientation evaluations
#pragma unroll
    for (int j = 0; j < HB; j++)
    {
        acc[j] = make_float2(0.f, 0.f);
        rim[j] = 0.f;
    }
    const int g0 = RANGE ? g0r : 0, g1 = RANGE ? g1r : NF / G;
    for (int gb = g0; gb < g1; gb += kScoreUnroll)
#pragma unroll
    for (int u = 0; u < kScoreUnroll; u++)
    {
        const int g = gb + u;
        float fa[4], fb[4], fc[4]; // the group's three plane values as FP32 (precise tiles: filled below, per hypothesis)
        double da[2], db[2], dc[2];
        if constexpr (PRECISE)
        {
            const double2 a = reinterpret_cast<const double2*>(stage)[g];
            const double2 b = reinterpret_cast<const double2*>(stage + 2 * NF)[g];
            const double2 c = reinterpret_cast<const double2*>(stage + 4 * NF)[g];
            da[0] = a.x; da[1] = a.y; db[0] = b.x; db[1] = b.y; dc[0] = c.x; dc[1] = c.y;
        }
        else
        {
            const float4 a = reinterpret_cast<const float4*>(stage)[g];
            const float4 b = reinterpret_cast<const float4*>(stage + NF)[g];
            const float4 c = reinterpret_cast<const float4*>(stage + 2 * NF)[g];
            fa[0] = a.x; fa[1] = a.y; fa[2] = a.z; fa[3] = a.w;
            fb[0] = b.x; fb[1] = b.y; fb[2] = b.z; fb[3] = b.w;
            fc[0] = c.x; fc[1] = c.y; fc[2] = c.z; fc[3] = c.w;
        }
#pragma unroll
        for (int j = 0; j < HB; j++)
        {
#pragma unroll
            for (int e = 0; e < G; e += 2)
            {
                float2 gq; // the pair's two contributions max(0, 1 - r^2 / tsq)
                if constexpr (CLS == 0)
                {
                    float2 v;
                    if constexpr (PRECISE)
                        v = make_float2(static_cast<float>(fma(-H[j].h8d, db[e], fma(-H[j].h7d, da[e], dc[e]))),
                                        static_cast<float>(fma(-H[j].h8d, db[e + 1], fma(-H[j].h7d, da[e + 1], dc[e + 1]))));
                    else
                        v = __ffma2_rn(make_float2(H[j].nh8, H[j].nh8), make_float2(fb[e], fb[e + 1]),
                                       __ffma2_rn(make_float2(H[j].nh7, H[j].nh7), make_float2(fa[e], fa[e + 1]), make_float2(fc[e], fc[e + 1])));
                    const float2 L = make_float2(lg2_approx(v.x), lg2_approx(v.y));
                    // dq = (A - lg2 v) / Tp in one rounding: H.A holds A / Tp here (k_sample_solve)
                    const float2 dq = __ffma2_rn(L, make_float2(-T.inv_Tp, -T.inv_Tp), make_float2(H[j].A, H[j].A));
                    gq = make_float2(fma_sat(-dq.x, dq.x, 1.0f), fma_sat(-dq.y, dq.y, 1.0f));
                }
                else
                {
                    float n[2], d[2], ri[2];
#pragma unroll
                    for (int q = 0; q < 2; q++)
                    {
                        if constexpr (PRECISE)
                        {
                            n[q] = static_cast<float>(fma(da[e + q], -H[j].h7d, db[e + q])); // sin - k h7
                            d[q] = static_cast<float>(fma(da[e + q], H[j].h8d, dc[e + q]));  // cos + k h8
                        }
                        else
                        {
                            n[q] = fmaf(fa[e + q], H[j].nh7, fb[e + q]);
                            d[q] = fmaf(-fa[e + q], H[j].nh8, fc[e + q]);
                        }
                    }
                    float gg[2], mn[2], mx[2];
                    float c1v[2], s1v[2];
                    if constexpr (GCR_WF2 && !PRECISE)
                    {
                        // numer, denom and the rotation by -phi of the PAIR as f32x2 operations (FFMA2 / FMUL2: same IEEE results, one issue
                        // slot for two features; the kernel is issue-bound, its FMA pipe has the cycles)
                        const float2 a2 = make_float2(fa[e], fa[e + 1]);
                        const float2 n2 = __ffma2_rn(a2, make_float2(H[j].nh7, H[j].nh7), make_float2(fb[e], fb[e + 1]));
                        const float2 d2 = __ffma2_rn(a2, make_float2(-H[j].nh8, -H[j].nh8), make_float2(fc[e], fc[e + 1]));
                        const float2 c2 = __ffma2_rn(n2, make_float2(H[j].sphi, H[j].sphi), __fmul2_rn(d2, make_float2(H[j].cphi, H[j].cphi)));
                        const float2 s2 = __ffma2_rn(n2, make_float2(H[j].cphi, H[j].cphi), __fmul2_rn(d2, make_float2(-H[j].sphi, -H[j].sphi)));
                        c1v[0] = c2.x; c1v[1] = c2.y; s1v[0] = s2.x; s1v[1] = s2.y;
                    }
                    else
                    {
#pragma unroll
                        for (int q = 0; q < 2; q++)
                        {
                            c1v[q] = fmaf(n[q], H[j].sphi, d[q] * H[j].cphi);
                            s1v[q] = fmaf(n[q], H[j].cphi, -(d[q] * H[j].sphi));
                        }
                    }
#pragma unroll
                    for (int q = 0; q < 2; q++)
                    {
                        const float c1 = c1v[q], s1 = s1v[q];
                        const float a1 = fabsf(c1), a2 = fabsf(s1);
                        // NaN-propagating maximum: a padding / irregular feature (NaN sine) gives ri = NaN, which the value (FFMA.SAT: NaN -> 0) and
                        // the running maximum below both ignore; an exact 0/0 of a real feature gives ri ~ 1e18, i.e. an unbounded error bound,
                        // and the hypothesis is then re-scored in FP64 whatever its value
                        mn[q] = fminf(a1, a2);
                        mx[q] = fmax_nan(fmax_nan(a1, a2), 1e-18f);
                    }
                    if constexpr (GCR_RCP2 && !RANGE)
                    {
                    // ONE reciprocal for the pair: 1/mx0 = mx1 / (mx0 mx1), 1/mx1 = mx0 / (mx0 mx1): the XU pipe (16 lanes per clock and SM) is
                    // what bounds this kernel, an FMUL is 8 times cheaper.  Padding / irregular features come in aligned pairs (k_prep_features),
                    // so a NaN never reaches a real feature's reciprocal.  mx in [1e-18, 1e6]: the product stays inside the FP32 range.
                    const float rp = rcp_approx(mx[0] * mx[1]);
                    const float2 r2 = __fmul2_rn(make_float2(rp, rp), make_float2(mx[1], mx[0]));
                    ri[0] = r2.x;
                    ri[1] = r2.y;
                    }
                    else // the resident kernel (one 1024-thread CTA per SM, small scenes) measured 6 % faster at C2 with two reciprocals
                    {
                        ri[0] = rcp_approx(mx[0]);
                        ri[1] = rcp_approx(mx[1]);
                    }
                    if constexpr (GCR_WF2 >= 2 && !PRECISE)
                    {
                        // t, z = t^2 and the polynomial of the pair as f32x2 operations too
                        const float2 t2 = __fmul2_rn(make_float2(mn[0], mn[1]), make_float2(ri[0], ri[1]));
                        const float2 z2 = __fmul2_rn(t2, t2);
                        float2 h2;
                        if (T.poly_deg == 2) // uniform branch (kernel parameter)
                            h2 = __ffma2_rn(__ffma2_rn(make_float2(T.pn[2], T.pn[2]), z2, make_float2(T.pn[1], T.pn[1])), z2, make_float2(T.pn[0], T.pn[0]));
                        else
                            h2 = __ffma2_rn(__ffma2_rn(__ffma2_rn(make_float2(T.pn[3], T.pn[3]), z2, make_float2(T.pn[2], T.pn[2])), z2,
                                                       make_float2(T.pn[1], T.pn[1])), z2, make_float2(T.pn[0], T.pn[0]));
                        gg[0] = fma_sat(-z2.x, h2.x, 1.0f);
                        gg[1] = fma_sat(-z2.y, h2.y, 1.0f);
                    }
                    else
                    {
#pragma unroll
                        for (int q = 0; q < 2; q++)
                        {
                            const float t = mn[q] * ri[q];
                            const float z = t * t;
                            float h;
                            if (T.poly_deg == 2) h = fmaf(fmaf(T.pn[2], z, T.pn[1]), z, T.pn[0]); // uniform branch (kernel parameter)
                            else h = fmaf(fmaf(fmaf(T.pn[3], z, T.pn[2]), z, T.pn[1]), z, T.pn[0]);
                            gg[q] = fma_sat(-z, h, 1.0f);
                        }
                    }
                    gq = make_float2(gg[0], gg[1]);
                    rim[j] = fmaxf(fmaxf(rim[j], ri[0]), ri[1]);
                }
                acc[j] = __fadd2_rn(acc[j], gq);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < HB; j++)
    {
        if (CLS == 0) H[j].sum_s += static_cast<double>(acc[j].x) + static_cast<double>(acc[j].y);
        else
        {
            H[j].sum_o += static_cast<double>(acc[j].x) + static_cast<double>(acc[j].y);
            H[j].rimax += rim[j] * static_cast<float>((g1 - g0) * G); // (round-off of this sum: 1e-7 relative, inside the bound's own margins)
        }
    }
}

// Hypothesis parameters of the HB hypotheses of a thread: compact index first_compact + j * stride + lane_in_tile of the FP32
// list (from the front of the compact arrays) or of the precise list (from the back).
template <bool PRECISE, bool SENTINEL = false, int HB = 1>
__device__ __forceinline__ void load_hyps(const Buffers& B, HypRegs (&H)[HB], int first_compact, int stride, int t, int n_list,
                                          bool from_back)
{
#pragma unroll
    for (int j = 0; j < HB; j++)
    {
        const int k = first_compact + j * stride + t;
        const bool active = k < n_list;
        const int c = from_back ? B.n_hyp - 1 - k : k;
        // inactive lanes: parameters for which every comparison is false
        const float4 p0 = active ? B.pf0[c] : make_float4(0.f, 0.f, 0.f, -1.f);
        const float4 p1 = active ? B.pf1[c] : make_float4(-1.f, __int_as_float(0x7fc00000), 0.f, 0.f);
        H[j].hyp = active ? B.fast_hyp[c] : (SENTINEL ? kNoHyp : 0u);
        H[j].active = SENTINEL ? true : active;
        H[j].nh7 = p0.x; H[j].nh8 = p0.y; H[j].A = p0.z; H[j].Tin = p0.w;
        H[j].Tout = p1.x; H[j].cphi = p1.y; H[j].sphi = p1.z; H[j].dpr = p1.w;
        H[j].h7d = H[j].h8d = 0.0;
        if (PRECISE && active)
        {
            const Model4 m = B.models[H[j].hyp];
            H[j].h7d = m.h7;
            H[j].h8d = m.h8;
        }
        H[j].sum_s = H[j].sum_o = 0.0;
        H[j].cnt_s = H[j].cnt_o = 0u;
        H[j].rimax = 0.f;
    }
}

// Partial sums / counts of a thread's hypotheses into the per-hypothesis arrays: plain stores when this item covered every
// feature, atomics when the feature range is split over several items.
template <bool SENTINEL = false, bool WINNER = false, int HB = 1>
__device__ __forceinline__ void flush_hyps(const Buffers& B, const Thresholds& T, HypRegs (&H)[HB], bool single_split)
{
#pragma unroll
    for (int j = 0; j < HB; j++)
    {
        if (SENTINEL ? H[j].hyp == kNoHyp : !H[j].active) continue;
        const size_t h = H[j].hyp;
        if constexpr (WINNER)
        {
            // class values (sums) and the bound of the sum of 1 / max over the orientation evaluations (a float in cnts[2h + 1])
            if (single_split)
            {
                B.sums[2 * h] = H[j].sum_s;
                B.sums[2 * h + 1] = H[j].sum_o;
                B.cnts[2 * h + 1] = __float_as_uint(H[j].rimax);
            }
            else
            {
                if (H[j].sum_s != 0.0) atomicAdd(&B.sums[2 * h], H[j].sum_s);
                if (H[j].sum_o != 0.0) atomicAdd(&B.sums[2 * h + 1], H[j].sum_o);
                if (H[j].rimax > 0.f) atomicAdd(reinterpret_cast<float*>(&B.cnts[2 * h + 1]), H[j].rimax);
            }
            continue;
        }
        const double ss = H[j].sum_s * static_cast<double>(T.scale_unit);
        // (reporting modes: the sum of 1 / max over the certain orientation inliers travels in the low word of the hypothesis' score slot,
        //  which k_sample_solve cleared and finalize_phase reads before it writes the score)
        float* risum = reinterpret_cast<float*>(B.score + h);
        if (single_split)
        {
            B.sums[2 * h] = ss;
            B.sums[2 * h + 1] = H[j].sum_o;
            B.cnts[2 * h] = H[j].cnt_s;
            B.cnts[2 * h + 1] = H[j].cnt_o;
            *risum = H[j].rimax;
        }
        else
        {
            if (H[j].cnt_s)
            {
                atomicAdd(&B.sums[2 * h], ss);
                atomicAdd(&B.cnts[2 * h], H[j].cnt_s);
            }
            if (H[j].cnt_o)
            {
                atomicAdd(&B.sums[2 * h + 1], H[j].sum_o);
                atomicAdd(&B.cnts[2 * h + 1], H[j].cnt_o);
                atomicAdd(risum, H[j].rimax);
            }
        }
    }
}

// Position in a tile flavour's chunk list -> (class, chunk of that class).  GCR_INTERLEAVE=1 (default) spreads the orientation chunks evenly
// among the scale chunks (Bresenham), so that every work item -- and, with items of many CTAs in flight, every SM at every moment -- carries
// the scene's mix of the two evaluation kinds: the scale evaluation is bound by the XU pipe (one MUFU per ~5 issue slots), the orientation
// evaluation by issue slots (one MUFU per ~16); side by side they fill both.  0: all scale chunks, then all orientation chunks.
#ifndef GCR_INTERLEAVE
#define GCR_INTERLEAVE 1
#endif
struct ChunkWalk // consecutive positions of the list, one division at the start (a 64-bit division per chunk cost ~10 % of a chunk's instructions)
{
    int ncs, nco, pos, o, acc; // o = floor(pos * nco / n), acc = (pos * nco) mod n
    __device__ __forceinline__ ChunkWalk(int ncs_, int nco_, int p) : ncs(ncs_), nco(nco_), pos(p)
    {
        const unsigned long long t = static_cast<unsigned long long>(p) * static_cast<unsigned int>(nco_);
        const unsigned int n = static_cast<unsigned int>(ncs_ + nco_);
        o = static_cast<int>(t / n);
        acc = static_cast<int>(t - static_cast<unsigned long long>(o) * n);
    }
    // class and chunk-of-class of the current position; advances
    __device__ __forceinline__ int next(bool& orient)
    {
#if GCR_INTERLEAVE
        acc += nco;
        orient = acc >= ncs + nco;
        const int ci = orient ? o : pos - o;
        if (orient)
        {
            acc -= ncs + nco;
            o++;
        }
        pos++;
        return ci;
#else
        orient = pos >= ncs;
        const int ci = orient ? pos - ncs : pos;
        pos++;
        return ci;
#endif
    }
};

// One work item: all chunks [c_begin, c_end) of this tile flavour's chunk list.
// MODE: 0 winner-only (score_chunk_w), 1 per-hypothesis counts and sums (score_chunk), 2 as 1 + packed inlier masks of every hypothesis
template <int MODE, bool PRECISE, int HB>
__device__ __forceinline__ void score_item(const Buffers& B, const Thresholds& T, ScoreSmem& S, uint32_t& ring,
                                           int first_compact, int n_list, bool from_back, int c_begin, int c_end,
                                           bool single_split)
{
    const int tid = threadIdx.x, lane = tid & 31;
    const int ncs = PRECISE ? B.ncsd : B.ncs, nco_list = PRECISE ? B.ncod : B.nco;
    const float* base_s = PRECISE ? reinterpret_cast<const float*>(B.fsd) : B.fs;
    const float* base_o = PRECISE ? reinterpret_cast<const float*>(B.fod) : B.fo;
    constexpr int NF = PRECISE ? kChunk / 2 : kChunk;
    const int nch = c_end - c_begin;

    ChunkWalk produce(ncs, nco_list, c_begin), consume(ncs, nco_list, c_begin); // the producer runs kStages - 1 chunks ahead
    auto issue = [&](uint32_t n) {
        if (tid == 0)
        {
            const uint32_t st = n % kStages;
            if (n >= kStages) mbar_wait(&S.empty[st], ((n / kStages) - 1) & 1);
            bool orient;
            const int ci = produce.next(orient);
            const float* src = orient ? base_o + static_cast<size_t>(ci) * kChunkFloats : base_s + static_cast<size_t>(ci) * kChunkFloats;
            mbar_expect_tx(&S.full[st], kChunkBytes);
            tma_load_1d(S.stage[st], src, kChunkBytes, &S.full[st]);
        }
    };
    for (int p = 0; p < min(kStages - 1, nch); p++) issue(ring + p);

    HypRegs H[HB];
    load_hyps<PRECISE, false, HB>(B, H, first_compact, kScoreThreads, tid, n_list, from_back);

    for (int i = 0; i < nch; i++)
    {
        if (i + kStages - 1 < nch) issue(ring + i + kStages - 1);
        const uint32_t n = ring + i;
        const uint32_t st = n % kStages;
        mbar_wait(&S.full[st], (n / kStages) & 1);
        bool orient;
        const int ci = consume.next(orient);
        if constexpr (MODE == 0)
        {
            if (!orient) score_chunk_w<PRECISE, 0, false, HB>(T, S.stage[st], H);
            else score_chunk_w<PRECISE, 1, false, HB>(T, S.stage[st], H);
        }
        else
        {
            if (!orient) score_chunk<MODE == 2, PRECISE, 0, false, HB>(B, T, S.stage[st], ci * NF, H);
            else score_chunk<MODE == 2, PRECISE, 1, false, HB>(B, T, S.stage[st], ci * NF, H);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&S.empty[st]);
    }
    ring += nch;

    flush_hyps<false, MODE == 0, HB>(B, T, H, single_split);
}

// hypotheses per thread / CTAs per SM of the ring kernel: 2 / GCR_MINB (6) in winner-only runs, 1 / 8 in the reporting modes
template <int MODE> constexpr int kRingHB = MODE == 0 ? GCR_HB : 1;
template <int MODE> constexpr int kRingMinB = MODE == 0 ? GCR_MINB : 8;
template <int MODE>
__global__ void __launch_bounds__(kScoreThreads, kRingMinB<MODE>) k_score_fast(const __grid_constant__ Buffers B,
                                                              const __grid_constant__ Thresholds T)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    ScoreSmem& S = *reinterpret_cast<ScoreSmem*>(smem_raw);
    if (threadIdx.x == 0)
    {
        if (blockIdx.x == 0) B.phase_ns[7] = global_ns(); // diagnostics
        for (int s = 0; s < kStages; s++)
        {
            mbar_init(&S.full[s], 1);
            mbar_init(&S.empty[s], kScoreThreads / 32);
        }
        fence_barrier_init();
    }
    GCR_TRACER(tr, blockIdx.x, threadIdx.x == 0);
    GCR_STAMP(tr);
    pdl_launch_dependents();
    __syncthreads();
    pdl_wait(); // compact lists and counters of k_sample_solve

    const int n_fast = static_cast<int>(B.ctr->n_fast), n_sens = static_cast<int>(B.ctr->n_sens);
    constexpr int HB = kRingHB<MODE>, kTileHyps = kScoreThreads * HB;
    const int tiles_a = (n_fast + kTileHyps - 1) / kTileHyps, tiles_b = (n_sens + kTileHyps - 1) / kTileHyps;
    const int chunks_a = B.ncs + B.nco, chunks_b = B.ncsd + B.ncod;
    // Work decomposition, identical in every CTA: the chunk lists are cut so that there are ~8 items per resident
    // CTA; items are handed out dynamically (atomic counter), costliest first: precise tiles, then FP32 tiles from
    // the orientation end of the chunk list (an orientation evaluation costs about twice a scale evaluation).
    const int tiles = tiles_a + 2 * tiles_b; // a precise tile has twice the chunks
    // (twice the items where they stay >= 4 chunks long: C3's 1 M hypotheses 4.920 -> 4.865 ms; shorter items cost more in hand-out and
    //  ring restarts than they level: a 125 000-hypothesis shard 0.640 -> 0.646 ms with 48 per CTA, same-box A/B)
    const bool long_items = static_cast<long long>(tiles) * chunks_a >= 8ll * kItemsPerCta * static_cast<long long>(gridDim.x);
    const int target = (long_items ? 2 : 1) * kItemsPerCta * static_cast<int>(gridDim.x);
    int splits_a = 1;
    if (tiles > 0 && tiles < target) splits_a = min(chunks_a, (target + tiles - 1) / tiles);
    const int cps_a = (chunks_a + splits_a - 1) / splits_a;
    splits_a = (chunks_a + cps_a - 1) / cps_a;
    const int cps_b = min(chunks_b, 2 * cps_a);
    const int splits_b = (chunks_b + cps_b - 1) / cps_b;
    const int items_b = tiles_b * splits_b, n_items = items_b + tiles_a * splits_a;

    GCR_STAMP(tr);
    __shared__ int s_item;
    uint32_t ring = 0; // chunks issued / consumed so far by this CTA (ring position, same in every thread)
    for (;;)
    {
        if (threadIdx.x == 0) s_item = static_cast<int>(atomicAdd(&B.ctr->next_item, 1u));
        __syncthreads();
        const int item = s_item;
        __syncthreads();
        if (item >= n_items) break;
        GCR_STAMP(tr);
        if (item >= items_b)
        {
            const int k = item - items_b;
            const int split = splits_a - 1 - k / tiles_a, tile = k % tiles_a;
            score_item<MODE, false, HB>(B, T, S, ring, tile * kTileHyps, n_fast, false, split * cps_a,
                                     min(chunks_a, (split + 1) * cps_a), splits_a == 1);
        }
        else
        {
            const int split = splits_b - 1 - item / tiles_b, tile = item % tiles_b;
            score_item<MODE, true, HB>(B, T, S, ring, tile * kTileHyps, n_sens, true, split * cps_b,
                                    min(chunks_b, (split + 1) * cps_b), splits_b == 1);
        }
        GCR_STAMP(tr);
    }
}

// ------------------------------------------------------------------------------------------------
// Resident form of the scoring kernel, for feature sets whose FP32 and FP64 working sets fit in one SM's shared
// memory together (<= kResidentMaxChunks chunks of 3 KB: ~5 000 features; configs C1, C2, C4).  One CTA of 1024
// threads per SM copies the whole working set into shared memory once (four TMA bulk copies on three mbarriers:
// FP32 scale, FP32 orientation, FP64 planes) and every WARP then works on its own, with no block barrier, ring
// protocol or work counter after the prologue.  The work is partitioned STATICALLY and by cost: hypotheses in tiles
// of 32 (precise tiles first), each tile a line of feature blocks (32 features) measured in cost units (scale block
// 10, orientation block cost_o, everything times cost_p / 10 in a precise tile); the lines of all tiles, laid end to
// end, are cut into one piece of equal cost per resident warp, so a warp scores at most a tail of one tile, some
// whole tiles and a head of another, and all warps finish together (the dynamic items of the ring kernel leave a
// 10 us drain at C2 because its costliest items last 30 of the kernel's 40 us).  Arithmetic, verdicts, worklist and
// mask emission are score_chunk's, i.e. bit-identical to the ring kernel.
constexpr int kResidentThreads = 1024;
// The winner-only form keeps two hypotheses per thread here too (every LDS.128 of the feature planes serves both, as in the ring kernel),
// in a CTA of fewer threads so that the registers are there: GCR_RES_HB hypotheses per thread, GCR_RES_THREADS threads.  Same-box A/B of
// (1, 1024) | (2, 768) | (2, 640) | (2, 512): C2 kernel 27.6 | 27.6 | 25.7 | 25.6 us, C4's image 50.2 us all four.
#ifndef GCR_RES_HB
#define GCR_RES_HB 2
#endif
#ifndef GCR_RES_THREADS
#define GCR_RES_THREADS 512
#endif
template <int MODE> constexpr int kResHB = MODE == 0 ? GCR_RES_HB : 1;
template <int MODE> constexpr int kResThreads = (MODE == 0 && GCR_RES_HB > 1) ? GCR_RES_THREADS : kResidentThreads;
constexpr int kResidentMaxChunks = 75; // 75 x 3 KB + barriers <= 227 KB
constexpr int kCostS = 10, kMinPieceCost = 40;

struct ResidentPlan // identical in every warp (pure function of the counters and the feature counts)
{
    int nbs, nbo;       // blocks of 32 features per class
    int cost_o;         // cost of an orientation block (scale block = kCostS)
    int total;          // cost of the feature line of one FP32 tile
    __device__ __forceinline__ int block_at(long long c) const // cost offset inside a tile's line -> block (monotone, onto)
    {
        const int cs = kCostS * nbs;
        return c <= cs ? static_cast<int>((c + kCostS / 2) / kCostS) : nbs + static_cast<int>((c - cs + cost_o / 2) / cost_o);
    }
};

template <int MODE, bool PRECISE>
__device__ __forceinline__ void resident_task(const Buffers& B, const Thresholds& T, const float* sm_s, const float* sm_o,
                                              const ResidentPlan& P, int tile, int b0, int b1, int n_list,
                                              uint64_t* bar_s, uint64_t* bar_o)
{
    constexpr int NF = PRECISE ? kChunk / 2 : kChunk; // features per chunk
    constexpr int G = PRECISE ? 2 : 4;
    constexpr int BPC = NF / 32, GPB = 32 / G;         // blocks per chunk, groups per block
    constexpr int HB = kResHB<MODE>; // (the reporting forms: one hypothesis per thread, 1024-thread CTAs leave 64 registers)
    HypRegs H[HB];
    load_hyps<PRECISE, true, HB>(B, H, tile * 32 * HB, 32, threadIdx.x & 31, n_list, PRECISE);
    const int s1 = min(b1, P.nbs);
    if (b0 < s1)
    {
        mbar_wait(bar_s, 0);
        for (int b = b0; b < s1;)
        {
            const int chunk = b / BPC, be = min(s1, (chunk + 1) * BPC);
            if constexpr (MODE == 0)
                score_chunk_w<PRECISE, 0, true, HB>(T, sm_s + chunk * kChunkFloats, H, (b - chunk * BPC) * GPB, (be - chunk * BPC) * GPB);
            else
                score_chunk<MODE == 2, PRECISE, 0, true>(B, T, sm_s + chunk * kChunkFloats, chunk * NF, H, (b - chunk * BPC) * GPB,
                                                         (be - chunk * BPC) * GPB);
            b = be;
        }
    }
    const int o0 = max(b0, P.nbs) - P.nbs, o1 = b1 - P.nbs;
    if (o0 < o1)
    {
        mbar_wait(bar_o, 0);
        for (int b = o0; b < o1;)
        {
            const int chunk = b / BPC, be = min(o1, (chunk + 1) * BPC);
            if constexpr (MODE == 0)
                score_chunk_w<PRECISE, 1, true, HB>(T, sm_o + chunk * kChunkFloats, H, (b - chunk * BPC) * GPB, (be - chunk * BPC) * GPB);
            else
                score_chunk<MODE == 2, PRECISE, 1, true>(B, T, sm_o + chunk * kChunkFloats, chunk * NF, H, (b - chunk * BPC) * GPB,
                                                         (be - chunk * BPC) * GPB);
            b = be;
        }
    }
    flush_hyps<true, MODE == 0, HB>(B, T, H, b0 == 0 && b1 == P.nbs + P.nbo); // plain stores when this warp covered the whole tile
}

template <int MODE>
__global__ void __launch_bounds__(kResThreads<MODE>, 1) k_score_resident(const __grid_constant__ Buffers B,
                                                                        const __grid_constant__ Thresholds T,
                                                                        const int cost_o, const int cost_p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* sm_fs = reinterpret_cast<float*>(smem_raw);
    float* sm_fo = sm_fs + B.ncs * kChunkFloats;
    float* sm_fsd = sm_fo + B.nco * kChunkFloats;
    float* sm_fod = sm_fsd + B.ncsd * kChunkFloats;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sm_fod + B.ncod * kChunkFloats); // [0] FP32 scale [1] FP32 orientation [2] FP64
    if (threadIdx.x == 0)
    {
        if (blockIdx.x == 0) B.phase_ns[7] = global_ns(); // diagnostics
        for (int i = 0; i < 3; i++) mbar_init(&bars[i], 1);
        fence_barrier_init();
        const uint32_t b_fs = B.ncs * kChunkBytes, b_fo = B.nco * kChunkBytes, b_fsd = B.ncsd * kChunkBytes,
                       b_fod = B.ncod * kChunkBytes;
        mbar_expect_tx(&bars[0], b_fs);
        if (b_fs) tma_load_1d(sm_fs, B.fs, b_fs, &bars[0]);
        mbar_expect_tx(&bars[1], b_fo);
        if (b_fo) tma_load_1d(sm_fo, B.fo, b_fo, &bars[1]);
        mbar_expect_tx(&bars[2], b_fsd + b_fod);
        if (b_fsd) tma_load_1d(sm_fsd, B.fsd, b_fsd, &bars[2]);
        if (b_fod) tma_load_1d(sm_fod, B.fod, b_fod, &bars[2]);
    }
    GCR_TRACER(tr, (threadIdx.x >> 5) * gridDim.x + blockIdx.x, (threadIdx.x & 31) == 0);
    GCR_STAMP(tr);
    pdl_launch_dependents();
    __syncthreads();
    pdl_wait(); // compact lists and counters of k_sample_solve (the features staged above are older than that kernel)

    ResidentPlan P;
    const int n_fast = static_cast<int>(B.ctr->n_fast), n_sens = static_cast<int>(B.ctr->n_sens);
    P.nbs = (B.ns + 31) / 32;
    P.nbo = (B.no + 31) / 32;
    P.cost_o = cost_o;
    P.total = kCostS * P.nbs + cost_o * P.nbo;
    constexpr int kTile = 32 * kResHB<MODE>; // hypotheses of a warp's tile
    const int tiles_a = (n_fast + kTile - 1) / kTile, tiles_b = (n_sens + kTile - 1) / kTile;
    // the line of all tiles in units of 1/10 cost: a precise tile is cost_p * total long, an FP32 tile 10 * total
    const long long len_b = static_cast<long long>(cost_p) * P.total, len_a = 10ll * P.total;
    const long long end_b = len_b * tiles_b, end_all = end_b + len_a * tiles_a;
    const int n_warps = static_cast<int>(gridDim.x) * (kResThreads<MODE> / 32);
    const long long piece = max((end_all + n_warps - 1) / n_warps, 10ll * kMinPieceCost);
    // consecutive pieces go to different SMs
    const int w = static_cast<int>(threadIdx.x >> 5) * static_cast<int>(gridDim.x) + static_cast<int>(blockIdx.x);
    long long pos = piece * w;
    const long long pos_end = min(pos + piece, end_all);
    GCR_STAMP(tr);
    while (pos < pos_end)
    {
        GCR_STAMP(tr);
        const bool precise = pos < end_b;
        const long long len = precise ? len_b : len_a, rel = precise ? pos : pos - end_b;
        const int tile = static_cast<int>(rel / len);
        const long long off0 = rel - len * tile, off1 = min(len, off0 + (pos_end - pos));
        // offsets in a tile's line -> blocks; the same rounding on both sides of a cut, so the pieces tile the line exactly
        const int b0 = P.block_at(off0 * P.total / len), b1 = P.block_at(off1 * P.total / len);
        if (b0 < b1)
        {
            if (precise) resident_task<MODE, true>(B, T, sm_fsd, sm_fod, P, tile, b0, b1, n_sens, &bars[2], &bars[2]);
            else resident_task<MODE, false>(B, T, sm_fs, sm_fo, P, tile, b0, b1, n_fast, &bars[0], &bars[1]);
        }
        pos += off1 - off0;
        GCR_STAMP(tr);
    }
    // no CTA may retire while its bulk copies are in flight (a CTA without work gets here at once)
    if (threadIdx.x == 0)
        for (int i = 0; i < 3; i++) mbar_wait(&bars[i], 0);
}

} // namespace gcr
