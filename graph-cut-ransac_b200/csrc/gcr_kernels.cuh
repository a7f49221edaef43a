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
#include <cuda_runtime.h>
#include <cstdint>
#include "gcr_exact.cuh"

namespace gcr
{

constexpr int kChunk = 512;             // features per shared-memory stage
constexpr int kChunkFloats = 3 * kChunk; // three SoA planes per chunk
constexpr int kChunkBytes = kChunkFloats * 4;
constexpr int kStages = 4;
constexpr int kScoreThreads = 128;
constexpr int kHB = 2; // hypotheses per thread in the fast kernel
constexpr int kTileHyps = kScoreThreads * kHB;
constexpr int kCandCap = 1024;
constexpr int kIrrCap = 4096;
constexpr double kLn2 = 0.693147180559945309417;
constexpr double kU = 5.9604644775390625e-8; // 2^-24, FP32 unit round-off

struct Counters
{
    unsigned int n_sample_valid, n_solved, n_fast, n_hard;
    unsigned int n_worklist, n_worklist_overflow, n_candidates, n_cand_overflow;
    unsigned int n_bad_sample, n_irregular_evals, blocks_done, pad0;
    unsigned long long best_key; // ordered bits of the best fast score among model-valid hypotheses
    unsigned long long pad1;
};

struct Scene // written by k_prep_features
{
    float xq_max, yq_max, iq_max, k_max; // maxima of |x/q|, |y/q|, |1/q| (scale class) and |k| (orientation class)
    unsigned int n_irr_s, n_irr_o;       // irregular features (kept out of the FP32 arrays)
    unsigned int irr_overflow, pad;
};

struct Thresholds
{
    double tsq[2];   // (2.25 t) t, S1
    float Tp;        // sqrt(tsq0) / (3 ln 2): scale threshold in log2 units of v
    float tanT;      // tan(sqrt(tsq1))
    float t_band0;   // constant part of the orientation band in tan units
    float pc[4];     // r^2 = z (pc0 + z (pc1 + z (pc2 + z pc3))), z = tan^2
    float scale_unit; // (3 ln 2)^2: converts sum of d^2 (log2 units) into sum of r^2
    int orient_all_exact; // orientation threshold outside the fast path's range -> every hypothesis exact
};

struct Buffers
{
    // features
    const double* scale;
    const double* orient;
    int ns, no;
    const float* fs; // chunked SoA, scale class: (x/q, y/q, 1/q)
    const float* fo; // chunked SoA, orientation class: (k, sin, cos)
    int ncs, nco;
    const Scene* scene;
    const unsigned int* irr_s;
    const unsigned int* irr_o;
    // hypotheses
    int n_hyp;
    unsigned long long first_hyp, seed;
    int* samples;       // [H,4]
    int have_samples;   // replay mode
    Model4* models;     // [H]
    unsigned char* flags; // [H]
    double* sums;       // [H,2]  sum of r^2 over inliers per class
    unsigned int* cnts; // [H,2]  raw inlier counts per class
    double* score;      // [H]
    unsigned int* n_inl; // [H,2] after zero-reset
    unsigned int* masks; // [H,W] or null
    int wmask, ws;       // words per hypothesis, words of the scale class
    // compact lists
    float4* pf0;         // (nh7, nh8, A, Tin)
    float4* pf1;         // (Tout, cphi, sphi, dprime)
    unsigned int* fast_hyp; // compact -> hypothesis
    unsigned int* hard_hyp;
    uint2* worklist;
    unsigned int worklist_cap;
    // candidates
    unsigned int* cand_hyp;  // [kCandCap]
    Model4* cand_model;
    double* cand_score;
    unsigned int* cand_cnt;  // [kCandCap,2]
    unsigned int* cand_masks; // [kCandCap, W]
    Counters* ctr;
    void* record;  // gcr_record (device), engine-owned
    void* record_user; // optional caller-owned copy target
    unsigned int* best_masks; // [W]
};

// ------------------------------------------------------------------------------------------------
// small PTX helpers (TMA bulk copy + mbarrier)

__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D TMA bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_barrier_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ float lg2_approx(float x)
{
    float r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float rcp_approx(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

__device__ __forceinline__ unsigned long long ordered_bits(double v)
{
    unsigned long long b = static_cast<unsigned long long>(__double_as_longlong(v));
    return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double from_ordered_bits(unsigned long long b)
{
    b = (b & 0x8000000000000000ull) ? (b & 0x7fffffffffffffffull) : ~b;
    return __longlong_as_double(static_cast<long long>(b));
}

// ------------------------------------------------------------------------------------------------
// K0: feature preparation.  Hoists what the reference recomputes per (model, feature):
//   scale       (x, y, s)     -> (x/q, y/q, 1/q), q = cbrt(s): then  s (1-h7x-h8y)^-3 = v^-3 with
//                                v = 1/q - h7 x/q - h8 y/q          (two_sift.hpp:621-645, model.h:194-204)
//   orientation (x, y, theta) -> (k = x sin - y cos, sin, cos): numer = sin - k h7, denom = cos + k h8
//                                                                  (model.h:156-165)
// Features whose prepared values are non-finite or absurdly large are "irregular": they are kept out of
// the FP32 arrays (sentinel = never an inlier) and always evaluated in FP64 (k_fixup).
__global__ void k_prep_features(const double* __restrict__ scale, int ns, const double* __restrict__ orient, int no,
                                float* __restrict__ fs, float* __restrict__ fo, int ncs, int nco, Scene* scene,
                                unsigned int* irr_s, unsigned int* irr_o)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const float inf = __int_as_float(0x7f800000);
    if (i < ncs * kChunk)
    {
        float a = 0.f, b = 0.f, c = inf; // pad: v = +inf -> never an inlier, never in the band
        if (i < ns)
        {
            const double x = scale[3 * i], y = scale[3 * i + 1], s = scale[3 * i + 2];
            const double q = cbrt(s);
            const double xa = x / q, ya = y / q, ia = 1.0 / q;
            const bool ok = isfinite(xa) && isfinite(ya) && isfinite(ia) && fabs(xa) < 1e12 && fabs(ya) < 1e12 &&
                            fabs(ia) < 1e12 && fabs(ia) > 1e-12;
            if (ok)
            {
                a = static_cast<float>(xa);
                b = static_cast<float>(ya);
                c = static_cast<float>(ia);
                atomicMax(reinterpret_cast<int*>(&scene->xq_max), __float_as_int(fabsf(a)));
                atomicMax(reinterpret_cast<int*>(&scene->yq_max), __float_as_int(fabsf(b)));
                atomicMax(reinterpret_cast<int*>(&scene->iq_max), __float_as_int(fabsf(c)));
            }
            else
            {
                const unsigned int slot = atomicAdd(&scene->n_irr_s, 1u);
                if (slot < kIrrCap) irr_s[slot] = i; else scene->irr_overflow = 1;
            }
        }
        float* base = fs + (i / kChunk) * kChunkFloats + (i % kChunk);
        base[0] = a;
        base[kChunk] = b;
        base[2 * kChunk] = c;
    }
    if (i < nco * kChunk)
    {
        float a = 0.f, b = 0.f, c = 0.f;
        bool pad = true;
        if (i < no)
        {
            const double x = orient[3 * i], y = orient[3 * i + 1], t = orient[3 * i + 2];
            double st, ct;
            sincos(t, &st, &ct);
            const double k = x * st - y * ct;
            const bool ok = isfinite(k) && isfinite(st) && isfinite(ct) && fabs(k) < 1e12;
            if (ok)
            {
                a = static_cast<float>(k);
                b = static_cast<float>(st);
                c = static_cast<float>(ct);
                pad = false;
                atomicMax(reinterpret_cast<int*>(&scene->k_max), __float_as_int(fabsf(a)));
            }
            else
            {
                const unsigned int slot = atomicAdd(&scene->n_irr_o, 1u);
                if (slot < kIrrCap) irr_o[slot] = i; else scene->irr_overflow = 1;
            }
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

__global__ void __launch_bounds__(128) k_sample_solve(Buffers B, Thresholds T, int all_exact)
{
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned lane = threadIdx.x & 31;
    bool solved = false, hard = false, sample_valid = false;
    Model4 m = {0.0, 0.0, 0.0, 0.0};
    float4 p0 = make_float4(0.f, 0.f, 0.f, 0.f), p1 = make_float4(0.f, 1.f, 0.f, 0.f);
    if (h < B.n_hyp)
    {
        int s0, s1, o0, o1;
        if (B.have_samples)
        {
            const int4 s = reinterpret_cast<const int4*>(B.samples)[h];
            s0 = s.x; s1 = s.y; o0 = s.z; o1 = s.w;
        }
        else
        {
            const unsigned long long g = B.first_hyp + static_cast<unsigned long long>(h);
            uint32_t c[4] = {static_cast<uint32_t>(g), static_cast<uint32_t>(g >> 32), 0x47435231u, 0u};
            philox4x32_10(c, static_cast<uint32_t>(B.seed), static_cast<uint32_t>(B.seed >> 32));
            draw_pair(c[0], c[1], static_cast<uint32_t>(B.ns), s0, s1);
            draw_pair(c[2], c[3], static_cast<uint32_t>(B.no), o0, o1);
            if (B.samples) reinterpret_cast<int4*>(B.samples)[h] = make_int4(s0, s1, o0, o1);
        }
        const bool in_range = static_cast<unsigned>(s0) < static_cast<unsigned>(B.ns) &&
                              static_cast<unsigned>(s1) < static_cast<unsigned>(B.ns) &&
                              static_cast<unsigned>(o0) < static_cast<unsigned>(B.no) &&
                              static_cast<unsigned>(o1) < static_cast<unsigned>(B.no);
        if (!in_range)
        {
            atomicAdd(&B.ctr->n_bad_sample, 1u);
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
            sample_valid = is_valid_sample(fs0, fs1, fo0, fo1, vp); // GCRANSAC.h:315
            if (sample_valid) solved = estimate_minimal(fs0, fs1, vp, m); // GCRANSAC.h:326
        }
        unsigned char f = 0;
        if (sample_valid) f |= 1;
        if (solved)
        {
            f |= 2;
            if (is_valid_model(m)) f |= 4; // consulted at selection time, GCRANSAC.h:440-446
            // ---- FP32 parameters and their error bands (DESIGN.md "error budget")
            const Scene sc = *B.scene;
            const double ah7 = fabs(m.h7), ah8 = fabs(m.h8);
            const double dv = 8.0 * kU * (ah7 * sc.xq_max + ah8 * sc.yq_max + sc.iq_max);
            const double A = log2(m.alpha);
            const double Tp = T.Tp;
            const double v_lo = m.alpha * exp2(-Tp), v_hi = m.alpha * exp2(Tp);
            double drho = 2.0 * (dv / ((v_lo - dv) * kLn2) + 2.384185791015625e-7 + 3.0 * kU * (fabs(A) + Tp + 1.0));
            const double kh = sc.k_max * (ah7 + ah8);
            const double dprime = 6.0 * kU * (2.0 + 3.0 * kh) * 1.0001;
            hard = all_exact || T.orient_all_exact || sc.irr_overflow || !isfinite(dv) || !isfinite(A) || !(v_lo > 4.0 * dv) ||
                   !(v_hi * 1.01 < 1000.0) || !(drho < 1e-3 * Tp) || !(kh < 1e6) || !isfinite(dprime) ||
                   !(ah7 < 1e6) || !(ah8 < 1e6);
            if (hard) f |= 8;
            double sphi, cphi;
            sincos(m.phi, &sphi, &cphi);
            p0 = make_float4(static_cast<float>(-m.h7), static_cast<float>(-m.h8), static_cast<float>(A),
                             __double2float_rd(Tp - drho));
            p1 = make_float4(__double2float_ru(Tp + drho), static_cast<float>(cphi), static_cast<float>(sphi),
                             __double2float_ru(dprime));
        }
        B.flags[h] = f;
        B.models[h] = m;
        B.sums[2 * h] = 0.0;
        B.sums[2 * h + 1] = 0.0;
        B.cnts[2 * h] = 0u;
        B.cnts[2 * h + 1] = 0u;
    }
    // warp-aggregated compaction of solved hypotheses into the fast / exact lists
    const unsigned m_valid = __ballot_sync(0xffffffffu, sample_valid);
    const unsigned m_fast = __ballot_sync(0xffffffffu, solved && !hard);
    const unsigned m_hard = __ballot_sync(0xffffffffu, solved && hard);
    unsigned base_fast = 0, base_hard = 0;
    if (lane == 0)
    {
        if (m_valid) atomicAdd(&B.ctr->n_sample_valid, __popc(m_valid));
        if (m_fast | m_hard) atomicAdd(&B.ctr->n_solved, __popc(m_fast) + __popc(m_hard));
        if (m_fast) base_fast = atomicAdd(&B.ctr->n_fast, __popc(m_fast));
        if (m_hard) base_hard = atomicAdd(&B.ctr->n_hard, __popc(m_hard));
    }
    base_fast = __shfl_sync(0xffffffffu, base_fast, 0);
    base_hard = __shfl_sync(0xffffffffu, base_hard, 0);
    const unsigned below = (1u << lane) - 1u;
    if (solved && !hard)
    {
        const unsigned c = base_fast + __popc(m_fast & below);
        B.pf0[c] = p0;
        B.pf1[c] = p1;
        B.fast_hyp[c] = h;
    }
    if (solved && hard) B.hard_hyp[base_hard + __popc(m_hard & below)] = h;
}

// ------------------------------------------------------------------------------------------------
// K3: the FP32 MSAC kernel.

struct ScaleEval
{
    bool in, unc;
    float d;
};
// One (hypothesis, scale feature) evaluation.  rectified scale = v^-3, residual r = 3 ln2 |A - lg2 v|
// (S4, two_sift.hpp:621-645).  `in`: certainly inside the truncated threshold (S1); `unc`: inside the
// proven FP32 error band around it -> decided in FP64 by k_fixup.  v <= 0 (beyond the horizon, DBL_MAX
// residual in the reference) gives NaN/inf and fails both tests.
__device__ __forceinline__ ScaleEval eval_scale(float xq, float yq, float iq, float nh7, float nh8, float A, float Tin,
                                                float Tout)
{
    const float v = fmaf(nh8, yq, fmaf(nh7, xq, iq));
    const float d = A - lg2_approx(v);
    ScaleEval e;
    e.in = fabsf(d) <= Tin;
    e.unc = (fabsf(d) <= Tout) && !e.in;
    e.d = d;
    return e;
}

struct OrientEval
{
    bool in, unc;
    float z, p; // r^2 = z * p
};
// One (hypothesis, orientation feature) evaluation.  (denom, numer) of rectifiedAngle (model.h:156-165)
// rotated by -phi; the residual (S5, two_sift.hpp:647-665) is the angle of that vector to the nearest
// multiple of pi/2 = atan(min/max) of its absolute components; r^2 by an even polynomial in tan.
__device__ __forceinline__ OrientEval eval_orient(float k, float st, float ct, float nh7, float nh8, float cphi,
                                                  float sphi, float dprime, const Thresholds& T)
{
    const float n = fmaf(k, nh7, st);  // sin - k h7
    const float d = fmaf(-k, nh8, ct); // cos + k h8
    const float c1 = fmaf(n, sphi, d * cphi);
    const float s1 = fmaf(n, cphi, -(d * sphi));
    const float a1 = fabsf(c1), a2 = fabsf(s1);
    // the 1e-30 floor keeps an exact FP32 0/0 (numer = denom = 0, where the reference evaluates atan2 of
    // round-off) finite: t = 0 with an enormous band -> guard band -> FP64
    const float mn = fminf(a1, a2), mx = fmaxf(fmaxf(a1, a2), 1e-30f);
    const float ri = rcp_approx(mx);
    const float t = mn * ri;
    const float band = fmaf(dprime, ri, T.t_band0);
    OrientEval e;
    e.in = t <= T.tanT - band;
    e.unc = (t <= T.tanT + band) && !e.in;
    e.z = t * t;
    e.p = fmaf(e.z, fmaf(e.z, fmaf(e.z, T.pc[3], T.pc[2]), T.pc[1]), T.pc[0]);
    return e;
}

struct __align__(16) ScoreSmem
{
    float stage[kStages][kChunkFloats];
    uint64_t full[kStages];
    uint64_t empty[kStages];
};

template <bool MASKS>
__global__ void __launch_bounds__(kScoreThreads) k_score_fast(const __grid_constant__ Buffers B,
                                                              const __grid_constant__ Thresholds T)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    ScoreSmem& S = *reinterpret_cast<ScoreSmem*>(smem_raw);
    const int tid = threadIdx.x;
    const int lane = tid & 31;
    if (tid == 0)
    {
        for (int s = 0; s < kStages; s++)
        {
            mbar_init(&S.full[s], 1);
            mbar_init(&S.empty[s], kScoreThreads / 32);
        }
        fence_barrier_init();
    }
    __syncthreads();

    const int n_fast = static_cast<int>(B.ctr->n_fast);
    const int n_tiles = (n_fast + kTileHyps - 1) / kTileHyps;
    const int n_chunks = B.ncs + B.nco;
    // work decomposition (identical in every CTA): split the feature chunks so that there are enough items
    int splits = 1;
    {
        const int target = 4 * static_cast<int>(gridDim.x);
        if (n_tiles > 0 && n_tiles < target) splits = min(n_chunks, (target + n_tiles - 1) / n_tiles);
    }
    const int cps = (n_chunks + splits - 1) / splits; // chunks per split
    splits = (n_chunks + cps - 1) / cps;
    const int n_items = n_tiles * splits;

    uint32_t ring = 0; // chunks issued/consumed so far by this CTA (ring position)
    for (int item = blockIdx.x; item < n_items; item += gridDim.x)
    {
        const int tile = item / splits, split = item - tile * splits;
        const int c_begin = split * cps, c_end = min(n_chunks, c_begin + cps);
        const int nch = c_end - c_begin;

        auto issue = [&](uint32_t n, int chunk) {
            if (tid == 0)
            {
                const uint32_t st = n % kStages;
                if (n >= kStages) mbar_wait(&S.empty[st], ((n / kStages) - 1) & 1);
                const float* src = chunk < B.ncs ? B.fs + static_cast<size_t>(chunk) * kChunkFloats
                                                 : B.fo + static_cast<size_t>(chunk - B.ncs) * kChunkFloats;
                mbar_expect_tx(&S.full[st], kChunkBytes);
                tma_load_1d(S.stage[st], src, kChunkBytes, &S.full[st]);
            }
        };
        for (int p = 0; p < min(kStages - 1, nch); p++) issue(ring + p, c_begin + p);

        // per-thread hypotheses
        float nh7[kHB], nh8[kHB], A[kHB], Tin[kHB], Tout[kHB], cphi[kHB], sphi[kHB], dpr[kHB];
        unsigned int hyp[kHB];
        bool active[kHB];
        double sum_s[kHB], sum_o[kHB];
        unsigned int cnt_s[kHB], cnt_o[kHB];
#pragma unroll
        for (int j = 0; j < kHB; j++)
        {
            const int c = tile * kTileHyps + j * kScoreThreads + tid;
            active[j] = c < n_fast;
            const float4 p0 = active[j] ? B.pf0[c] : make_float4(0.f, 0.f, 0.f, -1.f);
            const float4 p1 = active[j] ? B.pf1[c] : make_float4(-1.f, __int_as_float(0x7fc00000), 0.f, 0.f);
            hyp[j] = active[j] ? B.fast_hyp[c] : 0u;
            nh7[j] = p0.x; nh8[j] = p0.y; A[j] = p0.z; Tin[j] = p0.w;
            Tout[j] = p1.x; cphi[j] = p1.y; sphi[j] = p1.z; dpr[j] = p1.w;
            sum_s[j] = sum_o[j] = 0.0;
            cnt_s[j] = cnt_o[j] = 0u;
        }

        for (int i = 0; i < nch; i++)
        {
            if (i + kStages - 1 < nch) issue(ring + i + kStages - 1, c_begin + i + kStages - 1);
            const uint32_t n = ring + i;
            const uint32_t st = n % kStages;
            mbar_wait(&S.full[st], (n / kStages) & 1);
            const int chunk = c_begin + i;
            const float4* pa = reinterpret_cast<const float4*>(S.stage[st]);
            const float4* pb = reinterpret_cast<const float4*>(S.stage[st] + kChunk);
            const float4* pc = reinterpret_cast<const float4*>(S.stage[st] + 2 * kChunk);
            float acc[kHB], cntf[kHB]; // per-chunk FP32 partials (counts are exact integers < 2^24)
#pragma unroll
            for (int j = 0; j < kHB; j++) acc[j] = cntf[j] = 0.f;

            if (chunk < B.ncs)
            {
                const int fbase = chunk * kChunk;
                unsigned int mword[kHB];
#pragma unroll
                for (int j = 0; j < kHB; j++) mword[j] = 0u;
#pragma unroll 2
                for (int g = 0; g < kChunk / 4; g++)
                {
                    const float4 a = pa[g], b = pb[g], c = pc[g];
                    const float fa[4] = {a.x, a.y, a.z, a.w}, fb[4] = {b.x, b.y, b.z, b.w}, fc[4] = {c.x, c.y, c.z, c.w};
                    bool any_unc = false;
#pragma unroll
                    for (int j = 0; j < kHB; j++)
                    {
#pragma unroll
                        for (int e = 0; e < 4; e++)
                        {
                            const ScaleEval r = eval_scale(fa[e], fb[e], fc[e], nh7[j], nh8[j], A[j], Tin[j], Tout[j]);
                            if (r.in)
                            {
                                acc[j] = fmaf(r.d, r.d, acc[j]);
                                cntf[j] += 1.0f;
                                if (MASKS) mword[j] |= 1u << ((g * 4 + e) & 31);
                            }
                            any_unc |= r.unc;
                        }
                    }
                    if (__any_sync(0xffffffffu, any_unc))
                    {
                        // rare: append the guard-band pairs of this group to the FP64 worklist
#pragma unroll
                        for (int j = 0; j < kHB; j++)
#pragma unroll
                            for (int e = 0; e < 4; e++)
                            {
                                const ScaleEval r =
                                    eval_scale(fa[e], fb[e], fc[e], nh7[j], nh8[j], A[j], Tin[j], Tout[j]);
                                if (r.unc && active[j])
                                {
                                    const unsigned int slot = atomicAdd(&B.ctr->n_worklist, 1u);
                                    if (slot < B.worklist_cap)
                                        B.worklist[slot] = make_uint2(hyp[j], static_cast<unsigned>(fbase + g * 4 + e));
                                    else
                                        atomicAdd(&B.ctr->n_worklist_overflow, 1u);
                                }
                            }
                    }
                    if (MASKS && (g & 7) == 7)
                    {
#pragma unroll
                        for (int j = 0; j < kHB; j++)
                        {
                            if (active[j] && mword[j])
                                atomicOr(&B.masks[static_cast<size_t>(hyp[j]) * B.wmask + (fbase >> 5) + (g >> 3)], mword[j]);
                            mword[j] = 0u;
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < kHB; j++)
                {
                    sum_s[j] += static_cast<double>(acc[j]);
                    cnt_s[j] += static_cast<unsigned int>(cntf[j]);
                }
            }
            else
            {
                const int fbase = (chunk - B.ncs) * kChunk;
                unsigned int mword[kHB];
#pragma unroll
                for (int j = 0; j < kHB; j++) mword[j] = 0u;
#pragma unroll 2
                for (int g = 0; g < kChunk / 4; g++)
                {
                    const float4 a = pa[g], b = pb[g], c = pc[g];
                    const float fa[4] = {a.x, a.y, a.z, a.w}, fb[4] = {b.x, b.y, b.z, b.w}, fc[4] = {c.x, c.y, c.z, c.w};
                    bool any_unc = false;
#pragma unroll
                    for (int j = 0; j < kHB; j++)
                    {
#pragma unroll
                        for (int e = 0; e < 4; e++)
                        {
                            const OrientEval r =
                                eval_orient(fa[e], fb[e], fc[e], nh7[j], nh8[j], cphi[j], sphi[j], dpr[j], T);
                            if (r.in)
                            {
                                acc[j] = fmaf(r.z, r.p, acc[j]);
                                cntf[j] += 1.0f;
                                if (MASKS) mword[j] |= 1u << ((g * 4 + e) & 31);
                            }
                            any_unc |= r.unc;
                        }
                    }
                    if (__any_sync(0xffffffffu, any_unc))
                    {
#pragma unroll
                        for (int j = 0; j < kHB; j++)
#pragma unroll
                            for (int e = 0; e < 4; e++)
                            {
                                const OrientEval r =
                                    eval_orient(fa[e], fb[e], fc[e], nh7[j], nh8[j], cphi[j], sphi[j], dpr[j], T);
                                const int f = fbase + g * 4 + e;
                                if (r.unc && active[j] && f < B.no)
                                {
                                    const unsigned int slot = atomicAdd(&B.ctr->n_worklist, 1u);
                                    if (slot < B.worklist_cap)
                                        B.worklist[slot] = make_uint2(hyp[j], 0x80000000u | static_cast<unsigned>(f));
                                    else
                                        atomicAdd(&B.ctr->n_worklist_overflow, 1u);
                                }
                            }
                    }
                    if (MASKS && (g & 7) == 7)
                    {
#pragma unroll
                        for (int j = 0; j < kHB; j++)
                        {
                            if (active[j] && mword[j])
                                atomicOr(&B.masks[static_cast<size_t>(hyp[j]) * B.wmask + B.ws + (fbase >> 5) + (g >> 3)],
                                         mword[j]);
                            mword[j] = 0u;
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < kHB; j++)
                {
                    sum_o[j] += static_cast<double>(acc[j]);
                    cnt_o[j] += static_cast<unsigned int>(cntf[j]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&S.empty[st]);
        }
        ring += nch;

        // flush this item's partial sums
#pragma unroll
        for (int j = 0; j < kHB; j++)
        {
            if (!active[j]) continue;
            const size_t h = hyp[j];
            const double ss = sum_s[j] * static_cast<double>(T.scale_unit);
            if (splits == 1)
            {
                B.sums[2 * h] = ss;
                B.sums[2 * h + 1] = sum_o[j];
                B.cnts[2 * h] = cnt_s[j];
                B.cnts[2 * h + 1] = cnt_o[j];
            }
            else
            {
                if (cnt_s[j])
                {
                    atomicAdd(&B.sums[2 * h], ss);
                    atomicAdd(&B.cnts[2 * h], cnt_s[j]);
                }
                if (cnt_o[j])
                {
                    atomicAdd(&B.sums[2 * h + 1], sum_o[j]);
                    atomicAdd(&B.cnts[2 * h + 1], cnt_o[j]);
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K3x: FP64 re-evaluation of guard-band pairs and of irregular features, patched in with atomics.
__global__ void k_fixup(Buffers B, Thresholds T)
{
    const unsigned int n_wl = min(B.ctr->n_worklist, B.worklist_cap);
    const unsigned int stride = gridDim.x * blockDim.x;
    const unsigned int t0 = blockIdx.x * blockDim.x + threadIdx.x;
    for (unsigned int i = t0; i < n_wl; i += stride)
    {
        const uint2 w = B.worklist[i];
        const unsigned int h = w.x, cls = w.y >> 31, f = w.y & 0x7fffffffu;
        const Model4 m = B.models[h];
        double r;
        if (cls == 0) r = scale_residual(B.scale[3 * f], B.scale[3 * f + 1], B.scale[3 * f + 2], m);
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
    const unsigned long long total = static_cast<unsigned long long>(n_irr) * B.ctr->n_fast;
    for (unsigned long long i = t0; i < total; i += stride)
    {
        const unsigned int c = static_cast<unsigned int>(i / n_irr), k = static_cast<unsigned int>(i % n_irr);
        const unsigned int h = B.fast_hyp[c];
        const unsigned int cls = k >= n_is;
        const unsigned int f = cls ? B.irr_o[k - n_is] : B.irr_s[k];
        const Model4 m = B.models[h];
        double r;
        if (cls == 0) r = scale_residual(B.scale[3 * f], B.scale[3 * f + 1], B.scale[3 * f + 2], m);
        else r = orientation_residual(B.orient[3 * f], B.orient[3 * f + 1], B.orient[3 * f + 2], m);
        const double r2 = r * r;
        if (r2 <= T.tsq[cls])
        {
            atomicAdd(&B.sums[2 * h + cls], r2);
            atomicAdd(&B.cnts[2 * h + cls], 1u);
            if (B.masks) atomicOr(&B.masks[static_cast<size_t>(h) * B.wmask + (cls ? B.ws : 0) + (f >> 5)], 1u << (f & 31));
        }
    }
}

// ------------------------------------------------------------------------------------------------
// FP64 reference-order MSAC of whole hypotheses: MSACScoringFunction::getScore
// (MSAC_scoring_function.hpp:53-130) with one CTA per hypothesis; a warp covers 32 consecutive features so
// that __ballot_sync yields the packed mask word directly.
//   mode 0: the hard list (results into the per-hypothesis arrays)
//   mode 1: the candidate list (results into the candidate slots, value finalised)
//   mode 2: one caller-supplied model in cand slot 0
constexpr int kExactThreads = 256;

__device__ inline void exact_score_block(const Buffers& B, const Thresholds& T, const Model4& m, double& sum0,
                                         double& sum1, unsigned int& c0, unsigned int& c1, unsigned int* mask_out)
{
    __shared__ double red[2][kExactThreads / 32];
    __shared__ unsigned int redc[2][kExactThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double s[2] = {0.0, 0.0};
    unsigned int c[2] = {0u, 0u};
#pragma unroll
    for (int cls = 0; cls < 2; cls++)
    {
        const int n = cls ? B.no : B.ns;
        const double* F = cls ? B.orient : B.scale;
        const int nw = (n + 31) >> 5;
        for (int w = warp; w < nw; w += kExactThreads / 32)
        {
            const int f = w * 32 + lane;
            bool in = false;
            double r2 = 0.0;
            if (f < n)
            {
                const double r = cls ? orientation_residual(F[3 * f], F[3 * f + 1], F[3 * f + 2], m)
                                     : scale_residual(F[3 * f], F[3 * f + 1], F[3 * f + 2], m);
                r2 = r * r;
                in = r2 <= T.tsq[cls];
            }
            const unsigned int word = __ballot_sync(0xffffffffu, in);
            if (in)
            {
                s[cls] += r2;
                c[cls]++;
            }
            if (mask_out && lane == 0) mask_out[(cls ? B.ws : 0) + w] = word;
        }
    }
#pragma unroll
    for (int cls = 0; cls < 2; cls++)
    {
        double v = s[cls];
        unsigned int k = c[cls];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
        {
            v += __shfl_xor_sync(0xffffffffu, v, o);
            k += __shfl_xor_sync(0xffffffffu, k, o);
        }
        if (lane == 0)
        {
            red[cls][warp] = v;
            redc[cls][warp] = k;
        }
    }
    __syncthreads();
    double t0 = 0.0, t1 = 0.0;
    unsigned int k0 = 0, k1 = 0;
    for (int w = 0; w < kExactThreads / 32; w++)
    {
        t0 += red[0][w];
        t1 += red[1][w];
        k0 += redc[0][w];
        k1 += redc[1][w];
    }
    __syncthreads();
    sum0 = t0; sum1 = t1; c0 = k0; c1 = k1;
}

// S2/S3: per-class value = -sum r^2 / tsq + n; any class below the minimal sample size (2) zeroes everything
__device__ __forceinline__ double msac_value(const Thresholds& T, double s0, double s1, unsigned int c0, unsigned int c1,
                                             unsigned int& n0, unsigned int& n1)
{
    if (c0 < 2u || c1 < 2u)
    {
        n0 = n1 = 0u;
        return 0.0;
    }
    n0 = c0;
    n1 = c1;
    return (-s0 / T.tsq[0] + static_cast<double>(c0)) + (-s1 / T.tsq[1] + static_cast<double>(c1));
}

__global__ void __launch_bounds__(kExactThreads) k_exact_items(Buffers B, Thresholds T, int mode)
{
    const unsigned int n_items = mode == 0 ? B.ctr->n_hard : (mode == 1 ? min(B.ctr->n_candidates, static_cast<unsigned>(kCandCap)) : 1u);
    for (unsigned int it = blockIdx.x; it < n_items; it += gridDim.x)
    {
        double s0, s1;
        unsigned int c0, c1;
        if (mode == 0)
        {
            const unsigned int h = B.hard_hyp[it];
            const Model4 m = B.models[h];
            exact_score_block(B, T, m, s0, s1, c0, c1, B.masks ? B.masks + static_cast<size_t>(h) * B.wmask : nullptr);
            if (threadIdx.x == 0)
            {
                B.sums[2 * h] = s0;
                B.sums[2 * h + 1] = s1;
                B.cnts[2 * h] = c0;
                B.cnts[2 * h + 1] = c1;
            }
        }
        else
        {
            const Model4 m = B.cand_model[it];
            exact_score_block(B, T, m, s0, s1, c0, c1, B.cand_masks + static_cast<size_t>(it) * B.wmask);
            if (threadIdx.x == 0)
            {
                unsigned int n0, n1;
                B.cand_score[it] = msac_value(T, s0, s1, c0, c1, n0, n1);
                B.cand_cnt[2 * it] = n0;
                B.cand_cnt[2 * it + 1] = n1;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K4: per-hypothesis MSAC value + arg-max key over model-valid hypotheses.
__global__ void k_finalize(Buffers B, Thresholds T)
{
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    double v = 0.0;
    bool contender = false;
    if (h < B.n_hyp)
    {
        const unsigned char f = B.flags[h];
        unsigned int n0 = 0, n1 = 0;
        if (f & 2) v = msac_value(T, B.sums[2 * h], B.sums[2 * h + 1], B.cnts[2 * h], B.cnts[2 * h + 1], n0, n1);
        B.score[h] = v;
        B.n_inl[2 * h] = n0;
        B.n_inl[2 * h + 1] = n1;
        contender = (f & 4) && v > 0.0; // S10 + S11 (initial best is 0, strict '<')
    }
    unsigned long long key = contender ? ordered_bits(v) : 0ull;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other > key ? other : key;
    }
    if ((threadIdx.x & 31) == 0 && key) atomicMax(&B.ctr->best_key, key);
}

// K5: hypotheses whose FP32-path score is within the path's error of the best one are re-scored in FP64
// so that the selection rule sees reference-order scores.
__global__ void k_collect(Buffers B, double rel_tol)
{
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long bk = B.ctr->best_key;
    if (h >= B.n_hyp || bk == 0ull) return;
    const double best = from_ordered_bits(bk);
    const unsigned char f = B.flags[h];
    const double v = B.score[h];
    if ((f & 4) && v > 0.0 && v >= best - rel_tol * best)
    {
        const unsigned int slot = atomicAdd(&B.ctr->n_candidates, 1u);
        if (slot < kCandCap)
        {
            B.cand_hyp[slot] = h;
            B.cand_model[slot] = B.models[h];
        }
        else
            atomicAdd(&B.ctr->n_cand_overflow, 1u);
    }
}

struct Record // mirrors gcr_record
{
    double score;
    long long index;
    double model[4];
    unsigned int n_inl[2];
    unsigned long long reserved;
};

// K6: final selection (GCRANSAC.h:440-462): highest exact score, earliest hypothesis on ties; writes the
// winner record (also into the caller's buffer, e.g. an NCCL send buffer) and the winner's packed masks.
__global__ void k_select(Buffers B)
{
    __shared__ double s_best;
    __shared__ unsigned int s_hyp, s_slot;
    const unsigned int n = min(B.ctr->n_candidates, static_cast<unsigned>(kCandCap));
    if (threadIdx.x == 0)
    {
        double best = 0.0;
        unsigned int bh = 0xffffffffu, bs = 0;
        for (unsigned int i = 0; i < n; i++)
        {
            const double v = B.cand_score[i];
            const unsigned int h = B.cand_hyp[i];
            if (v > best || (v == best && v > 0.0 && h < bh))
            {
                best = v;
                bh = h;
                bs = i;
            }
        }
        s_best = best;
        s_hyp = bh;
        s_slot = bs;
        Record r;
        r.score = 0.0;
        r.index = -1;
        r.model[0] = r.model[1] = r.model[2] = r.model[3] = 0.0;
        r.n_inl[0] = r.n_inl[1] = 0;
        r.reserved = 0;
        if (bh != 0xffffffffu)
        {
            const Model4 m = B.cand_model[bs];
            r.score = best;
            r.index = static_cast<long long>(B.first_hyp + bh);
            r.model[0] = m.h7; r.model[1] = m.h8; r.model[2] = m.alpha; r.model[3] = m.phi;
            r.n_inl[0] = B.cand_cnt[2 * bs];
            r.n_inl[1] = B.cand_cnt[2 * bs + 1];
        }
        *reinterpret_cast<Record*>(B.record) = r;
        if (B.record_user) *reinterpret_cast<Record*>(B.record_user) = r;
    }
    __syncthreads();
    const bool have = s_hyp != 0xffffffffu;
    for (int w = threadIdx.x; w < B.wmask; w += blockDim.x)
        B.best_masks[w] = have ? B.cand_masks[static_cast<size_t>(s_slot) * B.wmask + w] : 0u;
}

// caller-supplied model -> candidate slot 0 (gcr_score_model)
__global__ void k_set_model(Buffers B, Model4 m)
{
    B.cand_model[0] = m;
    B.cand_hyp[0] = 0;
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
