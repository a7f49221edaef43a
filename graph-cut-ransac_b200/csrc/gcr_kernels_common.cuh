// Shared declarations of the device code: constants, buffer descriptors, PTX helpers (TMA bulk copy, mbarrier,
// MUFU approximations).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include "gcr_exact.cuh"

namespace gcr
{

constexpr int kChunk = 256;             // features per shared-memory stage
constexpr int kChunkFloats = 3 * kChunk; // three SoA planes per chunk
constexpr int kChunkBytes = kChunkFloats * 4;
constexpr int kStages = 6;
#ifndef GCR_THREADS
#define GCR_THREADS 128
#endif
// Two hypotheses per thread, six 128-thread CTAs per SM (80 registers): every LDS.128 of the feature planes serves two hypotheses, which
// relieves the MIO queue the MUFUs share.  Same-box A/B of the winner-only ring kernel at C3 (tools/compare_variants.py): 4.979 (1 hypothesis,
// 8 CTAs) / 4.919 (2, 6 CTAs) / 5.003 (2, 5 CTAs, 96 registers) / 5.046 ms (2, 4 CTAs, 100 registers); C5 0.2018 -> 0.1935 ms.  (Round 1
// measured the opposite for the reporting kernel, whose per-hypothesis state is twice as large.)
#ifndef GCR_HB
#define GCR_HB 2
#endif
#ifndef GCR_MINB
#define GCR_MINB 6
#endif
// Packed f32x2 arithmetic (FFMA2 / FMUL2: two IEEE operations per issue slot) in the FP32 tiles of the scoring kernel:
// 0 none, 1 scale class, 2 + the six leading FMAs of an orientation evaluation, 3 + the rest of the orientation evaluation.
// Interleaved A/B on one B200 (tools/compare_variants.py, C3 kernel): 9.94 / 9.60 / 9.80 / 9.75 ms for levels 0 / 1 / 2 / 3:
// the pair forms save issue slots but hold the FMA pipe for two cycles, so only the scale class (FMA-light) gains.
#ifndef GCR_F2
#define GCR_F2 1
#endif
constexpr int kResultHeaderWords = 48; // result block: record (16 words) + counters (16) + phase stamps (16), then the winner's masks
#ifndef GCR_ITEMS
#define GCR_ITEMS 24
#endif
constexpr int kItemsPerCta = GCR_ITEMS; // work items per resident CTA the chunk lists are cut for (k_score_fast)
#ifndef GCR_VOTE
#define GCR_VOTE 4
#endif
#ifndef GCR_UNROLL
#define GCR_UNROLL (GCR_VOTE < 4 ? GCR_VOTE : 4)
#endif
constexpr int kScoreUnroll = GCR_UNROLL;
constexpr int kVoteGroups = GCR_VOTE; // groups of 4 (FP32) / 2 (precise) features between two warp votes of the band check
template <int CLS>
constexpr bool kPacked = CLS == 0 ? (GCR_F2 >= 1) : (GCR_F2 >= 2);
constexpr int kScoreThreads = GCR_THREADS;
constexpr int kCandCap = 1024;
constexpr int kIrrCap = 4096;
constexpr int kMaxPeers = 16;
constexpr int kXchgSlotWords = 16;  // u64 words per slot: 8 record words, 1 sequence word, padding to 128 B
constexpr int kXchgWords = 2 * kMaxPeers * kXchgSlotWords; // two parities
constexpr double kLn2 = 0.693147180559945309417;
constexpr double kU = 5.9604644775390625e-8; // 2^-24, FP32 unit round-off

struct Counters
{
    unsigned int n_sample_valid, n_solved, n_fast, n_hard;
    unsigned int n_worklist, n_worklist_overflow, n_candidates, n_cand_overflow;
    unsigned int n_bad_sample, next_item, xchg_timeout, n_sens;
    unsigned long long best_key; // ordered bits of the best fast score among model-valid hypotheses (winner-only runs: of the best LOWER bound)
    unsigned int n_refined;      // winner-only runs: contenders re-scored in mid precision with exact inlier decisions (k_tail)
    unsigned int pad1;
};

constexpr int kContCap = 4096; // winner-only runs: contenders refined in mid precision per run (k_tail)
struct ContPar // one contender: parameters of its refinement (gcr_kernels.cuh, refine_phase)
{
    double h7, h8, cphi, sphi;
    float A, Tin, Tout, dpr;       // log2 alpha; |A - lg2 v| certainly in / possibly in; FP64 part of the orientation band (x 1/mx)
    float e_s, e_o, ri_cap, band0r; // error per certain inlier's term (scale, orientation); 1/mx beyond which FP64 decides; FP32 part of the band
    unsigned int hyp, exact;        // hypothesis; 1 = scored by the FP64 kernel already
    unsigned int pad_[2];
};
struct TailState // winner-only runs: state of the contender rounds (cleared with the run counters by the last kernel)
{
    unsigned long long lb_ref_key; // ordered bits of the best refined lower bound
    unsigned int n_cont, cont_overflow;
    unsigned int unproven, pad_;   // a contender whose class values do not prove the minimal inlier counts by themselves
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
    double exp2_neg_Tp, exp2_pos_Tp; // 2^-Tp, 2^+Tp: bounds of v / alpha for a scale inlier
    float Tp;        // sqrt(tsq0) / (3 ln 2): scale threshold in log2 units of v
    float tanT;      // tan(sqrt(tsq1))
    float t_band0;   // constant part of the orientation band in tan units
    float t_lo0, t_hi0; // tanT -/+ t_band0 (rounded outwards)
    int poly_deg;    // 2 or 3: degree of the r^2 polynomial that meets the fit tolerance
    float pc[4];     // r^2 = z (pc0 + z (pc1 + z (pc2 + z pc3))), z = tan^2
    float pm[3];     // monic form: P(z) = pc_lead * Q(z), Q = ((z + pm0) z + pm1) z + pm2 (degree 3) or (z + pm0) z + pm1 (degree 2):
    float pc_lead;   // every step has ONE constant operand (no register copies of constants in the hot loop); lead factor applied per chunk
    float scale_unit; // (3 ln 2)^2: converts sum of d^2 (log2 units) into sum of r^2
    int orient_all_exact; // orientation threshold outside the fast path's range -> every hypothesis exact
    int problem;          // 0 hybrid (two_sift), 1 scale-only (three_sift), 2 scale-only "original" (three_sift_original)
    int strict;           // gcr_score_models only: inlier iff r^2 < tsq
    unsigned int min_inl[2]; // minimal sample size per class: fewer inliers than this resets the score to zero (S3)
    // winner-only runs (gcr_score.cuh, score_chunk_w): MSAC as the truncated quadratic sum_f max(0, 1 - r_f^2 / tsq)
    float inv_Tp;    // 1 / Tp: dq = (A - lg2 v) / Tp, scale inlier <=> |dq| <= 1
    float pn[4];     // orientation: r^2 / tsq = z (pn0 + z (pn1 + z (pn2 + z pn3))), z = tan^2 (= pc / tsq1)
    float sqrt_tsq1; // sqrt(tsq[1]) (error bound of the orientation class value)
};

struct Buffers
{
    // features
    const double* scale;
    const double* orient;
    int per_hyp;         // reporting mode: per-hypothesis scores / inlier counts are outputs (GCR_RUN_KEEP_PER_HYP / EMIT_MASKS);
                         // 0 = winner-only run: truncated-quadratic class values, no counts / bands / worklist, invalid models unscored
    float2* err_ab;      // winner-only runs, [H]: error bound of the FP32 class values = err_ab.x + err_ab.y * (sum of 1/mx over the orientation evaluations, cnts[2h+1] as float) + 8e-6 value
    const double* aux_s; // [ns] pow(s, 1/3) (scale-only original: pow(s, -1/3)): the solvers' per-feature column
    const double* aux_o; // [no,2] (sin, cos) of the orientation
    int ns, no;
    const float* fs; // chunked SoA, scale class: (x/q, y/q, 1/q)
    const float* fo; // chunked SoA, orientation class: (k, sin, cos)
    int ncs, nco;
    const double* fsd; // FP64 chunked SoA (256 features per chunk) for the precise tiles: (x/q, y/q, 1/q)
    const double* fod; // FP64 chunked SoA: (k, sin, cos)
    int ncsd, ncod;
    const Scene* scene;
    const unsigned int* irr_s;
    const unsigned int* irr_o;
    // hypotheses
    int n_hyp;
    unsigned long long first_hyp, seed;
    int* samples;       // [H,4]
    int have_samples;   // replay mode
    int samples_u16;    // replay mode: the array holds uint16 indices (GCR_RUN_SAMPLES_U16)
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
    unsigned int* cand_cnt;  // [kCandCap,2] after the zero-reset rule
    unsigned long long* cand_qsum; // [kCandCap,2] fixed-point sums of r^2 (units of tsq 2^-44)
    unsigned int* cand_rawcnt;     // [kCandCap,2]
    unsigned int* cand_masks; // [kCandCap, W]
    ContPar* cont_par;              // [kContCap] winner-only runs: contenders
    unsigned long long* cont_q;     // [kContCap,2] fixed-point sums of r^2/tsq
    unsigned int* cont_n;           // [kContCap,2] exact inlier counts
    double* cont_val;               // [kContCap,2] refined value, error bound
    TailState* ts;
    unsigned int cont_cap;          // contender slots in use (kContCap; tests lower it to force the overflow path)
    double dbg_floor_bump;          // tests only: added to the first contender bar (forces the bar-lowering round); 0
    Counters* ctr;      // live counters of the run in flight
    Counters* ctr_snap; // snapshot for the host, written by the last kernel
    unsigned long long* phase_ns; // [8] %globaltimer at the phase boundaries of k_tail
    void* record;  // gcr_record (device), engine-owned
    void* record_user; // optional caller-owned copy target
    unsigned int* best_masks; // [W]
    unsigned int* host_block; // optional: device-mapped pinned host mirror of the whole result block, written by the tail
    // multi-GPU winner exchange over peer memory (NVLink P2P): every rank stores its 64-byte record + a sequence number
    // into slot [parity][rank] of every peer's exchange buffer, then reads the slots of its own buffer
    unsigned long long* xchg_peer[kMaxPeers]; // exchange buffers of all ranks (own one included), device-mapped
    int xchg_world, xchg_rank;                // xchg_world <= 1: no exchange
    unsigned long long xchg_seq;              // 1, 2, 3, ... identical on every rank for the same collective run
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


// Programmatic dependent launch (the three kernels of gcr_run are chained with
// cudaLaunchAttributeProgrammaticStreamSerialization): a kernel lets its successor start launching at once, and the successor
// runs its prologue (barrier set-up, feature staging) under the predecessor's tail, then waits here for the predecessor's
// results.  Both are no-ops in a plain launch.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

} // namespace gcr
