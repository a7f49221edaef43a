// Host side of libgcr_b200.so: buffer management, launch sequence and the extern "C" ABI declared in
// include/gcr_b200.h.  Compiled for sm_100a only, with -fmad=false (see gcr_exact.cuh).
#include "gcr_kernels.cuh"
#include "gcr_fit.cuh"
#include "gcr_lo.cuh"
#include "gcr_front.cuh"
#include "../../include/gcr_b200.h"

#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <vector>
#include <algorithm>
#include <new>
#include <unordered_map>

namespace
{
using namespace gcr;

static_assert(sizeof(Record) == sizeof(gcr_record), "record layout");
static_assert(sizeof(Model4) == 32, "model layout");
static_assert(sizeof(Counters) == 64, "counters layout");

template <typename T>
struct DevBuf
{
    T* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t n)
    {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        const size_t want = n + n / 4 + 64;
        cudaError_t e = cudaMalloc(&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

// pinned host staging that only grows
struct PinnedBuf
{
    unsigned char* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t n)
    {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        const size_t want = n + n / 4 + 256;
        cudaError_t e = cudaMallocHost(reinterpret_cast<void**>(&p), want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release()
    {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
};

} // namespace

struct gcr_engine_s
{
    int device = 0;
    int n_sm = 148;        // SMs every grid of this engine is sized for (all of the device's, or a share of them: gcr_set_sm_share)
    int n_sm_device = 148; // SMs of the device
    int problem = 0; // GCR_PROBLEM_*
    int score_grid[3] = {148, 148, 148}; // per scoring mode: winner-only, per-hypothesis counts, + masks
    bool pdl = true;            // programmatic dependent launch of the scoring kernel and the tail (GCR_PDL=0: plain launches)
    bool last_resident = false; // the last gcr_run used k_score_resident
    bool resident_ok = true;    // k_score_resident usable (GCR_RESIDENT=0 forces the ring kernel: A/B runs)
    long long resident_max_hyp = 262144; // ... for batches up to this many slots (GCR_RESIDENT_MAX_HYP) and slots x features (GCR_RESIDENT_MAX_EVALS).
    long long resident_max_evals = 300000000; // Measured on B200 against the ring kernel: step 59.8 -> 47.4 us at C1 (5k slots x 2 500 features),
                                              // 86.7 -> 85.4 us at C2 (100k x 2 500), 117.8 -> 120.2 us at C4's image size (100k x 5 000), kernel + 4 %
                                              // at C5 (1M x 2 500); with several engines in flight (C4 batches) the ring kernel's small CTAs interleave
                                              // with the other engines' kernels and win by 10 %
    int resident_cost_o = 20;   // cost model of k_score_resident's static partition: orientation block (scale block = 10) ...
    int resident_cost_p = 20;   // ... and a precise tile relative to an FP32 tile (x 10); GCR_COST_O / GCR_COST_P for tuning runs
    int tail_grid = 148;
    long long n_phi_fallbacks = 0;          // gcr_fit_score_models: fits whose device-side phi differed from the host's map replay (diagnostics)
    unsigned int worklist_cap_override = 0; // GCR_WORKLIST_CAP (tests of the overflow report)
    unsigned int cont_cap_override = 0;     // GCR_CONT_CAP (tests of the contender-overflow path of the selection cascade)
    double dbg_floor_bump = 0.0;            // GCR_DEBUG_FLOOR_BUMP (tests of the bar-lowering contender round)
    cudaStream_t own_stream = nullptr, stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, evk0 = nullptr, evk1 = nullptr, ev_wait = nullptr, ev_stage = nullptr;
    PinnedBuf h_feat_stage;    // small pageable feature arrays are copied here and read in place by the preparation kernel
    bool stage_in_use = false; // ... ev_stage marks the end of that kernel
    std::string err;

    // features
    int ns = 0, no = 0, ncs = 0, nco = 0;
    DevBuf<double> d_scale, d_orient;
    const double* prep_src_scale = nullptr; // set_features from mapped host memory: what k_prep_features reads (else null)
    const double* prep_src_orient = nullptr;
    const double* scale_ptr = nullptr; // adopted or owned
    const double* orient_ptr = nullptr;
    DevBuf<float> d_fs, d_fo;
    DevBuf<double> d_fsd, d_fod; // FP64 SoA for the precise tiles
    DevBuf<double> d_aux_s, d_aux_o; // per-feature inputs of the minimal solvers (k_prep_features)
    std::vector<double> h_aux_s, h_aux_o; // GCR_FEATURES_HOST_LIBM: the same columns evaluated by the host's C library
    int ncsd = 0, ncod = 0;
    DevBuf<unsigned int> d_irr_s, d_irr_o;
    DevBuf<Scene> d_scene;

    // hypotheses
    int n_hyp = 0;
    unsigned long long first_hyp = 0;
    bool have_masks = false, have_run = false, have_per_hyp = false;
    DevBuf<int> d_samples;
    const int* samples_view = nullptr; // where the samples of the last run live (own buffer or the caller's device array)
    bool samples_u16 = false;          // ... as uint16 indices (GCR_RUN_SAMPLES_U16)
    DevBuf<Model4> d_models;
    DevBuf<unsigned char> d_flags;
    DevBuf<double> d_sums, d_score;
    DevBuf<unsigned int> d_cnts, d_ninl, d_masks, d_fast_hyp, d_hard_hyp;
    DevBuf<float4> d_pf0, d_pf1;
    DevBuf<ContPar> d_cont_par; // winner-only runs: contenders of the selection cascade (k_tail)
    DevBuf<unsigned long long> d_cont_q;
    DevBuf<unsigned int> d_cont_n;
    DevBuf<double> d_cont_val;
    DevBuf<TailState> d_ts;
    DevBuf<float2> d_err_ab; // winner-only runs: per-hypothesis error bound of the FP32 class values (k_sample_solve -> k_tail)
    DevBuf<uint2> d_worklist;
    DevBuf<unsigned int> d_cand_hyp, d_cand_cnt, d_cand_masks;
    DevBuf<Model4> d_cand_model;
    DevBuf<double> d_cand_score;
    DevBuf<unsigned long long> d_cand_qsum;
    DevBuf<unsigned int> d_cand_rawcnt;
    DevBuf<Counters> d_ctr;
    // result block of a run, one contiguous D2H copy: [gcr_record 64 B][Counters snapshot 64 B][8 phase stamps 64 B][winner masks W words]
    DevBuf<unsigned int> d_best;
    unsigned int* h_best = nullptr; // pinned staging of the same layout
    size_t h_best_words = 0;
    DevBuf<double> d_dbg_in, d_dbg_out;
    DevBuf<Model4> d_user_models;

    // non-minimal fits (gcr_fit_models)
    DevBuf<unsigned char> d_fit_in, d_fit_out;
    DevBuf<double> d_fit_lines, d_fit_partial;
    DevBuf<unsigned int> d_fit_done; // per-fit counters of k_fit_rows, zero between calls
    PinnedBuf h_fit_in, h_fit_out, h_score_io;
    // one round of the local optimisation on the device (gcr_lo_round)
    DevBuf<unsigned int> d_lo_in, d_lo_info, d_lo_idx;
    DevBuf<FitDesc> d_lo_desc;
    PinnedBuf h_lo_out;

    // keypoint front end (gcr_set_keypoints)
    DevBuf<unsigned char> d_kp;
    DevBuf<unsigned int> d_front_table, d_front_counts;
    DevBuf<double> d_front_sizes;
    // rows gcr_get_features returns: the resident features, or everything the last gcr_set_keypoints built
    const double* fetch_scale = nullptr;
    const double* fetch_orient = nullptr;
    int64_t fetch_ns = 0, fetch_no = 0;

    // multi-GPU winner exchange (peer memory)
    DevBuf<unsigned long long> d_xchg;
    void* peer_ptr[kMaxPeers] = {};
    bool peer_ipc[kMaxPeers] = {};
    int xchg_world = 0, xchg_rank = 0;
    unsigned long long xchg_seq = 0;
    bool last_run_exchanged = false, last_run_mirrored = false;
    double last_thr[2] = {-1.0, -1.0};

    int fail(int code, const std::string& msg)
    {
        err = msg;
        return code;
    }
    int cuda_fail(cudaError_t e, const char* what)
    {
        err = std::string(what) + ": " + cudaGetErrorString(e);
        return GCR_ERR_CUDA;
    }
};

// Every ABI entry point works on the engine's device and leaves the caller's current device as it found it (a torch process reads
// the thread's current device with cudaGetDevice: an engine on device 1 must not move later torch allocations there).
struct DeviceGuard
{
    int prev = -1, dev;
    cudaError_t err;
    explicit DeviceGuard(int d) : dev(d)
    {
        err = cudaGetDevice(&prev);
        if (err == cudaSuccess && prev != dev) err = cudaSetDevice(dev);
    }
    ~DeviceGuard()
    {
        if (prev >= 0 && prev != dev) cudaSetDevice(prev);
    }
};
#define GCR_DEVICE(h)                                                      \
    DeviceGuard device_guard_((h)->device);                                \
    if (device_guard_.err != cudaSuccess) return (h)->cuda_fail(device_guard_.err, "cudaSetDevice")

#define GCR_CUDA(call)                                             \
    do                                                             \
    {                                                              \
        cudaError_t e_ = (call);                                   \
        if (e_ != cudaSuccess) return h->cuda_fail(e_, #call);     \
    } while (0)

namespace
{

// r^2 = z P(z), z = tan^2: degree-3 interpolation of P(z) = (atan(sqrt z)/sqrt z)^2 at Chebyshev nodes of
// [0, zmax]; returns the maximum relative error of z P(z) against atan^2 on a dense grid.
double fit_atan2_poly(double zmax, float pc[4], int deg)
{
    auto P = [](double z) {
        if (z < 1e-12) return 1.0 - 2.0 * z / 3.0;
        const double s = std::sqrt(z), a = std::atan(s) / s;
        return a * a;
    };
    const int n = deg + 1; // interpolation nodes
    double x[4], dd[4];
    for (int i = 0; i < n; i++)
    {
        x[i] = 0.5 * zmax * (1.0 + std::cos(M_PI * (2.0 * i + 1.0) / (2.0 * n)));
        dd[i] = P(x[i]);
    }
    // Newton divided differences -> monomial coefficients
    for (int j = 1; j < n; j++)
        for (int i = n - 1; i >= j; i--) dd[i] = (dd[i] - dd[i - 1]) / (x[i] - x[i - j]);
    double c[4] = {dd[n - 1], 0, 0, 0}; // Horner expansion of the Newton form
    int d = 0;
    for (int k = n - 2; k >= 0; k--)
    {
        double nc[4] = {0, 0, 0, 0}; // c(z) <- c(z) * (z - x[k]) + dd[k]
        for (int i = 0; i <= d; i++)
        {
            nc[i + 1] += c[i];
            nc[i] -= c[i] * x[k];
        }
        nc[0] += dd[k];
        d++;
        for (int i = 0; i <= d; i++) c[i] = nc[i];
    }
    for (int i = 0; i < 4; i++) pc[i] = i < n ? static_cast<float>(c[i]) : 0.f;
    double worst = 0.0;
    for (int i = 1; i <= 2000; i++)
    {
        const double z = zmax * i / 2000.0;
        const double approx = ((static_cast<double>(pc[3]) * z + pc[2]) * z + pc[1]) * z + pc[0];
        worst = std::max(worst, std::fabs(approx - P(z)) / P(z));
    }
    return worst;
}

Thresholds make_thresholds_uncached(const double thr_in[2], int problem);

// The orientation polynomial is fitted and verified on a 2000-point grid (~30 us of host time): callers repeat the same
// thresholds run after run, so the last result is kept.
Thresholds make_thresholds(const double thr_in[2], int problem)
{
    struct Entry { double t0, t1; int problem; bool valid; Thresholds T; };
    static thread_local Entry cache = {0.0, 0.0, 0, false, {}};
    if (cache.valid && cache.t0 == thr_in[0] && cache.t1 == thr_in[1] && cache.problem == problem) return cache.T;
    cache.T = make_thresholds_uncached(thr_in, problem);
    cache.t0 = thr_in[0];
    cache.t1 = thr_in[1];
    cache.problem = problem;
    cache.valid = true;
    return cache.T;
}

Thresholds make_thresholds_uncached(const double thr_in[2], int problem)
{
    Thresholds T;
    std::memset(&T, 0, sizeof(T));
    // scale-only problems have one class: the orientation entries are placeholders that are never compared against
    const double thr[2] = {thr_in[0], problem == 0 ? thr_in[1] : 0.05};
    T.problem = problem;
    T.min_inl[0] = problem == 0 ? 2u : 3u; // sampleSize(): two_sift.hpp:28-32 {2,2}; three_sift.hpp:30 {3}
    T.min_inl[1] = problem == 0 ? 2u : 0u;
    T.tsq[0] = (2.25 * thr[0]) * thr[0]; // S1: MSAC_scoring_function.hpp:64
    T.tsq[1] = (2.25 * thr[1]) * thr[1];
    const double tr0 = std::sqrt(T.tsq[0]), tr1 = std::sqrt(T.tsq[1]);
    T.Tp = static_cast<float>(tr0 / (3.0 * kLn2));
    T.scale_unit = static_cast<float>((3.0 * kLn2) * (3.0 * kLn2));
    T.exp2_neg_Tp = std::exp2(-static_cast<double>(T.Tp));
    T.exp2_pos_Tp = std::exp2(static_cast<double>(T.Tp));
    T.orient_all_exact = 0;
    if (!(tr1 < 0.6)) // beyond ~34 degrees the tan-domain test loses its margin: exact path
    {
        T.orient_all_exact = 1;
        T.tanT = 0.5f;
    }
    else
    {
        const double tt = std::tan(tr1);
        T.tanT = static_cast<float>(tt);
        const double zmax = (tt * 1.01 + 1e-6) * (tt * 1.01 + 1e-6);
        T.poly_deg = 2;
        if (fit_atan2_poly(zmax, T.pc, 2) > 5e-7)
        {
            T.poly_deg = 3;
            if (fit_atan2_poly(zmax, T.pc, 3) > 5e-7) T.orient_all_exact = 1;
        }
    }
    {
        const int d = T.poly_deg == 2 ? 2 : 3;
        T.pc_lead = T.pc[d];
        for (int i = 0; i < d; i++) T.pm[i] = static_cast<float>(static_cast<double>(T.pc[d - 1 - i]) / static_cast<double>(T.pc[d]));
        if (d == 2) T.pm[2] = 0.f;
        if (!(std::fabs(T.pc_lead) > 1e-12f)) T.orient_all_exact = 1; // degenerate fit: no monic form
    }
    // rounding of tanT to float + rcp (1 ulp) + two multiplies + polynomial-free compare
    T.t_band0 = static_cast<float>(26.0 * kU + 4.0 * kU * T.tanT);
    T.t_lo0 = std::nextafter(T.tanT - T.t_band0, -1.0f);
    T.t_hi0 = std::nextafter(T.tanT + T.t_band0, 2.0f);
    // winner-only runs (score_chunk_w): dq = (A - lg2 v) / Tp and r^2 / tsq1 = z (pn0 + z (pn1 + z (pn2 + z pn3)))
    T.inv_Tp = static_cast<float>(1.0 / static_cast<double>(T.Tp));
    T.sqrt_tsq1 = static_cast<float>(tr1);
    for (int i = 0; i < 4; i++) T.pn[i] = T.tsq[1] > 0.0 ? static_cast<float>(static_cast<double>(T.pc[i]) / T.tsq[1]) : 0.f;
    if (problem == 0 && !T.orient_all_exact)
    {
        // beyond the fitted range the polynomial only has to keep the term at zero: z P(z) / tsq >= 1 for every tan^2 up to 1
        const double tt = static_cast<double>(T.tanT) * 1.01 + 1e-6, z0 = tt * tt;
        for (int i = 0; i <= 4000 && !T.orient_all_exact; i++)
        {
            const double z = z0 + (1.0 - z0) * i / 4000.0;
            const double hz = ((static_cast<double>(T.pn[3]) * z + T.pn[2]) * z + T.pn[1]) * z + T.pn[0];
            if (!(z * hz >= 1.0 + 1e-3)) T.orient_all_exact = 1;
        }
    }
    return T;
}

Buffers make_buffers(gcr_engine_s* h, unsigned long long seed, bool have_samples, void* record_user)
{
    Buffers B;
    std::memset(&B, 0, sizeof(B));
    B.scale = h->scale_ptr;
    B.orient = h->orient_ptr;
    B.aux_s = h->d_aux_s.p;
    B.aux_o = h->d_aux_o.p;
    B.ns = h->ns;
    B.no = h->no;
    B.fs = h->d_fs.p;
    B.fo = h->d_fo.p;
    B.ncs = h->ncs;
    B.nco = h->nco;
    B.fsd = h->d_fsd.p;
    B.fod = h->d_fod.p;
    B.ncsd = h->ncsd;
    B.ncod = h->ncod;
    B.scene = h->d_scene.p;
    B.irr_s = h->d_irr_s.p;
    B.irr_o = h->d_irr_o.p;
    B.n_hyp = h->n_hyp;
    B.first_hyp = h->first_hyp;
    B.seed = seed;
    B.samples = const_cast<int*>(h->samples_view);
    B.have_samples = have_samples ? 1 : 0;
    B.samples_u16 = have_samples && h->samples_u16 ? 1 : 0;
    B.models = h->d_models.p;
    B.flags = h->d_flags.p;
    B.sums = h->d_sums.p;
    B.cnts = h->d_cnts.p;
    B.score = h->d_score.p;
    B.n_inl = h->d_ninl.p;
    B.masks = h->have_masks ? h->d_masks.p : nullptr;
    B.ws = (h->ns + 31) / 32;
    B.wmask = B.ws + (h->no + 31) / 32;
    B.pf0 = h->d_pf0.p;
    B.pf1 = h->d_pf1.p;
    B.fast_hyp = h->d_fast_hyp.p;
    B.hard_hyp = h->d_hard_hyp.p;
    B.worklist = h->d_worklist.p;
    B.worklist_cap = static_cast<unsigned int>(std::min<size_t>(h->d_worklist.cap, 0x7fffffffu));
    if (h->worklist_cap_override) B.worklist_cap = std::min(B.worklist_cap, h->worklist_cap_override); // tests: force an overflow
    B.cand_hyp = h->d_cand_hyp.p;
    B.cand_model = h->d_cand_model.p;
    B.cand_score = h->d_cand_score.p;
    B.cand_cnt = h->d_cand_cnt.p;
    B.cand_qsum = h->d_cand_qsum.p;
    B.cand_rawcnt = h->d_cand_rawcnt.p;
    B.cand_masks = h->d_cand_masks.p;
    B.ctr = h->d_ctr.p;
    B.record = h->d_best.p;
    B.ctr_snap = reinterpret_cast<Counters*>(h->d_best.p + 16);
    B.phase_ns = reinterpret_cast<unsigned long long*>(h->d_best.p + 32);
    B.record_user = record_user;
    B.best_masks = h->d_best.p + kResultHeaderWords;
    B.err_ab = h->d_err_ab.p;
    B.cont_par = h->d_cont_par.p;
    B.cont_q = h->d_cont_q.p;
    B.cont_n = h->d_cont_n.p;
    B.cont_val = h->d_cont_val.p;
    B.ts = h->d_ts.p;
    B.cont_cap = h->cont_cap_override ? std::min<unsigned int>(h->cont_cap_override, kContCap) : kContCap;
    B.dbg_floor_bump = h->dbg_floor_bump;
    B.host_block = nullptr;
    for (int p = 0; p < kMaxPeers; p++) B.xchg_peer[p] = static_cast<unsigned long long*>(h->peer_ptr[p]);
    B.xchg_world = 0; // set by gcr_run for exchanging runs
    B.xchg_rank = h->xchg_rank;
    B.xchg_seq = 0;
    return B;
}

int ensure_common(gcr_engine_s* h)
{
    const size_t W = static_cast<size_t>((h->ns + 31) / 32 + (h->no + 31) / 32);
    if (h->d_ctr.cap < 1)
    {
        // live run counters: cleared by the last kernel of every run after it has snapshotted them into the result block
        GCR_CUDA(h->d_ctr.ensure(1));
        GCR_CUDA(cudaMemset(h->d_ctr.p, 0, sizeof(Counters)));
    }
    if (h->d_ts.cap < 1)
    {
        GCR_CUDA(h->d_ts.ensure(1));
        GCR_CUDA(cudaMemset(h->d_ts.p, 0, sizeof(TailState)));
    }
    GCR_CUDA(h->d_cont_par.ensure(kContCap));
    GCR_CUDA(h->d_cont_q.ensure(2 * kContCap));
    GCR_CUDA(h->d_cont_n.ensure(2 * kContCap));
    GCR_CUDA(h->d_cont_val.ensure(2 * kContCap));
    GCR_CUDA(h->d_cand_hyp.ensure(kCandCap));
    GCR_CUDA(h->d_cand_model.ensure(kCandCap));
    GCR_CUDA(h->d_cand_score.ensure(kCandCap));
    GCR_CUDA(h->d_cand_cnt.ensure(2 * kCandCap));
    GCR_CUDA(h->d_cand_qsum.ensure(2 * kCandCap));
    GCR_CUDA(h->d_cand_rawcnt.ensure(2 * kCandCap));
    GCR_CUDA(h->d_cand_masks.ensure(W * kCandCap));
    GCR_CUDA(h->d_best.ensure(kResultHeaderWords + W));
    if (h->h_best_words < kResultHeaderWords + W)
    {
        if (h->h_best) cudaFreeHost(h->h_best);
        h->h_best = nullptr;
        h->h_best_words = 0;
        GCR_CUDA(cudaMallocHost(&h->h_best, sizeof(unsigned int) * (kResultHeaderWords + W + 64)));
        h->h_best_words = kResultHeaderWords + W + 64;
    }
    return GCR_OK;
}

} // namespace

static int peer_release(gcr_engine_s* h)
{
    for (int p = 0; p < kMaxPeers; p++)
    {
        if (h->peer_ipc[p] && h->peer_ptr[p]) cudaIpcCloseMemHandle(h->peer_ptr[p]);
        h->peer_ptr[p] = nullptr;
        h->peer_ipc[p] = false;
    }
    h->xchg_world = 0;
    h->xchg_rank = 0;
    h->xchg_seq = 0;
    return GCR_OK;
}

static int peer_buffer(gcr_engine_s* h)
{
    if (!h->d_xchg.p)
    {
        GCR_DEVICE(h);
        GCR_CUDA(h->d_xchg.ensure(kXchgWords));
        GCR_CUDA(cudaMemset(h->d_xchg.p, 0, sizeof(unsigned long long) * kXchgWords));
    }
    return GCR_OK;
}

// device-side address of a page-locked, device-mapped host array (nullptr for pageable memory or when disabled)
static const double* mapped_host_pointer(const void* p)
{
    static const bool disabled = std::getenv("GCR_NO_ZEROCOPY") != nullptr;
    cudaPointerAttributes attr;
    if (!disabled && p && cudaPointerGetAttributes(&attr, p) == cudaSuccess && attr.type == cudaMemoryTypeHost && attr.devicePointer &&
        reinterpret_cast<uintptr_t>(attr.devicePointer) % 8 == 0)
        return static_cast<const double*>(attr.devicePointer);
    cudaGetLastError();
    return nullptr;
}

// Derived feature planes of the rows at h->scale_ptr / h->orient_ptr (device memory), sizes in h->ns / h->no
static int finish_features(gcr_handle h)
{
    h->ncs = (h->ns + kChunk - 1) / kChunk;
    h->nco = (h->no + kChunk - 1) / kChunk;
    h->have_run = false;
    GCR_CUDA(h->d_fs.ensure(static_cast<size_t>(h->ncs) * kChunkFloats));
    GCR_CUDA(h->d_fo.ensure(static_cast<size_t>(h->nco) * kChunkFloats));
    h->ncsd = (h->ns + kChunk / 2 - 1) / (kChunk / 2);
    h->ncod = (h->no + kChunk / 2 - 1) / (kChunk / 2);
    GCR_CUDA(h->d_fsd.ensure(static_cast<size_t>(h->ncsd) * (kChunkFloats / 2)));
    GCR_CUDA(h->d_fod.ensure(static_cast<size_t>(h->ncod) * (kChunkFloats / 2)));
    GCR_CUDA(h->d_aux_s.ensure(static_cast<size_t>(std::max(h->ns, 1))));
    GCR_CUDA(h->d_aux_o.ensure(2 * static_cast<size_t>(std::max(h->no, 1))));
    GCR_CUDA(h->d_irr_s.ensure(kIrrCap));
    GCR_CUDA(h->d_irr_o.ensure(kIrrCap));
    GCR_CUDA(h->d_scene.ensure(1));
    GCR_CUDA(cudaMemsetAsync(h->d_scene.p, 0, sizeof(Scene), h->stream));
    const int n = std::max(h->ncs, h->nco) * kChunk;
    const bool in_place = h->prep_src_scale != nullptr;
    k_prep_features<<<(n + 255) / 256, 256, 0, h->stream>>>(in_place ? h->prep_src_scale : h->scale_ptr, h->ns,
                                                           in_place && h->prep_src_orient ? h->prep_src_orient : h->orient_ptr, h->no, h->d_fs.p,
                                                           h->d_fo.p, h->ncs, h->nco, h->d_fsd.p, h->d_fod.p, h->ncsd, h->ncod, h->d_scene.p, h->d_irr_s.p,
                                                           h->d_irr_o.p, h->d_aux_s.p, h->d_aux_o.p, h->problem, in_place ? h->d_scale.p : nullptr,
                                                           in_place && h->prep_src_orient ? h->d_orient.p : nullptr);
    h->prep_src_scale = h->prep_src_orient = nullptr;
    GCR_CUDA(cudaGetLastError());
    return ensure_common(h);
}


// Launch of a kernel that follows another kernel of the same call: programmatic stream serialisation, the kernel waits at its top
// (griddepcontrol.wait) for its predecessor, so only the launch latency overlaps.  Kernels launched this way touch nothing before the wait.
template <typename... KArgs, typename... Args>
static cudaError_t launch_chained(gcr_engine_s* h, void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t s, Args... args)
{
    cudaLaunchAttribute at;
    at.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at.val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.stream = s;
    cfg.attrs = &at;
    cfg.numAttrs = h->pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ------------------------------------------------------------------------------------------------
// gcr_score_models_static: scratch of the batched FP64 model scoring in a file-scope device array instead of engine-owned allocations
// (callers that must not allocate: CUDA graph capture, tight serving loops).  1 MiB per GPU that loads the library, held until exit.
// In HOST code the array's name is a host-side stand-in, not a device address: the address the kernels get comes from
// cudaGetSymbolAddress, resolved once per device with that device current (static_scratch_address).
constexpr size_t kStaticScratchBytes = size_t(1) << 20;
constexpr int kStaticMaxModels = 1024;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];
// fixed part of the layout, for kStaticMaxModels models: [Model4 32 KB][hyp u32 4 KB][qsum u64 x2 16 KB][rawcnt u32 x2 8 KB]
// [score f64 8 KB][cnt u32 x2 8 KB] = 76 KB; the rest holds the packed masks of the n models (n x W words)
constexpr size_t kStaticOffModel = 0, kStaticOffHyp = kStaticOffModel + 32 * kStaticMaxModels, kStaticOffQsum = kStaticOffHyp + 4 * kStaticMaxModels,
                 kStaticOffRaw = kStaticOffQsum + 16 * kStaticMaxModels, kStaticOffScore = kStaticOffRaw + 8 * kStaticMaxModels,
                 kStaticOffCnt = kStaticOffScore + 8 * kStaticMaxModels, kStaticOffMasks = kStaticOffCnt + 8 * kStaticMaxModels;
static_assert(kStaticOffMasks % 256 == 0 && kStaticOffMasks < kStaticScratchBytes, "static scratch layout");
constexpr int kStaticMaxDevices = 64;
static unsigned char* g_static_scratch_ptr[kStaticMaxDevices] = {};

// device address of g_static_scratch on the CURRENT device (= `device`); no allocation
static cudaError_t static_scratch_address(int device, unsigned char** out)
{
    if (!g_static_scratch_ptr[device])
    {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        g_static_scratch_ptr[device] = static_cast<unsigned char*>(p); // (racing first calls resolve the same address)
    }
    *out = g_static_scratch_ptr[device];
    return cudaSuccess;
}

// Device-side failures of the last run, from the counter snapshot its last kernel wrote into the result block.  The reference throws
// rather than continue on bad input (GCRANSAC.h:222-228); an asynchronous caller (gcr_run with out == NULL) learns of them from
// gcr_best / gcr_fetch.
static int check_run_counters(gcr_engine_s* h, const Counters& c)
{
    if (c.xchg_timeout) return h->fail(GCR_ERR_STATE, "winner exchange timed out: a peer rank did not run the same collective call");
    if (c.n_bad_sample) return h->fail(GCR_ERR_INVALID, "replayed sample indices out of range");
    if (c.n_worklist_overflow) return h->fail(GCR_ERR_STATE, "guard-band worklist overflow: masks / counts of this run are incomplete");
    return GCR_OK;
}

extern "C"
{

const char* gcr_version(void) { return "gcr_b200 0.1 sm_100a"; }

// grids of the persistent scoring kernel (per mode) and of the cooperative tail for h->n_sm SMs
static cudaError_t size_grids(gcr_engine_s* h)
{
    cudaError_t e = cudaSuccess;
    for (int m = 0; m < 3 && e == cudaSuccess; m++)
    {
        int per_sm = 1;
        e = m == 0 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_score_fast<0>, kScoreThreads, sizeof(ScoreSmem))
                   : (m == 1 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_score_fast<1>, kScoreThreads, sizeof(ScoreSmem))
                             : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_score_fast<2>, kScoreThreads, sizeof(ScoreSmem)));
        h->score_grid[m] = h->n_sm * std::max(1, per_sm);
    }
    if (e == cudaSuccess)
    {
        int per_sm = 1;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_tail, kExactThreads, 0);
        h->tail_grid = h->n_sm * std::max(1, std::min(per_sm, 2));
        if (const char* v = std::getenv("GCR_TAIL_GRID")) h->tail_grid = std::max(1, std::min(std::atoi(v), h->n_sm * std::max(1, per_sm)));
    }
    return e;
}

int gcr_create(int device, gcr_handle* out)
{
    if (!out) return GCR_ERR_INVALID;
    *out = nullptr;
    gcr_engine_s* h = new (std::nothrow) gcr_engine_s();
    if (!h) return GCR_ERR_NOMEM;
    h->device = device;
    DeviceGuard device_guard_(device);
    cudaError_t e = device_guard_.err;
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreate(&h->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&h->ev1);
    if (e == cudaSuccess) e = cudaEventCreate(&h->evk0);
    if (e == cudaSuccess) e = cudaEventCreate(&h->evk1);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_wait, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_stage, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, device);
    h->n_sm_device = h->n_sm;
    if (const char* v = std::getenv("GCR_SM_SHARE")) // (experiments: the default share of engines created from here on, see gcr_set_sm_share)
        if (e == cudaSuccess && std::atoi(v) > 1) h->n_sm = std::max(1, h->n_sm_device / std::atoi(v));
    const int resident_smem = kResidentMaxChunks * kChunkBytes + 64;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_score_fast<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScoreSmem));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_score_fast<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScoreSmem));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_score_fast<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScoreSmem));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_score_resident<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, resident_smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_score_resident<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, resident_smem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_score_resident<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, resident_smem);
    if (const char* v = std::getenv("GCR_RESIDENT")) h->resident_ok = std::atoi(v) != 0;
    if (const char* v = std::getenv("GCR_PDL")) h->pdl = std::atoi(v) != 0;
    if (const char* v = std::getenv("GCR_CONT_CAP")) h->cont_cap_override = static_cast<unsigned int>(std::max(1, std::atoi(v)));
    if (const char* v = std::getenv("GCR_DEBUG_FLOOR_BUMP")) h->dbg_floor_bump = std::atof(v);
    if (const char* v = std::getenv("GCR_WORKLIST_CAP")) h->worklist_cap_override = static_cast<unsigned int>(std::max(1, std::atoi(v)));
    if (const char* v = std::getenv("GCR_RESIDENT_MAX_HYP")) h->resident_max_hyp = h->resident_max_evals = std::atoll(v);
    if (const char* v = std::getenv("GCR_RESIDENT_MAX_EVALS")) h->resident_max_evals = std::atoll(v);
    if (const char* v = std::getenv("GCR_COST_O")) h->resident_cost_o = std::max(1, std::atoi(v));
    if (const char* v = std::getenv("GCR_COST_P")) h->resident_cost_p = std::max(1, std::atoi(v));
    if (e == cudaSuccess) e = size_grids(h);
    if (e == cudaSuccess)
    {
        int coop = 0;
        e = cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device);
        if (e == cudaSuccess && !coop)
        {
            h->err = "device does not support cooperative launches";
            *out = h;
            return GCR_ERR_CUDA;
        }
    }
    if (e != cudaSuccess)
    {
        // keep the handle so that gcr_last_error can be read; every later call fails with GCR_ERR_CUDA
        h->cuda_fail(e, "gcr_create");
        *out = h;
        return GCR_ERR_CUDA;
    }
    h->stream = h->own_stream;
    *out = h;
    return GCR_OK;
}

int gcr_destroy(gcr_handle h)
{
    if (!h) return GCR_ERR_INVALID;
    DeviceGuard device_guard_(h->device);
    if (h->own_stream) cudaStreamSynchronize(h->own_stream);
    peer_release(h);
    h->d_xchg.release();
    h->d_scale.release(); h->d_orient.release(); h->d_fs.release(); h->d_fo.release(); h->d_fsd.release(); h->d_fod.release();
    h->d_aux_s.release(); h->d_aux_o.release();
    h->d_irr_s.release(); h->d_irr_o.release(); h->d_scene.release(); h->d_samples.release();
    h->d_models.release(); h->d_flags.release(); h->d_sums.release(); h->d_score.release();
    h->d_cnts.release(); h->d_ninl.release(); h->d_masks.release(); h->d_fast_hyp.release();
    h->d_hard_hyp.release(); h->d_pf0.release(); h->d_pf1.release(); h->d_err_ab.release(); h->d_cont_par.release(); h->d_cont_q.release(); h->d_cont_n.release();
    h->d_cont_val.release(); h->d_ts.release(); h->d_worklist.release();
    h->d_cand_hyp.release(); h->d_cand_cnt.release(); h->d_cand_masks.release(); h->d_best.release();
    if (h->h_best) cudaFreeHost(h->h_best);
    h->d_cand_model.release(); h->d_cand_score.release(); h->d_cand_qsum.release(); h->d_cand_rawcnt.release(); h->d_ctr.release();
    h->d_dbg_in.release(); h->d_dbg_out.release(); h->d_user_models.release();
    h->d_fit_in.release(); h->d_fit_out.release(); h->d_fit_lines.release(); h->d_fit_partial.release(); h->d_fit_done.release();
    h->h_fit_in.release(); h->h_fit_out.release(); h->h_score_io.release();
    h->d_lo_in.release(); h->d_lo_info.release(); h->d_lo_idx.release(); h->d_lo_desc.release(); h->h_lo_out.release();
    h->d_kp.release(); h->d_front_table.release(); h->d_front_counts.release(); h->d_front_sizes.release();
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->evk0) cudaEventDestroy(h->evk0);
    if (h->evk1) cudaEventDestroy(h->evk1);
    if (h->ev_wait) cudaEventDestroy(h->ev_wait);
    if (h->ev_stage) cudaEventDestroy(h->ev_stage);
    h->h_feat_stage.release();
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    delete h;
    return GCR_OK;
}

#ifdef GCR_TRACE
// diagnostic builds only (tools/trace_score.py): copies the scoring kernels' timeline rows and clears them
int gcr_debug_read_trace(void* dst, size_t bytes)
{
    cudaDeviceSynchronize();
    if (cudaMemcpyFromSymbol(dst, gcr::g_trace, std::min(bytes, sizeof(gcr::g_trace))) != cudaSuccess) return -1;
    void* p = nullptr;
    cudaGetSymbolAddress(&p, gcr::g_trace);
    cudaMemset(p, 0, sizeof(gcr::g_trace));
    return 0;
}
#endif

const char* gcr_last_error(gcr_handle h) { return h ? h->err.c_str() : "null handle"; }

int gcr_set_stream(gcr_handle h, void* cuda_stream)
{
    if (!h) return GCR_ERR_INVALID;
    h->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : h->own_stream;
    return GCR_OK;
}

int gcr_wait_stream(gcr_handle h, void* producer_stream)
{
    if (!h) return GCR_ERR_INVALID;
    if (!h->stream) return h->fail(GCR_ERR_STATE, "engine was not created successfully");
    GCR_DEVICE(h);
    cudaStream_t p = static_cast<cudaStream_t>(producer_stream);
    if (p == h->stream) return GCR_OK; // same stream: already ordered
    GCR_CUDA(cudaEventRecord(h->ev_wait, p));
    GCR_CUDA(cudaStreamWaitEvent(h->stream, h->ev_wait, 0));
    return GCR_OK;
}

int gcr_set_features(gcr_handle h, const double* scale, int64_t ns, const double* orient, int64_t no, int where)
{
    if (!h) return GCR_ERR_INVALID;
    const bool on_device = (where & GCR_FEATURES_ON_DEVICE) != 0, host_libm = (where & GCR_FEATURES_HOST_LIBM) != 0;
    if (on_device && host_libm) return h->fail(GCR_ERR_INVALID, "GCR_FEATURES_HOST_LIBM needs host arrays");
    if (!h->stream) return h->fail(GCR_ERR_STATE, "engine was not created successfully");
    if (h->problem == 0)
    {
        if (!scale || !orient) return h->fail(GCR_ERR_INVALID, "null feature pointer");
        // bindings.cpp:206-224: at least 2 rows per class (the minimal sample size)
        if (ns < 2 || no < 2) return h->fail(GCR_ERR_INVALID, "each feature class needs at least 2 rows of 3 columns");
    }
    else
    {
        // scale-only problems (bindings.cpp:19-179): one class, minimal sample of 3 rows; the orientation class is empty
        if (!scale) return h->fail(GCR_ERR_INVALID, "null feature pointer");
        if (ns < 3) return h->fail(GCR_ERR_INVALID, "scale features need at least 3 rows of 3 columns");
        if (no != 0) return h->fail(GCR_ERR_INVALID, "scale-only problems take no orientation features (no = 0)");
        orient = nullptr;
    }
    if (ns > (1ll << 30) || no > (1ll << 30)) return h->fail(GCR_ERR_INVALID, "feature count too large");
    GCR_DEVICE(h);
    h->ns = static_cast<int>(ns);
    h->no = static_cast<int>(no);
    bool staged = false;
    if (on_device)
    {
        h->scale_ptr = scale;
        h->orient_ptr = orient;
    }
    else
    {
        GCR_CUDA(h->d_scale.ensure(3 * static_cast<size_t>(ns)));
        GCR_CUDA(h->d_orient.ensure(3 * static_cast<size_t>(no)));
        // page-locked, device-mapped host arrays are read in place by the preparation kernel, which then also writes the
        // device copy (no separate host-to-device copies in front of it)
        h->prep_src_scale = mapped_host_pointer(scale);
        h->prep_src_orient = no > 0 ? mapped_host_pointer(orient) : nullptr;
        if (!h->prep_src_scale || (no > 0 && !h->prep_src_orient))
        {
            h->prep_src_scale = h->prep_src_orient = nullptr;
            // Pageable arrays.  Small ones (an image's few thousand features) are copied into the engine's own page-locked block and take
            // the in-place route: two cudaMemcpyAsync calls from pageable memory cost 54 us of host time at C1's 2 500 features, this
            // copy 3 us.  The block is reused by the next call, which first waits for the preparation kernel that reads it.
            constexpr size_t kStageBytes = size_t(1) << 20;
            const size_t bs = sizeof(double) * 3 * static_cast<size_t>(ns), bo = sizeof(double) * 3 * static_cast<size_t>(no);
            static const bool no_zero_copy = std::getenv("GCR_NO_ZEROCOPY") != nullptr;
            if (!no_zero_copy && bs + bo <= kStageBytes)
            {
                if (h->stage_in_use) GCR_CUDA(cudaEventSynchronize(h->ev_stage));
                GCR_CUDA(h->h_feat_stage.ensure(bs + bo));
                std::memcpy(h->h_feat_stage.p, scale, bs);
                if (no > 0) std::memcpy(h->h_feat_stage.p + bs, orient, bo);
                h->prep_src_scale = mapped_host_pointer(h->h_feat_stage.p);
                h->prep_src_orient = no > 0 && h->prep_src_scale ? mapped_host_pointer(h->h_feat_stage.p + bs) : nullptr;
                staged = h->prep_src_scale && (no == 0 || h->prep_src_orient);
                if (!staged) h->prep_src_scale = h->prep_src_orient = nullptr;
            }
            if (!staged)
            {
                GCR_CUDA(cudaMemcpyAsync(h->d_scale.p, scale, sizeof(double) * 3 * ns, cudaMemcpyHostToDevice, h->stream));
                if (no > 0) GCR_CUDA(cudaMemcpyAsync(h->d_orient.p, orient, sizeof(double) * 3 * no, cudaMemcpyHostToDevice, h->stream));
            }
        }
        h->scale_ptr = h->d_scale.p;
        h->orient_ptr = h->d_orient.p;
    }
    h->fetch_scale = h->scale_ptr;
    h->fetch_orient = h->orient_ptr;
    h->fetch_ns = h->ns;
    h->fetch_no = h->no;
    const int rc = finish_features(h);
    if (rc == GCR_OK && staged)
    {
        GCR_CUDA(cudaEventRecord(h->ev_stage, h->stream));
        h->stage_in_use = true;
    }
    if (rc != GCR_OK || !host_libm) return rc;
    // Bit-exact replay mode: the solvers' per-FEATURE transcendental columns -- pow(s, 1/3) of two_sift.hpp:288-299 (three_sift.hpp:153-160,
    // three_sift_original.hpp: pow(s, -1/3)) and the sine / cosine of lineFromPointAndAngle (math_utils.hpp:104-109) -- evaluated by the host's
    // C library, i.e. by the very functions the reference calls, and written over the device-evaluated columns.  isValidSample and the
    // 3 x 4 system are then IEEE-basic operations on identical inputs: flags are the reference's bit for bit also where the vanishing point
    // sits on a vertex of the sample hull (the knife edge that is the last ulp of sin / cos).
    h->h_aux_s.resize(static_cast<size_t>(ns));
    h->h_aux_o.resize(2 * static_cast<size_t>(no));
    for (int64_t i = 0; i < ns; ++i) h->h_aux_s[i] = h->problem == 2 ? std::pow(scale[3 * i + 2], -1.0 / 3.0) : std::pow(scale[3 * i + 2], 1.0 / 3.0);
    for (int64_t i = 0; i < no; ++i)
    {
        const double t = orient[3 * i + 2];
        h->h_aux_o[2 * i] = std::sin(t);
        h->h_aux_o[2 * i + 1] = std::cos(t);
    }
    GCR_CUDA(cudaMemcpyAsync(h->d_aux_s.p, h->h_aux_s.data(), sizeof(double) * ns, cudaMemcpyHostToDevice, h->stream));
    if (no > 0) GCR_CUDA(cudaMemcpyAsync(h->d_aux_o.p, h->h_aux_o.data(), sizeof(double) * 2 * no, cudaMemcpyHostToDevice, h->stream));
    GCR_CUDA(cudaStreamSynchronize(h->stream)); // pageable staging: the vectors may be reused by the next call
    return GCR_OK;
}

int gcr_set_keypoints(gcr_handle h, const void* keypoints, int64_t n, int is_f32, int on_device, int64_t* n_scale,
                      int64_t* n_orient, double* orient_sizes)
{
    if (!h) return GCR_ERR_INVALID;
    if (!h->stream) return h->fail(GCR_ERR_STATE, "engine was not created successfully");
    if (!keypoints || n < 1 || n > (1ll << 30)) return h->fail(GCR_ERR_INVALID, "gcr_set_keypoints: 1 <= n <= 2^30 rows of (x, y, size, angle)");
    GCR_DEVICE(h);
    const size_t row_bytes = is_f32 ? 16 : 32;
    const void* d_kp = keypoints;
    cudaStream_t s = h->stream;
    if (on_device)
    {
        if (reinterpret_cast<uintptr_t>(keypoints) % 16 != 0) return h->fail(GCR_ERR_INVALID, "gcr_set_keypoints: device rows must be 16-byte aligned");
    }
    else
    {
        GCR_CUDA(h->d_kp.ensure(row_bytes * static_cast<size_t>(n)));
        GCR_CUDA(cudaMemcpyAsync(h->d_kp.p, keypoints, row_bytes * static_cast<size_t>(n), cudaMemcpyHostToDevice, s));
        d_kp = h->d_kp.p;
    }
    size_t cap = 1024;
    while (cap < 2 * static_cast<size_t>(n)) cap <<= 1;
    const unsigned int nb = static_cast<unsigned int>((n + kFrontThreads - 1) / kFrontThreads);
    GCR_CUDA(h->d_front_table.ensure(cap));
    GCR_CUDA(h->d_front_counts.ensure(2 * static_cast<size_t>(nb) + 2 + 2 * ((static_cast<size_t>(n) + 31) / 32)));
    GCR_CUDA(h->d_scale.ensure(3 * static_cast<size_t>(n)));
    GCR_CUDA(h->d_orient.ensure(3 * static_cast<size_t>(n)));
    GCR_CUDA(h->d_front_sizes.ensure(static_cast<size_t>(n)));
    GCR_CUDA(cudaMemsetAsync(h->d_front_table.p, 0xff, sizeof(unsigned int) * cap, s));
    FrontBuffers F;
    F.kp = d_kp;
    F.is_f32 = is_f32 ? 1 : 0;
    F.n = static_cast<unsigned int>(n);
    F.table = h->d_front_table.p;
    F.cap_mask = static_cast<unsigned int>(cap - 1);
    F.block_counts = h->d_front_counts.p;
    F.n_blocks = nb;
    F.scale = h->d_scale.p;
    F.orient = h->d_orient.p;
    F.sizes = h->d_front_sizes.p;
    F.votes = h->d_front_counts.p + 2 * static_cast<size_t>(nb) + 2;
    k_front_insert<<<nb, kFrontThreads, 0, s>>>(F);
    k_front_compact<false><<<nb, kFrontThreads, 0, s>>>(F);
    k_front_scan<<<1, kFrontScanThreads, 0, s>>>(F);
    k_front_compact<true><<<nb, kFrontThreads, 0, s>>>(F);
    GCR_CUDA(cudaGetLastError());
    unsigned int totals[2] = {0, 0};
    GCR_CUDA(cudaMemcpyAsync(totals, h->d_front_counts.p + 2 * static_cast<size_t>(nb), sizeof(totals), cudaMemcpyDeviceToHost, s));
    GCR_CUDA(cudaStreamSynchronize(s));
    if (n_scale) *n_scale = totals[0];
    if (n_orient) *n_orient = totals[1];
    if (orient_sizes && totals[1])
        GCR_CUDA(cudaMemcpy(orient_sizes, h->d_front_sizes.p, sizeof(double) * totals[1], cudaMemcpyDeviceToHost));
    h->ns = h->no = 0;
    h->have_run = false;
    h->fetch_scale = h->d_scale.p;
    h->fetch_orient = h->d_orient.p;
    h->fetch_ns = totals[0];
    h->fetch_no = totals[1];
    const bool hybrid = h->problem == 0;
    // fewer rows than a minimal sample: the rows can be fetched, but the engine holds no features to run on
    if (totals[0] < (hybrid ? 2u : 3u) || (hybrid && totals[1] < 2u)) return GCR_OK;
    h->ns = static_cast<int>(totals[0]);
    h->no = hybrid ? static_cast<int>(totals[1]) : 0;
    h->scale_ptr = h->d_scale.p;
    h->orient_ptr = hybrid ? h->d_orient.p : nullptr;
    return finish_features(h);
}

int gcr_set_sm_share(gcr_handle h, int share)
{
    if (!h) return GCR_ERR_INVALID;
    if (!h->stream) return h->fail(GCR_ERR_STATE, "engine was not created successfully");
    if (share < 1 || share > h->n_sm_device) return h->fail(GCR_ERR_INVALID, "gcr_set_sm_share: 1 <= share <= SMs of the device");
    GCR_DEVICE(h);
    h->n_sm = std::max(1, h->n_sm_device / share);
    GCR_CUDA(size_grids(h));
    return GCR_OK;
}

int gcr_set_problem(gcr_handle h, int problem)
{
    if (!h) return GCR_ERR_INVALID;
    if (problem < GCR_PROBLEM_HYBRID || problem > GCR_PROBLEM_SCALE_ONLY_ORIGINAL) return h->fail(GCR_ERR_INVALID, "unknown problem id");
    if (problem != h->problem)
    {
        h->problem = problem;
        h->ns = h->no = 0; // features must be uploaded again for the new problem
        h->have_run = false;
    }
    return GCR_OK;
}

int gcr_peer_export(gcr_handle h, void* ipc_handle_64)
{
    if (!h || !ipc_handle_64) return GCR_ERR_INVALID;
    static_assert(sizeof(cudaIpcMemHandle_t) == GCR_IPC_HANDLE_BYTES, "ipc handle size");
    int rc = peer_buffer(h);
    if (rc != GCR_OK) return rc;
    cudaIpcMemHandle_t mh;
    GCR_CUDA(cudaIpcGetMemHandle(&mh, h->d_xchg.p));
    std::memcpy(ipc_handle_64, &mh, sizeof(mh));
    return GCR_OK;
}

int gcr_peer_buffer(gcr_handle h, void** device_ptr)
{
    if (!h || !device_ptr) return GCR_ERR_INVALID;
    int rc = peer_buffer(h);
    if (rc != GCR_OK) return rc;
    *device_ptr = h->d_xchg.p;
    return GCR_OK;
}

int gcr_peer_attach_ptrs(gcr_handle h, void* const* peer_buffers, int world, int rank)
{
    if (!h || !peer_buffers) return GCR_ERR_INVALID;
    if (world < 2 || world > kMaxPeers || rank < 0 || rank >= world) return h->fail(GCR_ERR_INVALID, "2 <= world <= 16, 0 <= rank < world");
    int rc = peer_buffer(h);
    if (rc != GCR_OK) return rc;
    peer_release(h);
    GCR_DEVICE(h);
    GCR_CUDA(cudaStreamSynchronize(h->stream));
    GCR_CUDA(cudaMemset(h->d_xchg.p, 0, sizeof(unsigned long long) * kXchgWords));
    for (int p = 0; p < world; p++) h->peer_ptr[p] = p == rank ? static_cast<void*>(h->d_xchg.p) : peer_buffers[p];
    h->xchg_world = world;
    h->xchg_rank = rank;
    return GCR_OK;
}

int gcr_peer_attach(gcr_handle h, const void* ipc_handles, int world, int rank)
{
    if (!h || !ipc_handles) return GCR_ERR_INVALID;
    if (world < 2 || world > kMaxPeers || rank < 0 || rank >= world) return h->fail(GCR_ERR_INVALID, "2 <= world <= 16, 0 <= rank < world");
    int rc = peer_buffer(h);
    if (rc != GCR_OK) return rc;
    peer_release(h);
    GCR_DEVICE(h);
    GCR_CUDA(cudaStreamSynchronize(h->stream));
    GCR_CUDA(cudaMemset(h->d_xchg.p, 0, sizeof(unsigned long long) * kXchgWords));
    for (int p = 0; p < world; p++)
    {
        if (p == rank)
        {
            h->peer_ptr[p] = h->d_xchg.p;
            continue;
        }
        cudaIpcMemHandle_t mh;
        std::memcpy(&mh, static_cast<const char*>(ipc_handles) + static_cast<size_t>(p) * sizeof(mh), sizeof(mh));
        void* ptr = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&ptr, mh, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess)
        {
            peer_release(h);
            return h->cuda_fail(e, "cudaIpcOpenMemHandle (peer exchange buffer)");
        }
        h->peer_ptr[p] = ptr;
        h->peer_ipc[p] = true;
    }
    h->xchg_world = world;
    h->xchg_rank = rank;
    return GCR_OK;
}

int gcr_peer_detach(gcr_handle h)
{
    if (!h) return GCR_ERR_INVALID;
    DeviceGuard device_guard_(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    return peer_release(h);
}

int gcr_get_features(gcr_handle h, double* scale, double* orient)
{
    if (!h) return GCR_ERR_INVALID;
    if (!h->fetch_scale) return h->fail(GCR_ERR_STATE, "gcr_get_features before gcr_set_features / gcr_set_keypoints");
    GCR_DEVICE(h);
    if (scale && h->fetch_ns > 0)
        GCR_CUDA(cudaMemcpyAsync(scale, h->fetch_scale, sizeof(double) * 3 * h->fetch_ns, cudaMemcpyDeviceToHost, h->stream));
    if (orient && h->fetch_no > 0 && h->fetch_orient)
        GCR_CUDA(cudaMemcpyAsync(orient, h->fetch_orient, sizeof(double) * 3 * h->fetch_no, cudaMemcpyDeviceToHost, h->stream));
    GCR_CUDA(cudaStreamSynchronize(h->stream));
    return GCR_OK;
}

int gcr_run(gcr_handle h, int64_t n_hyp, uint64_t seed, uint64_t first_hyp, const int32_t* samples,
            const double thr[2], uint32_t flags, void* d_record, gcr_result* out)
{
    if (!h) return GCR_ERR_INVALID;
    if (h->ns < 2) return h->fail(GCR_ERR_STATE, "gcr_run before gcr_set_features");
    if ((h->problem == 0) != (h->no >= 2)) return h->fail(GCR_ERR_STATE, "gcr_set_problem changed after gcr_set_features: upload the features again");
    if (n_hyp < 1 || n_hyp > (1ll << 30)) return h->fail(GCR_ERR_INVALID, "n_hyp out of range");
    if (!thr || !(thr[0] >= 0.0) || !std::isfinite(thr[0]) || (h->problem == 0 && (!(thr[1] >= 0.0) || !std::isfinite(thr[1]))))
        return h->fail(GCR_ERR_INVALID, "thresholds must be finite and non-negative");
    GCR_DEVICE(h);
    const size_t H = static_cast<size_t>(n_hyp);
    const size_t W = static_cast<size_t>((h->ns + 31) / 32 + (h->no + 31) / 32);
    h->n_hyp = static_cast<int>(n_hyp);
    h->first_hyp = first_hyp;
    h->have_masks = (flags & GCR_RUN_EMIT_MASKS) != 0;
    GCR_CUDA(h->d_samples.ensure(4 * H));
    GCR_CUDA(h->d_models.ensure(H));
    GCR_CUDA(h->d_flags.ensure(H));
    GCR_CUDA(h->d_sums.ensure(2 * H));
    GCR_CUDA(h->d_cnts.ensure(2 * H));
    GCR_CUDA(h->d_score.ensure(H));
    GCR_CUDA(h->d_ninl.ensure(2 * H));
    GCR_CUDA(h->d_fast_hyp.ensure(H));
    GCR_CUDA(h->d_hard_hyp.ensure(H));
    GCR_CUDA(h->d_pf0.ensure(H));
    GCR_CUDA(h->d_pf1.ensure(H));
    GCR_CUDA(h->d_err_ab.ensure(H));
    GCR_CUDA(h->d_worklist.ensure(std::max<size_t>(1u << 16, 4 * H)));
    if (h->have_masks) GCR_CUDA(h->d_masks.ensure(H * W));
    int rc = ensure_common(h);
    if (rc != GCR_OK) return rc;

    cudaStream_t s = h->stream;
    h->samples_view = h->d_samples.p;
    h->samples_u16 = samples != nullptr && (flags & GCR_RUN_SAMPLES_U16) != 0;
    if (h->samples_u16 && (h->ns > 65535 || h->no > 65535))
        return h->fail(GCR_ERR_INVALID, "GCR_RUN_SAMPLES_U16: a class has more than 65 535 features");
    const size_t sample_bytes = (h->samples_u16 ? sizeof(uint16_t) : sizeof(int)) * 4 * H;
    if (samples)
    {
        if (flags & GCR_RUN_SAMPLES_ON_DEVICE)
            h->samples_view = samples; // read in place: the caller keeps the array alive until the run has finished
        else
        {
            // Page-locked, device-mapped host memory is read IN PLACE by the sampler/solver kernel (one coalesced 16-byte read
            // per hypothesis): the PCIe transfer then overlaps that kernel instead of preceding it.  As with any asynchronous
            // copy from pinned memory, the caller must leave the array alone until the run has finished.
            const int* mapped = nullptr;
            cudaPointerAttributes attr;
            static const bool no_zero_copy = std::getenv("GCR_NO_ZEROCOPY") != nullptr;
            if (!no_zero_copy && cudaPointerGetAttributes(&attr, samples) == cudaSuccess && attr.type == cudaMemoryTypeHost &&
                attr.devicePointer != nullptr && reinterpret_cast<uintptr_t>(attr.devicePointer) % (h->samples_u16 ? 8 : 16) == 0)
                mapped = static_cast<const int*>(attr.devicePointer);
            else
                cudaGetLastError(); // pageable memory: cudaPointerGetAttributes may leave an error state on old drivers
            if (mapped)
                h->samples_view = mapped;
            else
                GCR_CUDA(cudaMemcpyAsync(h->d_samples.p, samples, sample_bytes, cudaMemcpyHostToDevice, s));
        }
    }
    const Thresholds T = make_thresholds(thr, h->problem);
    Buffers B = make_buffers(h, seed, samples != nullptr, d_record);
    {
        // the tail writes the result block straight into the pinned staging buffer (device-mapped): no copy after the run
        static const bool no_mirror = std::getenv("GCR_NO_ZEROCOPY") != nullptr;
        void* dp = nullptr;
        // (only for synchronous runs, out != NULL: an asynchronous caller that reads nothing back should not pay the ~2 us the
        //  host writes add to the tail)
        h->last_run_mirrored = out != nullptr && !no_mirror && cudaHostGetDevicePointer(&dp, h->h_best, 0) == cudaSuccess && dp != nullptr;
        if (h->last_run_mirrored) B.host_block = static_cast<unsigned int*>(dp);
        else cudaGetLastError();
    }
    h->have_per_hyp = (flags & (GCR_RUN_KEEP_PER_HYP | GCR_RUN_EMIT_MASKS)) != 0;
    B.per_hyp = h->have_per_hyp ? 1 : 0;
    const int all_exact = (flags & GCR_RUN_ALL_EXACT) ? 1 : 0;
    h->last_thr[0] = thr[0];
    h->last_thr[1] = h->problem == 0 ? thr[1] : 0.0;
    h->last_run_exchanged = false;
    if (flags & GCR_RUN_EXCHANGE)
    {
        if (h->xchg_world < 2) return h->fail(GCR_ERR_STATE, "GCR_RUN_EXCHANGE without attached peers (gcr_peer_attach)");
        B.xchg_world = h->xchg_world;
        B.xchg_seq = ++h->xchg_seq;
        h->last_run_exchanged = true;
    }

    // launch sequence: sample+solve -> FP32 scoring -> fused cooperative tail (3 kernels; the run counters were
    // cleared by the previous run's tail)
    const bool time_kernels = out != nullptr && (flags & GCR_RUN_TIME_KERNELS) != 0; // opt-in: every event record is an in-stream operation
    if (time_kernels) GCR_CUDA(cudaEventRecord(h->ev0, s));
    if (h->have_masks) GCR_CUDA(cudaMemsetAsync(h->d_masks.p, 0, sizeof(unsigned int) * H * W, s));
    const int hb = static_cast<int>((H + 127) / 128);
    {
        // chained to the kernel before it: behind k_prep_features (gcr_set_features + gcr_run back to back) the sample fetch, over
        // PCIe for pinned replay arrays, runs under the feature preparation; behind any other kernel it is a plain launch
        cudaLaunchAttribute at;
        at.id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at.val.programmaticStreamSerializationAllowed = 1;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(hb);
        cfg.blockDim = dim3(128);
        cfg.stream = s;
        cfg.attrs = &at;
        cfg.numAttrs = h->pdl && !time_kernels && !h->have_masks ? 1 : 0;
        GCR_CUDA(cudaLaunchKernelEx(&cfg, k_sample_solve, B, T, all_exact));
    }
    {
        const int mode = h->have_masks ? 2 : (h->have_per_hyp ? 1 : 0); // scoring mode: winner-only | per-hypothesis counts | + masks
        const int grid = h->score_grid[mode]; // persistent: one CTA per resident slot
        // small feature sets (both working sets fit in one SM's shared memory): resident kernel, one CTA per SM
        const int chunks = B.ncs + B.nco + B.ncsd + B.ncod;
        const bool resident = h->last_resident = h->resident_ok && chunks <= kResidentMaxChunks && n_hyp <= h->resident_max_hyp &&
                                                   n_hyp * (static_cast<long long>(B.ns) + B.no) <= h->resident_max_evals;
        const size_t rsmem = static_cast<size_t>(chunks) * kChunkBytes + 64;
        // chained launches: the scoring kernel and the tail start launching while their predecessor runs and wait for its
        // results on the device (griddepcontrol.wait); not when the scoring kernel is bracketed by timing events
        const bool pdl = h->pdl && !time_kernels;
        cudaLaunchAttribute at[2];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        at[1].id = cudaLaunchAttributeCooperative;
        at[1].val.cooperative = 1;
        cudaLaunchConfig_t cfg = {};
        cfg.stream = s;
        cfg.attrs = at;
        cfg.numAttrs = pdl ? 1 : 0;
        if (time_kernels) GCR_CUDA(cudaEventRecord(h->evk0, s));
        if (resident)
        {
            cfg.gridDim = dim3(h->n_sm);
            cfg.blockDim = dim3(mode == 0 ? kResThreads<0> : kResidentThreads);
            cfg.dynamicSmemBytes = rsmem;
            GCR_CUDA(cudaLaunchKernelEx(&cfg, mode == 2 ? k_score_resident<2> : (mode == 1 ? k_score_resident<1> : k_score_resident<0>), B, T,
                                        h->resident_cost_o, h->resident_cost_p));
        }
        else
        {
            cfg.gridDim = dim3(grid);
            cfg.blockDim = dim3(kScoreThreads);
            cfg.dynamicSmemBytes = sizeof(ScoreSmem);
            GCR_CUDA(cudaLaunchKernelEx(&cfg, mode == 2 ? k_score_fast<2> : (mode == 1 ? k_score_fast<1> : k_score_fast<0>), B, T));
        }
        if (time_kernels) GCR_CUDA(cudaEventRecord(h->evk1, s));
        cfg.gridDim = dim3(h->tail_grid);
        cfg.blockDim = dim3(kExactThreads);
        cfg.dynamicSmemBytes = 0;
        if (pdl)
            cfg.numAttrs = 2; // cooperative + programmatic
        else
        {
            cfg.attrs = at + 1; // cooperative only
            cfg.numAttrs = 1;
        }
        GCR_CUDA(cudaLaunchKernelEx(&cfg, k_tail, B, T));
    }
    if (time_kernels) GCR_CUDA(cudaEventRecord(h->ev1, s));
    if (flags & (GCR_RUN_KEEP_PER_HYP | GCR_RUN_EMIT_MASKS))
    {
        // reporting modes: FP64 re-scoring of the low-margin hypotheses (see k_refine_collect); off the timed stage
        // tighter-than-default thresholds amplify the FP32 path's relative score error past the 1e-5 reporting bar for
        // every hypothesis, not only the low-margin ones: re-score them all
        const int refine_all = (T.Tp < 0.7f * 0.096f || (h->problem == 0 && std::sqrt(T.tsq[1]) < 0.7 * 0.0785)) ? 1 : 0;
        k_refine_collect<<<hb, 128, 0, s>>>(B, refine_all);
        k_exact_items<<<h->n_sm * 8, kExactThreads, 0, s>>>(B, T, 0);
        k_refine_finalize<<<h->n_sm, 256, 0, s>>>(B, T);
        k_refine_done<<<1, 1, 0, s>>>(B);
    }
    GCR_CUDA(cudaGetLastError());
    h->have_run = true;

    if (out)
    {
        Counters c;
        Record r;
        if (!h->last_run_mirrored) GCR_CUDA(cudaMemcpyAsync(h->h_best, h->d_best.p, 4 * kResultHeaderWords, cudaMemcpyDeviceToHost, s));
        GCR_CUDA(cudaStreamSynchronize(s));
        std::memcpy(&r, h->h_best, sizeof(r));
        std::memcpy(&c, h->h_best + 16, sizeof(c));
        unsigned long long ph[8];
        std::memcpy(ph, h->h_best + 32, sizeof(ph));
        std::memset(out, 0, sizeof(*out));
        out->best_index = r.index;
        out->best_score = r.score;
        for (int i = 0; i < 4; i++) out->best_model[i] = r.model[i];
        out->best_n_inl[0] = r.n_inl[0];
        out->best_n_inl[1] = r.n_inl[1];
        out->n_hyp = n_hyp;
        out->n_sample_valid = c.n_sample_valid;
        out->n_solved = c.n_solved;
        out->n_rechecked = c.n_worklist;
        out->n_exact_hyp = c.n_hard;
        out->n_refined = c.n_refined;
        out->n_scored = static_cast<int64_t>(c.n_fast) + c.n_sens + c.n_hard;
        out->n_candidates = c.n_candidates;
        if (time_kernels)
        {
            GCR_CUDA(cudaEventElapsedTime(&out->device_ms, h->ev0, h->ev1));
            GCR_CUDA(cudaEventElapsedTime(&out->score_kernel_ms, h->evk0, h->evk1));
        }
        for (int i = 0; i < 5; i++) out->tail_phase_us[i] = static_cast<float>(static_cast<double>(ph[i + 1] - ph[i]) * 1e-3);
        out->tail_phase_us[5] = static_cast<float>(static_cast<double>(ph[0] - ph[7]) * 1e-3); // k_score_fast start -> k_tail start
        out->stage_phase_us[0] = static_cast<float>(static_cast<double>(ph[7] - ph[6]) * 1e-3); // k_sample_solve start -> k_score_fast start
        out->stage_phase_us[1] = static_cast<float>(static_cast<double>(ph[5] - ph[6]) * 1e-3); // k_sample_solve start -> end of k_tail
        out->winner_rank = static_cast<int32_t>(r.reserved);
        out->score_kernel = h->last_resident ? 1 : 0;
        return check_run_counters(h, c);
    }
    return GCR_OK;
}

int gcr_best(gcr_handle h, double model[4], double* score, uint32_t n_inl[2], uint32_t* scale_mask_words,
             uint32_t* orient_mask_words, int64_t* hyp_index)
{
    if (!h) return GCR_ERR_INVALID;
    if (!h->have_run) return h->fail(GCR_ERR_STATE, "gcr_best before gcr_run");
    const size_t ws = (h->ns + 31) / 32, wo = (h->no + 31) / 32;
    const bool want_masks = scale_mask_words || orient_mask_words;
    if (!h->last_run_mirrored)
    {
        GCR_DEVICE(h);
        GCR_CUDA(cudaMemcpyAsync(h->h_best, h->d_best.p, sizeof(unsigned int) * (want_masks ? kResultHeaderWords + ws + wo : 32),
                                 cudaMemcpyDeviceToHost, h->stream));
        GCR_CUDA(cudaStreamSynchronize(h->stream));
    }
    // (a mirrored run was synchronous: the staging block is complete and the stream idle)
    Record r;
    std::memcpy(&r, h->h_best, sizeof(r));
    {
        Counters c;
        std::memcpy(&c, h->h_best + 16, sizeof(c));
        const int rc = check_run_counters(h, c);
        if (rc != GCR_OK) return rc;
    }
    if (model) for (int i = 0; i < 4; i++) model[i] = r.model[i];
    if (score) *score = r.score;
    if (n_inl) { n_inl[0] = r.n_inl[0]; n_inl[1] = r.n_inl[1]; }
    if (hyp_index) *hyp_index = r.index;
    if (want_masks && h->last_run_exchanged && r.index >= 0 && static_cast<int>(r.reserved) != h->xchg_rank)
    {
        // the winning hypothesis lives on another rank: its masks are recomputed here from the replicated features
        if (h->last_thr[0] < 0.0) return h->fail(GCR_ERR_STATE, "no thresholds recorded");
        std::vector<uint32_t> words(ws + wo);
        const int rc = gcr_score_models(h, r.model, 1, h->last_thr, 0, nullptr, nullptr, words.data());
        if (rc != GCR_OK) return rc;
        if (scale_mask_words) std::memcpy(scale_mask_words, words.data(), 4 * ws);
        if (orient_mask_words) std::memcpy(orient_mask_words, words.data() + ws, 4 * wo);
        return GCR_OK;
    }
    if (scale_mask_words) std::memcpy(scale_mask_words, h->h_best + kResultHeaderWords, 4 * ws);
    if (orient_mask_words) std::memcpy(orient_mask_words, h->h_best + kResultHeaderWords + ws, 4 * wo);
    return GCR_OK;
}

int gcr_fetch(gcr_handle h, uint8_t* flags, double* models, double* score, uint32_t* n_inl, uint32_t* masks,
              int32_t* samples)
{
    if (!h) return GCR_ERR_INVALID;
    if (!h->have_run) return h->fail(GCR_ERR_STATE, "gcr_fetch before gcr_run");
    GCR_DEVICE(h);
    if (!h->last_run_mirrored) GCR_CUDA(cudaMemcpyAsync(h->h_best, h->d_best.p, sizeof(unsigned int) * 32, cudaMemcpyDeviceToHost, h->stream));
    GCR_CUDA(cudaStreamSynchronize(h->stream));
    {
        Counters c;
        std::memcpy(&c, h->h_best + 16, sizeof(c));
        const int rc = check_run_counters(h, c);
        if (rc != GCR_OK) return rc;
    }
    const size_t H = static_cast<size_t>(h->n_hyp);
    const size_t W = static_cast<size_t>((h->ns + 31) / 32 + (h->no + 31) / 32);
    if ((score || n_inl) && !h->have_per_hyp)
        return h->fail(GCR_ERR_STATE, "per-hypothesis scores were not kept: run with GCR_RUN_KEEP_PER_HYP or GCR_RUN_EMIT_MASKS");
    if (flags) GCR_CUDA(cudaMemcpy(flags, h->d_flags.p, H, cudaMemcpyDeviceToHost));
    if (models) GCR_CUDA(cudaMemcpy(models, h->d_models.p, H * sizeof(Model4), cudaMemcpyDeviceToHost));
    if (score) GCR_CUDA(cudaMemcpy(score, h->d_score.p, H * sizeof(double), cudaMemcpyDeviceToHost));
    if (n_inl) GCR_CUDA(cudaMemcpy(n_inl, h->d_ninl.p, 2 * H * sizeof(unsigned int), cudaMemcpyDeviceToHost));
    if (samples)
    {
        if (h->samples_u16) // (replayed from a uint16 array: widened here, 0xFFFF = the unused index -1)
        {
            std::vector<uint16_t> narrow(4 * H);
            GCR_CUDA(cudaMemcpy(narrow.data(), h->samples_view, 4 * H * sizeof(uint16_t), cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < 4 * H; ++i) samples[i] = (i % 4 == 3 && narrow[i] == 0xffffu) ? -1 : static_cast<int>(narrow[i]);
        }
        else
            GCR_CUDA(cudaMemcpy(samples, h->samples_view, 4 * H * sizeof(int), cudaMemcpyDeviceToHost));
    }
    if (masks)
    {
        if (!h->have_masks) return h->fail(GCR_ERR_STATE, "masks were not kept: run with GCR_RUN_EMIT_MASKS");
        GCR_CUDA(cudaMemcpy(masks, h->d_masks.p, H * W * sizeof(unsigned int), cudaMemcpyDeviceToHost));
    }
    return GCR_OK;
}

int gcr_score_models(gcr_handle h, const double* models, int64_t n, const double thr[2], uint32_t flags, double* score,
                     uint32_t* n_inl, uint32_t* mask_words)
{
    if (!h || !models || !thr) return GCR_ERR_INVALID;
    if (h->ns < 2) return h->fail(GCR_ERR_STATE, "gcr_score_models before gcr_set_features");
    if (n < 1 || n > kCandCap) return h->fail(GCR_ERR_INVALID, "gcr_score_models: 1 <= n <= 1024 models per call");
    if (!(thr[0] >= 0.0) || !std::isfinite(thr[0]) || (h->problem == 0 && (!(thr[1] >= 0.0) || !std::isfinite(thr[1]))))
        return h->fail(GCR_ERR_INVALID, "thresholds must be finite and non-negative");
    GCR_DEVICE(h);
    int rc = ensure_common(h);
    if (rc != GCR_OK) return rc;
    GCR_CUDA(h->d_user_models.ensure(kCandCap));
    Thresholds T = make_thresholds(thr, h->problem);
    if (flags & GCR_SCORE_TRUNCATED_SQ) // the relabelling threshold of the local optimisation, GCRANSAC.h:207-208,931-937
    {
        T.tsq[0] = (1.5 * thr[0]) * (1.5 * thr[0]);
        T.tsq[1] = (1.5 * thr[1]) * (1.5 * thr[1]);
    }
    T.strict = (flags & GCR_SCORE_STRICT) ? 1 : 0;
    const Buffers B = make_buffers(h, 0, false, nullptr);
    cudaStream_t s = h->stream;
    const unsigned int un = static_cast<unsigned int>(n);
    const size_t W = static_cast<size_t>((h->ns + 31) / 32 + (h->no + 31) / 32);
    const unsigned long long units =
        static_cast<unsigned long long>(un) * ((h->ns + kExactThreads - 1) / kExactThreads + (h->no + kExactThreads - 1) / kExactThreads);
    const int grid = static_cast<int>(std::min<unsigned long long>(units, static_cast<unsigned long long>(h->n_sm) * 8));
    // models in and results out through ONE page-locked staging block that the kernels read / write in place
    auto align16 = [](size_t x) { return (x + 15) & ~static_cast<size_t>(15); };
    const size_t off_score = align16(sizeof(Model4) * n), off_cnt = align16(off_score + 8 * n), off_masks = align16(off_cnt + 8 * n),
                 total = off_masks + (mask_words ? 4 * W * n : 0);
    GCR_CUDA(h->h_score_io.ensure(total));
    void* dp = nullptr;
    static const bool no_zero_copy = std::getenv("GCR_NO_ZEROCOPY") != nullptr;
    if (!no_zero_copy && cudaHostGetDevicePointer(&dp, h->h_score_io.p, 0) == cudaSuccess && dp)
    {
        unsigned char* d = static_cast<unsigned char*>(dp);
        std::memcpy(h->h_score_io.p, models, sizeof(Model4) * n);
        k_set_models<<<(un + 127) / 128, 128, 0, s>>>(B, reinterpret_cast<const Model4*>(d), un);
        GCR_CUDA(launch_chained(h, k_exact_items, dim3(grid), dim3(kExactThreads), s, B, T, 2, un));
        GCR_CUDA(launch_chained(h, k_score_out, dim3(std::max(1, std::min<int>(h->n_sm, static_cast<int>((W * n + 255) / 256)))), dim3(256), s, B, T,
                                un, reinterpret_cast<double*>(d + off_score), reinterpret_cast<unsigned int*>(d + off_cnt),
                                mask_words ? reinterpret_cast<unsigned int*>(d + off_masks) : nullptr));
        GCR_CUDA(cudaGetLastError());
        GCR_CUDA(cudaStreamSynchronize(s));
        if (score) std::memcpy(score, h->h_score_io.p + off_score, sizeof(double) * n);
        if (n_inl) std::memcpy(n_inl, h->h_score_io.p + off_cnt, 2 * sizeof(unsigned int) * n);
        if (mask_words) std::memcpy(mask_words, h->h_score_io.p + off_masks, 4 * W * n);
        return GCR_OK;
    }
    cudaGetLastError();
    GCR_CUDA(cudaMemcpyAsync(h->d_user_models.p, models, sizeof(Model4) * n, cudaMemcpyHostToDevice, s));
    k_set_models<<<(un + 127) / 128, 128, 0, s>>>(B, h->d_user_models.p, un);
    GCR_CUDA(launch_chained(h, k_exact_items, dim3(grid), dim3(kExactThreads), s, B, T, 2, un));
    GCR_CUDA(launch_chained(h, k_cand_value, dim3((un + 127) / 128), dim3(128), s, B, T, un));
    GCR_CUDA(cudaGetLastError());
    if (score) GCR_CUDA(cudaMemcpyAsync(score, h->d_cand_score.p, sizeof(double) * n, cudaMemcpyDeviceToHost, s));
    if (n_inl) GCR_CUDA(cudaMemcpyAsync(n_inl, h->d_cand_cnt.p, 2 * sizeof(unsigned int) * n, cudaMemcpyDeviceToHost, s));
    if (mask_words) GCR_CUDA(cudaMemcpyAsync(mask_words, h->d_cand_masks.p, 4 * W * n, cudaMemcpyDeviceToHost, s));
    GCR_CUDA(cudaStreamSynchronize(s));
    return GCR_OK;
}

// findWeightedMode (two_sift.hpp:354-394) on the histogram entries a fit returned (bins in order of first appearance): the heaviest
// bin, the first strictly heavier one in the iteration order of an unordered_map<int, double> filled in input order; with unit weights
// the weight order is the count order.  The REAL map of this host: the check of the device's restatement of its order.
static double replay_phi(const FitEntry* e, unsigned int n)
{
    std::unordered_map<int, double> bin_weight;
    for (unsigned int i = 0; i < n; ++i) bin_weight[e[i].bin] += static_cast<double>(e[i].count);
    int mode_bin = 0;
    double heaviest = -1.0;
    for (const auto& kv : bin_weight)
        if (kv.second > heaviest)
        {
            heaviest = kv.second;
            mode_bin = kv.first;
        }
    double phi = 0.0;
    for (unsigned int i = 0; i < n; ++i)
        if (e[i].bin == mode_bin) phi = static_cast<double>(e[i].moment) / kPhiFixedPoint / static_cast<double>(e[i].count);
    return phi;
}

// bucket-count schedule of this host's std::unordered_map<int, double> while it grows to kPhiBins keys (fit_phi_block)
static void fill_rehash_schedule(FitBuffers& F)
{
    struct Sched { int n; int at[8], to[8]; };
    static const Sched sched = [] {
        Sched sc{0, {}, {}};
        std::unordered_map<int, double> probe;
        size_t last = probe.bucket_count();
        for (int i = 0; i < kPhiBins; ++i)
        {
            probe[i] += 1.0;
            if (probe.bucket_count() != last && sc.n < 8)
            {
                last = probe.bucket_count();
                sc.at[sc.n] = i + 1;
                sc.to[sc.n++] = static_cast<int>(last);
            }
        }
        return sc;
    }();
    F.n_rehash = sched.n;
    for (int i = 0; i < 8; ++i)
    {
        F.rehash_at[i] = sched.at[i];
        F.rehash_to[i] = sched.to[i];
    }
}

// gcr_fit_models, and with `thr` the scoring of the fitted models in the same launch sequence (gcr_fit_score_models)
static int fit_models_impl(gcr_handle h, int64_t n_fits, const uint32_t* scale_idx, const int64_t* scale_off, const uint32_t* orient_idx,
                           const int64_t* orient_off, double* models, int32_t* ok, const double* thr, uint32_t score_flags, double* score,
                           uint32_t* n_inl, uint32_t* mask_words)
{
    if (!h || !scale_idx || !scale_off || !models || !ok) return GCR_ERR_INVALID;
    const bool do_score = thr != nullptr;
    if (do_score && (!(thr[0] >= 0.0) || !std::isfinite(thr[0]) || (h->problem == 0 && (!(thr[1] >= 0.0) || !std::isfinite(thr[1])))))
        return h->fail(GCR_ERR_INVALID, "thresholds must be finite and non-negative");
    if (h->ns < 2) return h->fail(GCR_ERR_STATE, "gcr_fit_models before gcr_set_features");
    const bool hybrid = h->problem == 0;
    if (hybrid && (!orient_idx || !orient_off)) return h->fail(GCR_ERR_INVALID, "gcr_fit_models: the hybrid problem needs orientation index lists");
    if (n_fits < 1 || n_fits > kCandCap) return h->fail(GCR_ERR_INVALID, "gcr_fit_models: 1 <= n_fits <= 1024 per call");
    const size_t K = static_cast<size_t>(n_fits);
    if (scale_off[0] != 0 || (hybrid && orient_off[0] != 0)) return h->fail(GCR_ERR_INVALID, "gcr_fit_models: offsets must start at 0");
    for (size_t k = 0; k < K; ++k)
        if (scale_off[k + 1] < scale_off[k] || (hybrid && orient_off[k + 1] < orient_off[k]) ||
            scale_off[k + 1] - scale_off[k] > h->ns || (hybrid && orient_off[k + 1] - orient_off[k] > h->no))
            return h->fail(GCR_ERR_INVALID, "gcr_fit_models: offsets must be non-decreasing, a list no longer than its class");
    const size_t S = static_cast<size_t>(scale_off[K]), O = hybrid ? static_cast<size_t>(orient_off[K]) : 0;
    for (size_t i = 0; i < S; ++i)
        if (scale_idx[i] >= static_cast<uint32_t>(h->ns)) return h->fail(GCR_ERR_INVALID, "gcr_fit_models: scale index out of range");
    for (size_t i = 0; i < O; ++i)
        if (orient_idx[i] >= static_cast<uint32_t>(h->no)) return h->fail(GCR_ERR_INVALID, "gcr_fit_models: orientation index out of range");
    GCR_DEVICE(h);

    auto align16 = [](size_t x) { return (x + 15) & ~static_cast<size_t>(15); };
    const size_t in_desc = 0, in_s = align16(K * sizeof(FitDesc)), in_o = align16(in_s + 4 * S), in_bytes = align16(in_o + 4 * O);
    GCR_CUDA(h->h_fit_in.ensure(in_bytes));
    GCR_CUDA(h->d_fit_in.ensure(in_bytes));
    FitDesc* desc = reinterpret_cast<FitDesc*>(h->h_fit_in.p + in_desc);
    unsigned long long units = 0, entries = 0;
    int longest_o = 0;
    for (size_t k = 0; k < K; ++k)
    {
        FitDesc& D = desc[k];
        D.s_off = scale_off[k];
        D.ns = static_cast<int>(scale_off[k + 1] - scale_off[k]);
        D.o_off = hybrid ? orient_off[k] : 0;
        D.no = hybrid ? static_cast<int>(orient_off[k + 1] - orient_off[k]) : 0;
        // two_sift.hpp:431-447,603-609 / three_sift.hpp:203-211: fewer constraints than a minimal sample -> no model; exactly a
        // minimal sample belongs to the minimal solver, which neither the local optimisation nor the refit presents
        D.valid = hybrid ? ((D.ns >= 2 && D.no >= 2 && !(D.ns == 2 && D.no == 2)) ? 1 : 0) : (D.ns > 3 ? 1 : 0);
        longest_o = std::max(longest_o, D.no);
        const unsigned long long rows =
            static_cast<unsigned long long>(D.ns) + (hybrid ? static_cast<unsigned long long>(D.no) * (D.no > 0 ? D.no - 1 : 0) / 2 : 0);
        D.unit_first = static_cast<unsigned int>(units);
        D.ent_off = static_cast<unsigned int>(entries);
        D.unit_rows = 1024;
        while (D.unit_rows < kFitUnitRows && (rows + D.unit_rows - 1) / D.unit_rows > static_cast<unsigned long long>(kFitUnitsAim)) D.unit_rows *= 2;
        units += std::max<unsigned long long>(1, (rows + D.unit_rows - 1) / D.unit_rows);
        entries += static_cast<unsigned long long>(std::min(D.no, kPhiBins));
        if (units > 0x7fffffffull) return h->fail(GCR_ERR_INVALID, "gcr_fit_models: too many rows");
    }
    std::memcpy(h->h_fit_in.p + in_s, scale_idx, 4 * S);
    if (O) std::memcpy(h->h_fit_in.p + in_o, orient_idx, 4 * O);
    const size_t Wm = static_cast<size_t>((h->ns + 31) / 32 + (h->no + 31) / 32);
    const size_t out_models = 0, out_ok = align16(K * 32), out_nent = align16(out_ok + 4 * K), out_ent = align16(out_nent + 4 * K),
                 out_score = align16(out_ent + entries * sizeof(FitEntry)), out_cnt = align16(out_score + (do_score ? 8 * K : 0)),
                 out_masks = align16(out_cnt + (do_score ? 8 * K : 0)),
                 out_bytes = align16(out_masks + (do_score && mask_words ? 4 * Wm * K : 0));
    if (do_score)
    {
        const int rc = ensure_common(h);
        if (rc != GCR_OK) return rc;
    }
    GCR_CUDA(h->h_fit_out.ensure(out_bytes));
    GCR_CUDA(h->d_fit_out.ensure(out_bytes));
    const bool lines_in_smem = longest_o <= kFitSmemLines;
    if (!lines_in_smem) GCR_CUDA(h->d_fit_lines.ensure(3 * std::max<size_t>(O, 1)));
    if (!h->d_fit_done.p)
    {
        GCR_CUDA(h->d_fit_done.ensure(kCandCap));
        GCR_CUDA(cudaMemsetAsync(h->d_fit_done.p, 0, sizeof(unsigned int) * kCandCap, h->stream));
    }
    GCR_CUDA(h->d_fit_partial.ensure(static_cast<size_t>(units) * 2 * kFitGram));

    cudaStream_t s = h->stream;
    // Small calls (a batch of local-optimisation fits, the refit) are chains of microsecond kernels: their results are written into the
    // page-locked output block by the kernels that produce them (posted writes), larger ones go through the copy engine.  The INPUT
    // always does: descriptors and index lists read in place over PCIe put a ~1.5 us round trip into every dependent access of the
    // row kernel (binary search over the descriptors, index, feature) -- measured 5-10 us slower per call than the one copy.
    constexpr size_t kFitInPlaceBytes = size_t(256) << 10;
    const bool out_place = out_bytes <= kFitInPlaceBytes;
    const unsigned char* in_base = h->d_fit_in.p;
    GCR_CUDA(cudaMemcpyAsync(h->d_fit_in.p, h->h_fit_in.p, in_bytes, cudaMemcpyHostToDevice, s));
    FitBuffers F;
    F.scale = h->scale_ptr;
    F.orient = h->orient_ptr;
    F.desc = reinterpret_cast<const FitDesc*>(in_base + in_desc);
    F.idx_s = reinterpret_cast<const unsigned int*>(in_base + in_s);
    F.idx_o = reinterpret_cast<const unsigned int*>(in_base + in_o);
    F.lines = lines_in_smem ? nullptr : h->d_fit_lines.p;
    F.done = h->d_fit_done.p;
    F.partial = h->d_fit_partial.p;
    F.models = reinterpret_cast<double*>(h->d_fit_out.p + out_models);
    F.ok = reinterpret_cast<int*>(h->d_fit_out.p + out_ok);
    F.n_ent = reinterpret_cast<unsigned int*>(h->d_fit_out.p + out_nent);
    F.ent = reinterpret_cast<FitEntry*>(h->d_fit_out.p + out_ent);
    F.h_models = out_place ? reinterpret_cast<double*>(h->h_fit_out.p + out_models) : nullptr;
    F.h_ok = out_place ? reinterpret_cast<int*>(h->h_fit_out.p + out_ok) : nullptr;
    F.h_n_ent = out_place ? reinterpret_cast<unsigned int*>(h->h_fit_out.p + out_nent) : nullptr;
    F.h_ent = out_place ? reinterpret_cast<FitEntry*>(h->h_fit_out.p + out_ent) : nullptr;
    F.n_fits = static_cast<int>(K);
    F.n_units = static_cast<unsigned int>(units);
    F.n_lines = static_cast<long long>(O);
    F.problem = h->problem;
    fill_rehash_schedule(F);
    FitSlots slots{nullptr, nullptr, nullptr, nullptr};
    if (do_score)
    {
        const Buffers B0 = make_buffers(h, 0, false, nullptr);
        slots = FitSlots{B0.cand_model, B0.cand_hyp, B0.cand_qsum, B0.cand_rawcnt};
    }
    if (!lines_in_smem)
    {
        k_fit_lines<<<static_cast<unsigned int>((O + kFitThreads - 1) / kFitThreads), kFitThreads, 0, s>>>(F);
        GCR_CUDA(launch_chained(h, k_fit_rows, dim3(F.n_units), dim3(kFitThreads), s, F, slots));
    }
    else k_fit_rows<<<F.n_units, kFitThreads, 0, s>>>(F, slots); // first kernel of the call, behind the input copy: plain
    if (do_score)
    {
        // the fitted models (phi: the device's histogram mode) scored against every feature in the same stream, no round trip: the
        // rescoring step of the local optimisation (GCRANSAC.h:1004-1046) and of the final refit (:645-661) behind their fits
        Thresholds T = make_thresholds(thr, h->problem);
        if (score_flags & GCR_SCORE_TRUNCATED_SQ)
        {
            T.tsq[0] = (1.5 * thr[0]) * (1.5 * thr[0]);
            T.tsq[1] = (1.5 * thr[1]) * (1.5 * thr[1]);
        }
        T.strict = (score_flags & GCR_SCORE_STRICT) ? 1 : 0;
        const Buffers B = make_buffers(h, 0, false, nullptr);
        unsigned char* out_base = out_place ? h->h_fit_out.p : h->d_fit_out.p;
        const unsigned int un = static_cast<unsigned int>(K);
        const unsigned long long su =
            static_cast<unsigned long long>(un) * ((h->ns + kExactThreads - 1) / kExactThreads + (h->no + kExactThreads - 1) / kExactThreads);
        const int sgrid = static_cast<int>(std::min<unsigned long long>(su, static_cast<unsigned long long>(h->n_sm) * 8));
        GCR_CUDA(launch_chained(h, k_exact_items, dim3(sgrid), dim3(kExactThreads), s, B, T, 2, un)); // (slots set by k_fit_finish)
        GCR_CUDA(launch_chained(h, k_score_out, dim3(std::max(1, std::min<int>(h->n_sm, static_cast<int>((Wm * K + 255) / 256)))), dim3(256), s, B, T,
                                un, reinterpret_cast<double*>(out_base + out_score), reinterpret_cast<unsigned int*>(out_base + out_cnt),
                                mask_words ? reinterpret_cast<unsigned int*>(out_base + out_masks) : nullptr));
    }
    GCR_CUDA(cudaGetLastError());
    if (!out_place) GCR_CUDA(cudaMemcpyAsync(h->h_fit_out.p, h->d_fit_out.p, out_bytes, cudaMemcpyDeviceToHost, s));
    GCR_CUDA(cudaStreamSynchronize(s));

    const double* m = reinterpret_cast<const double*>(h->h_fit_out.p + out_models);
    const int* okd = reinterpret_cast<const int*>(h->h_fit_out.p + out_ok);
    const unsigned int* nent = reinterpret_cast<const unsigned int*>(h->h_fit_out.p + out_nent);
    const FitEntry* ent = reinterpret_cast<const FitEntry*>(h->h_fit_out.p + out_ent);
    std::vector<size_t> phi_differs;
    for (size_t k = 0; k < K; ++k)
    {
        std::memcpy(models + 4 * k, m + 4 * k, 32);
        ok[k] = okd[k];
        if (!hybrid || !okd[k]) continue;
        const double phi = replay_phi(ent + desc[k].ent_off, nent[k]);
        models[4 * k + 3] = phi;
        if (do_score && phi != m[4 * k + 3]) phi_differs.push_back(k); // equally heavy bins, resolved differently: scored again below
    }
    if (do_score)
    {
        const double* sc = reinterpret_cast<const double*>(h->h_fit_out.p + out_score);
        const uint32_t* cn = reinterpret_cast<const uint32_t*>(h->h_fit_out.p + out_cnt);
        for (size_t k = 0; k < K; ++k)
        {
            if (score) score[k] = okd[k] ? sc[k] : 0.0;
            if (n_inl)
            {
                n_inl[2 * k] = okd[k] ? cn[2 * k] : 0u;
                n_inl[2 * k + 1] = okd[k] ? cn[2 * k + 1] : 0u;
            }
        }
        if (mask_words) std::memcpy(mask_words, h->h_fit_out.p + out_masks, 4 * Wm * K);
        h->n_phi_fallbacks += static_cast<long long>(phi_differs.size());
        if (!phi_differs.empty())
        {
            // the device's restatement of the map order disagreed with the real map (another standard library's policy): those models are
            // scored again with the host's phi, all of them in one pass
            const size_t nd = phi_differs.size();
            std::vector<double> mm(4 * nd), ss(nd);
            std::vector<uint32_t> cc(2 * nd), ww(mask_words ? Wm * nd : 0);
            for (size_t i = 0; i < nd; ++i) std::memcpy(&mm[4 * i], models + 4 * phi_differs[i], 32);
            const int rc = gcr_score_models(h, mm.data(), static_cast<int64_t>(nd), thr, score_flags, ss.data(), cc.data(), mask_words ? ww.data() : nullptr);
            if (rc != GCR_OK) return rc;
            for (size_t i = 0; i < nd; ++i)
            {
                const size_t k = phi_differs[i];
                if (score) score[k] = ss[i];
                if (n_inl)
                {
                    n_inl[2 * k] = cc[2 * i];
                    n_inl[2 * k + 1] = cc[2 * i + 1];
                }
                if (mask_words) std::memcpy(mask_words + Wm * k, &ww[Wm * i], 4 * Wm);
            }
        }
    }
    return GCR_OK;
}

int gcr_fit_models(gcr_handle h, int64_t n_fits, const uint32_t* scale_idx, const int64_t* scale_off,
                   const uint32_t* orient_idx, const int64_t* orient_off, double* models, int32_t* ok)
{
    return fit_models_impl(h, n_fits, scale_idx, scale_off, orient_idx, orient_off, models, ok, nullptr, 0, nullptr, nullptr, nullptr);
}

long long gcr_debug_phi_fallbacks(gcr_handle h) { return h ? h->n_phi_fallbacks : -1; }

int gcr_fit_score_models(gcr_handle h, int64_t n_fits, const uint32_t* scale_idx, const int64_t* scale_off, const uint32_t* orient_idx,
                         const int64_t* orient_off, const double thr[2], uint32_t flags, double* models, int32_t* ok, double* score,
                         uint32_t* n_inl, uint32_t* mask_words)
{
    if (!thr) return GCR_ERR_INVALID;
    return fit_models_impl(h, n_fits, scale_idx, scale_off, orient_idx, orient_off, models, ok, thr, flags, score, n_inl, mask_words);
}

int gcr_lo_round(gcr_handle h, gcr_lo_round_io* io, const double thr[2], uint32_t label_flags, int32_t n_trials, const int32_t lim[2],
                 const int32_t min_inl[2], uint32_t* mask_words)
{
    if (!h || !io || !thr || !lim || !min_inl) return GCR_ERR_INVALID;
    if (h->ns < 2) return h->fail(GCR_ERR_STATE, "gcr_lo_round before gcr_set_features");
    const bool hybrid = h->problem == 0;
    if (!(thr[0] >= 0.0) || !std::isfinite(thr[0]) || (hybrid && (!(thr[1] >= 0.0) || !std::isfinite(thr[1]))))
        return h->fail(GCR_ERR_INVALID, "thresholds must be finite and non-negative");
    if (n_trials < 1 || n_trials > kCandCap) return h->fail(GCR_ERR_INVALID, "gcr_lo_round: 1 <= n_trials <= 1024");
    if (lim[0] < 1 || lim[0] > kLoMaxSubset || (hybrid && (lim[1] < 1 || lim[1] > kLoMaxSubset)) || min_inl[0] < 0 || min_inl[1] < 0)
        return h->fail(GCR_ERR_INVALID, "gcr_lo_round: sub-sample sizes must be 1 ... 32 per class");
    GCR_DEVICE(h);
    const int rc = ensure_common(h);
    if (rc != GCR_OK) return rc;
    const size_t K = static_cast<size_t>(n_trials);
    const size_t lim_s = static_cast<size_t>(lim[0]), lim_o = hybrid ? static_cast<size_t>(lim[1]) : 1;
    const size_t Wm = static_cast<size_t>((h->ns + 31) / 32 + (h->no + 31) / 32);
    auto align16 = [](size_t x) { return (x + 15) & ~static_cast<size_t>(15); };
    const size_t ent_per_fit = hybrid ? std::min<size_t>(lim_o, kPhiBins) : 0;
    const size_t out_models = 0, out_ok = align16(K * 32), out_nent = align16(out_ok + 4 * K), out_ent = align16(out_nent + 4 * K),
                 out_bytes = align16(out_ent + K * ent_per_fit * sizeof(FitEntry));
    GCR_CUDA(h->d_fit_out.ensure(out_bytes));
    GCR_CUDA(h->h_fit_out.ensure(out_bytes));
    GCR_CUDA(h->d_fit_partial.ensure(K * 2 * kFitGram));
    GCR_CUDA(h->d_lo_in.ensure(static_cast<size_t>(h->ns) + static_cast<size_t>(std::max(h->no, 0)) + 1));
    GCR_CUDA(h->d_lo_info.ensure(4));
    GCR_CUDA(h->d_lo_idx.ensure(K * (lim_s + lim_o)));
    GCR_CUDA(h->d_lo_desc.ensure(K));
    const size_t lo_words = align16(sizeof(LoResult));
    GCR_CUDA(h->h_lo_out.ensure(lo_words + 4 * Wm));
    if (!h->d_fit_done.p)
    {
        GCR_CUDA(h->d_fit_done.ensure(kCandCap));
        GCR_CUDA(cudaMemsetAsync(h->d_fit_done.p, 0, sizeof(unsigned int) * kCandCap, h->stream));
    }
    cudaStream_t s = h->stream;

    LoRound R;
    R.model = Model4{io->model[0], io->model[1], io->model[2], io->model[3]};
    R.score = io->score;
    R.rng_state = io->rng_state;
    R.n_trials = n_trials;
    R.lim[0] = lim[0];
    R.lim[1] = hybrid ? lim[1] : 0;
    R.min_inl[0] = min_inl[0];
    R.min_inl[1] = min_inl[1];
    R.hybrid = hybrid ? 1 : 0;
    LoDevice L;
    L.in_idx = h->d_lo_in.p;
    L.info = h->d_lo_info.p;
    L.desc = h->d_lo_desc.p;
    L.idx_s = h->d_lo_idx.p;
    L.idx_o = h->d_lo_idx.p + K * lim_s;
    L.out = reinterpret_cast<LoResult*>(h->h_lo_out.p);
    L.out_words = reinterpret_cast<unsigned int*>(h->h_lo_out.p + lo_words);

    FitBuffers F;
    F.scale = h->scale_ptr;
    F.orient = h->orient_ptr;
    F.desc = L.desc;
    F.idx_s = L.idx_s;
    F.idx_o = L.idx_o;
    F.lines = nullptr; // (<= 32 orientation features per fit: lines in shared memory)
    F.done = h->d_fit_done.p;
    F.partial = h->d_fit_partial.p;
    F.models = reinterpret_cast<double*>(h->d_fit_out.p + out_models);
    F.ok = reinterpret_cast<int*>(h->d_fit_out.p + out_ok);
    F.n_ent = reinterpret_cast<unsigned int*>(h->d_fit_out.p + out_nent);
    F.ent = reinterpret_cast<FitEntry*>(h->d_fit_out.p + out_ent);
    F.h_models = reinterpret_cast<double*>(h->h_fit_out.p + out_models);
    F.h_ok = reinterpret_cast<int*>(h->h_fit_out.p + out_ok);
    F.h_n_ent = reinterpret_cast<unsigned int*>(h->h_fit_out.p + out_nent);
    F.h_ent = reinterpret_cast<FitEntry*>(h->h_fit_out.p + out_ent);
    F.n_fits = n_trials;
    F.n_units = static_cast<unsigned int>(K);
    F.n_lines = 0;
    F.problem = h->problem;
    fill_rehash_schedule(F);

    Thresholds T = make_thresholds(thr, h->problem), TL = T;
    if (label_flags & GCR_SCORE_TRUNCATED_SQ)
    {
        TL.tsq[0] = (1.5 * thr[0]) * (1.5 * thr[0]);
        TL.tsq[1] = (1.5 * thr[1]) * (1.5 * thr[1]);
    }
    TL.strict = (label_flags & GCR_SCORE_STRICT) ? 1 : 0;
    const Buffers B = make_buffers(h, 0, false, nullptr);
    const FitSlots slots{B.cand_model, B.cand_hyp, B.cand_qsum, B.cand_rawcnt};
    const unsigned long long upi = (h->ns + kExactThreads - 1) / kExactThreads + (h->no + kExactThreads - 1) / kExactThreads;
    const int grid1 = static_cast<int>(std::min<unsigned long long>(upi, static_cast<unsigned long long>(h->n_sm) * 8));
    const int gridK = static_cast<int>(std::min<unsigned long long>(upi * K, static_cast<unsigned long long>(h->n_sm) * 8));
    k_lo_label<<<grid1, kExactThreads, 0, s>>>(B, TL, R);
    GCR_CUDA(launch_chained(h, k_lo_draw, dim3(1), dim3(kLoThreads), s, B, R, L));
    GCR_CUDA(launch_chained(h, k_fit_rows, dim3(F.n_units), dim3(kFitThreads), s, F, slots));
    GCR_CUDA(launch_chained(h, k_exact_items, dim3(gridK), dim3(kExactThreads), s, B, T, 2, static_cast<unsigned int>(K)));
    GCR_CUDA(launch_chained(h, k_lo_select, dim3(1), dim3(256), s, B, T, R, L, static_cast<const double*>(F.models), static_cast<const int*>(F.ok)));
    GCR_CUDA(cudaGetLastError());
    GCR_CUDA(cudaStreamSynchronize(s));

    const LoResult& r = *reinterpret_cast<const LoResult*>(h->h_lo_out.p);
    io->phi_mismatch = 0;
    if (hybrid && r.sampled)
    {
        // the device's restatement of the map order against this host's real map, on every trial's histogram entries
        const double* m = reinterpret_cast<const double*>(h->h_fit_out.p + out_models);
        const int* okd = reinterpret_cast<const int*>(h->h_fit_out.p + out_ok);
        const unsigned int* nent = reinterpret_cast<const unsigned int*>(h->h_fit_out.p + out_nent);
        const FitEntry* ent = reinterpret_cast<const FitEntry*>(h->h_fit_out.p + out_ent);
        for (size_t k = 0; k < K; ++k)
            if (okd[k] && replay_phi(ent + k * ent_per_fit, nent[k]) != m[4 * k + 3]) io->phi_mismatch += 1;
        h->n_phi_fallbacks += io->phi_mismatch;
    }
    io->updated = r.updated;
    io->sampled = r.sampled;
    io->n_models_scored = r.n_models_scored;
    io->n_labelled[0] = r.n_in[0];
    io->n_labelled[1] = r.n_in[1];
    if (io->phi_mismatch) return GCR_OK; // the caller repeats the round through gcr_score_models / gcr_fit_score_models; nothing else is reported
    io->rng_state += r.draws * 0x9E3779B97F4A7C15ull;
    if (r.updated)
    {
        std::memcpy(io->model, r.model, sizeof(io->model));
        io->score = r.score;
        io->n_inl[0] = r.n_inl[0];
        io->n_inl[1] = r.n_inl[1];
        if (mask_words) std::memcpy(mask_words, h->h_lo_out.p + lo_words, 4 * Wm);
    }
    return GCR_OK;
}

int gcr_score_model(gcr_handle h, const double model[4], const double thr[2], double* score, uint32_t n_inl[2],
                    uint32_t* scale_mask_words, uint32_t* orient_mask_words)
{
    if (!h || !model || !thr) return GCR_ERR_INVALID;
    if (h->ns < 2) return h->fail(GCR_ERR_STATE, "gcr_score_model before gcr_set_features");
    const size_t ws = (h->ns + 31) / 32, wo = (h->no + 31) / 32;
    std::vector<uint32_t> words;
    const bool want_masks = scale_mask_words || orient_mask_words;
    if (want_masks) words.resize(ws + wo);
    const int rc = gcr_score_models(h, model, 1, thr, 0, score, n_inl, want_masks ? words.data() : nullptr);
    if (rc != GCR_OK) return rc;
    if (scale_mask_words) std::memcpy(scale_mask_words, words.data(), 4 * ws);
    if (orient_mask_words) std::memcpy(orient_mask_words, words.data() + ws, 4 * wo);
    return GCR_OK;
}

int gcr_debug_eval(gcr_handle h, int which, const double* in, int64_t n, const double model[4], double* out)
{
    if (!h || !in || !out || n < 1 || which < 0 || which > 4) return GCR_ERR_INVALID;
    if (!h->stream) return h->fail(GCR_ERR_STATE, "engine was not created successfully");
    GCR_DEVICE(h);
    GCR_CUDA(h->d_dbg_in.ensure(3 * static_cast<size_t>(n)));
    GCR_CUDA(h->d_dbg_out.ensure(static_cast<size_t>(n)));
    Model4 m = {0, 0, 1, 0};
    if (model) m = Model4{model[0], model[1], model[2], model[3]};
    GCR_CUDA(cudaMemcpyAsync(h->d_dbg_in.p, in, sizeof(double) * 3 * n, cudaMemcpyHostToDevice, h->stream));
    k_debug_eval<<<static_cast<int>((n + 127) / 128), 128, 0, h->stream>>>(which, h->d_dbg_in.p, static_cast<int>(n), m,
                                                                           h->d_dbg_out.p);
    GCR_CUDA(cudaGetLastError());
    GCR_CUDA(cudaMemcpyAsync(out, h->d_dbg_out.p, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
    GCR_CUDA(cudaStreamSynchronize(h->stream));
    return GCR_OK;
}

int gcr_score_models_static(int device, const double* d_scale, int64_t ns, const double* d_orient, int64_t no, const double* d_models, int64_t n,
                            const double thr[2], uint32_t flags, double* d_score, uint32_t* d_n_inl, uint32_t* d_mask_words, void* cuda_stream,
                            char* err, size_t err_cap)
{
    auto fail = [&](int code, const char* msg) {
        if (err && err_cap) std::snprintf(err, err_cap, "%s", msg);
        return code;
    };
    if (err && err_cap) err[0] = 0;
    if (!d_scale || !d_orient || !d_models || !thr) return fail(GCR_ERR_INVALID, "gcr_score_models_static: null input pointer");
    if (device < 0 || device >= kStaticMaxDevices) return fail(GCR_ERR_INVALID, "gcr_score_models_static: device index out of range (0..63)");
    if (ns < 2 || no < 2 || ns > (1ll << 30) || no > (1ll << 30))
        return fail(GCR_ERR_INVALID, "gcr_score_models_static: each feature class needs 2 .. 2^30 rows (hybrid problem only)");
    if (n < 1 || n > kStaticMaxModels) return fail(GCR_ERR_INVALID, "gcr_score_models_static: 1 <= n <= 1024 models per call");
    if (!(thr[0] >= 0.0) || !std::isfinite(thr[0]) || !(thr[1] >= 0.0) || !std::isfinite(thr[1]))
        return fail(GCR_ERR_INVALID, "gcr_score_models_static: thresholds must be finite and non-negative");
    const size_t W = static_cast<size_t>((ns + 31) / 32 + (no + 31) / 32);
    if (static_cast<size_t>(n) * W * 4 > kStaticScratchBytes - kStaticOffMasks)
    {
        char msg[256];
        std::snprintf(msg, sizeof(msg), "gcr_score_models_static: n x mask words = %lld x %zu exceeds the static scratch (%zu words: 1 MiB array, 76 KB of it "
                      "accumulators); score fewer models per call or use gcr_score_models", static_cast<long long>(n), W, (kStaticScratchBytes - kStaticOffMasks) / 4);
        return fail(GCR_ERR_INVALID, msg);
    }
    DeviceGuard device_guard_(device);
    if (device_guard_.err != cudaSuccess) return fail(GCR_ERR_CUDA, cudaGetErrorString(device_guard_.err));
    unsigned char* sp = nullptr;
    cudaError_t e = static_scratch_address(device, &sp);
    if (e != cudaSuccess) return fail(GCR_ERR_CUDA, cudaGetErrorString(e));

    Thresholds T = make_thresholds(thr, GCR_PROBLEM_HYBRID);
    if (flags & GCR_SCORE_TRUNCATED_SQ)
    {
        T.tsq[0] = (1.5 * thr[0]) * (1.5 * thr[0]);
        T.tsq[1] = (1.5 * thr[1]) * (1.5 * thr[1]);
    }
    T.strict = (flags & GCR_SCORE_STRICT) ? 1 : 0;
    Buffers B;
    std::memset(&B, 0, sizeof(B));
    B.scale = d_scale;
    B.orient = d_orient;
    B.ns = static_cast<int>(ns);
    B.no = static_cast<int>(no);
    B.ws = (B.ns + 31) / 32;
    B.wmask = static_cast<int>(W);
    B.cand_model = reinterpret_cast<Model4*>(sp + kStaticOffModel);
    B.cand_hyp = reinterpret_cast<unsigned int*>(sp + kStaticOffHyp);
    B.cand_qsum = reinterpret_cast<unsigned long long*>(sp + kStaticOffQsum);
    B.cand_rawcnt = reinterpret_cast<unsigned int*>(sp + kStaticOffRaw);
    B.cand_score = reinterpret_cast<double*>(sp + kStaticOffScore);
    B.cand_cnt = reinterpret_cast<unsigned int*>(sp + kStaticOffCnt);
    B.cand_masks = reinterpret_cast<unsigned int*>(sp + kStaticOffMasks);

    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
    const unsigned int un = static_cast<unsigned int>(n);
    const unsigned long long units =
        static_cast<unsigned long long>(un) * ((B.ns + kExactThreads - 1) / kExactThreads + (B.no + kExactThreads - 1) / kExactThreads);
    int n_sm = 148;
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, device);
    const int grid = static_cast<int>(std::min<unsigned long long>(units, static_cast<unsigned long long>(n_sm) * 8));
    // the same three kernels as gcr_score_models, plain launches in stream order: models -> scratch and cleared accumulators (the array keeps
    // the last call's contents) | FP64 reference-order scoring (every mask word of every model is written) | values, counts, masks -> caller
    k_set_models<<<(un + 127) / 128, 128, 0, s>>>(B, reinterpret_cast<const Model4*>(d_models), un);
    k_exact_items<<<grid, kExactThreads, 0, s>>>(B, T, 2, un);
    // (outputs the caller did not ask for land in the scratch's own value / count slots)
    k_score_out<<<std::max(1, std::min<int>(n_sm, static_cast<int>((W * n + 255) / 256))), 256, 0, s>>>(
        B, T, un, d_score ? d_score : B.cand_score, d_n_inl ? d_n_inl : B.cand_cnt, d_mask_words);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(GCR_ERR_CUDA, cudaGetErrorString(e));
    return GCR_OK;
}

} // extern "C"
