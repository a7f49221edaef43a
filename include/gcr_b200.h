/*
 * gcr_b200.h -- C ABI of the B200-native hybrid-RANSAC hypothesis engine (libgcr_b200.so).
 *
 * One batched device call replaces the reference's per-iteration hypothesise/verify loop
 *     sample -> isValidSample -> estimateModel -> getScore -> best-model update
 * of yuvalnis/graph-cut-ransac, src/pygcransac/include/GCRANSAC.h:286-485, for the hybrid
 * (scale-change + orientation-change) rectifying-homography problem:
 *     sampler      GCRANSAC.h:53-80,160-189      (get_random_subset / GCRANSAC::sample)
 *     degeneracy   estimators/solver_rectifying_homography_two_sift.hpp:141-215  (isValidSample)
 *     solver       ...two_sift.hpp:261-344       (estimateMinimalModel), :45-61 (isValidModel)
 *     residuals    ...two_sift.hpp:621-713, model.h:128-204, math_utils.hpp:78-102
 *     MSAC         scoring_functions/MSAC_scoring_function.hpp:22-130, score.hpp:11-102
 *     selection    GCRANSAC.h:440-462 (strict '<' on Score::value, earliest wins ties)
 *
 * Conventions: plain pointers and sizes, no C++/torch types; every call returns 0 (GCR_OK) or a
 * negative gcr_status and never throws; the caller owns every buffer it passes, the handle owns
 * all device memory; one CUDA stream per handle (replaceable with gcr_set_stream so the engine
 * can be ordered against a caller's NCCL / PyTorch work).  Feature rows are the reference's:
 * scale (x, y, s) and orientation (x, y, theta[rad]), float64, row-major [N,3]
 * (bindings.cpp:181-250, gcransac_python.cpp:284-296).
 *
 * Packed masks: uint32 words, bit j of word w <-> feature 32*w + j, scale class first
 * (ceil(Ns/32) words) then orientation class (ceil(No/32) words).
 */
#ifndef GCR_B200_H
#define GCR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gcr_engine_s* gcr_handle;

enum gcr_status
{
    GCR_OK = 0,
    GCR_ERR_INVALID = -1, /* bad argument (the binding maps it to std::invalid_argument) */
    GCR_ERR_CUDA = -2,    /* a CUDA call failed; see gcr_last_error */
    GCR_ERR_STATE = -3,   /* call order violated (e.g. run before set_features) */
    GCR_ERR_NOMEM = -4
};

/* per-hypothesis flag bits (gcr_fetch) */
#define GCR_FLAG_SAMPLE_VALID 1u /* isValidSample accepted the minimal sample */
#define GCR_FLAG_SOLVED 2u       /* estimateMinimalModel produced a model */
#define GCR_FLAG_MODEL_VALID 4u  /* isValidModel (max(|h7|,|h8|) < 1e-3) */
#define GCR_FLAG_EXACT 8u        /* scored entirely in FP64 (outside the FP32 fast path's proven range) */
/* (bits 16 and 32 are internal to the selection: deferred candidate; contender of a winner-only run) */

/* gcr_run flags */
#define GCR_RUN_EMIT_MASKS 1u        /* keep the packed inlier mask of EVERY hypothesis (parity / replay mode) */
#define GCR_RUN_SAMPLES_ON_DEVICE 2u /* `samples` is a device pointer */
#define GCR_RUN_KEEP_PER_HYP 4u      /* keep per-hypothesis models/scores/counts for gcr_fetch (implied by EMIT_MASKS) */
#define GCR_RUN_ALL_EXACT 8u         /* score every hypothesis with the FP64 reference-order kernel (validation) */
#define GCR_RUN_EXCHANGE 16u         /* multi-GPU: exchange winner records with the attached peers inside the last kernel */
#define GCR_RUN_SAMPLES_U16 64u      /* `samples` holds uint16_t[n_hyp,4] instead of int32_t[n_hyp,4] (both classes <= 65 535 features; the unused
                                      * fourth index of the scale-only problems is 0xFFFF): half the bytes over PCIe for a replayed host array */
#define GCR_RUN_TIME_KERNELS 32u     /* diagnostics: CUDA events around the whole stage and around the scoring kernel (gcr_result.device_ms,
                                      * score_kernel_ms; both 0 without it): four more in-stream operations, ~2.5 us each */

typedef struct gcr_result
{
    int64_t best_index;    /* GLOBAL hypothesis index (first_hyp + local) of the selected model, -1 if none */
    double best_score;     /* its MSAC score, FP64 reference-order evaluation */
    double best_model[4];  /* h7, h8, alpha, phi; the vanishing line is (h7, h8, 1) */
    uint32_t best_n_inl[2];
    uint32_t pad_[2];
    int64_t n_hyp;          /* hypothesis slots processed by this call */
    int64_t n_sample_valid; /* ... that passed isValidSample */
    int64_t n_solved;       /* ... that yielded a model (these are scored against every feature) */
    int64_t n_rechecked;    /* (hypothesis, feature) pairs inside the FP32 guard band, re-evaluated in FP64 */
    int64_t n_exact_hyp;    /* hypotheses scored entirely in FP64 */
    int64_t n_candidates;   /* near-best hypotheses re-scored in FP64 before the final selection */
    float device_ms;        /* CUDA-event time of the whole hypothesis stage on the engine stream (GCR_RUN_TIME_KERNELS; else 0) */
    float score_kernel_ms;  /* CUDA-event time of the dominant FP32 scoring kernel alone (GCR_RUN_TIME_KERNELS; else 0) */
    int32_t winner_rank;    /* rank owning best_index (0 unless the run exchanged records with peers) */
    int32_t score_kernel;   /* which scoring kernel ran: 0 = k_score_fast (TMA ring), 1 = k_score_resident (working set in shared memory) */
    float tail_phase_us[6]; /* diagnostics (device timer): fix-up+FP64 items | finalize | collect | candidate re-scoring | select | start of the scoring kernel -> start of the tail */
    float stage_phase_us[2]; /* diagnostics (device timer): start of the sampler/solver kernel -> start of the scoring kernel | -> end of the tail */
    int64_t n_refined;      /* winner-only runs: contenders (upper bound of the FP32 value >= best lower bound) re-scored in mid precision with
                             * exact inlier decisions before the FP64 candidates were chosen; 0 in the reporting modes */
    int64_t n_scored;       /* hypotheses scored against every feature (FP32 tiles + precise tiles + FP64 list); == n_solved in the reporting
                             * modes, fewer in winner-only runs, which skip models that isValidModel rejects (they can never be selected) */
} gcr_result;

/* Winner record written to device memory by the last kernel of gcr_run when `d_record` is given:
 * ready to be the send buffer of an NCCL all-gather (one 64-byte record per rank). */
typedef struct gcr_record
{
    double score;        /* 0 when no model */
    int64_t index;       /* global hypothesis index, -1 when no model */
    double model[4];
    uint32_t n_inl[2];
    uint64_t reserved;   /* rank that owns the winning hypothesis (exchanging runs), else 0 */
} gcr_record;

int gcr_create(int device, gcr_handle* out);
int gcr_destroy(gcr_handle h);
/* Size every launch of this engine for 1/share of the device's SMs (default 1: all of them).  For batches of small problems -- the
 * reference's per-image entry point called in a loop, src/gcransac_python.cpp:256-406 -- with one engine per stream: a full-size engine's
 * kernels each fill the GPU, so streams only queue behind one another and an image costs the SUM of its kernels' latencies; `share`
 * engines of a 1/share size run their launch sequences side by side (256 images of 5 000 features x 100 000 hypotheses: 19.1 -> 11.4 ms
 * with 8 engines of share 8).  Results do not depend on the share.  Call between runs; it stays until changed.  Keep the engines in flight
 * at or below `share`: the last kernel of a run is a cooperative launch (two CTAs per SM of the engine's share), and the cooperative grids
 * of all runs in flight should fit the device together. */
int gcr_set_sm_share(gcr_handle h, int share);
const char* gcr_last_error(gcr_handle h);


/* Feature front end on the device: the reference's examples/utils.py:5-49 between cv2.SIFT and pygcransac, for keypoint
 * rows (x, y, size, angle in degrees; angle == -1: no orientation), float32 (is_f32 = 1, what cv2.KeyPoint holds) or
 * float64, in host memory or, on_device = 1, device memory of the engine's device (16-byte aligned).
 *   scale features       one (x, y, size) row per unique (int(x), int(y)): the first keypoint of each coordinate, in
 *                        order of first appearance                          (scale_features_from_sift, utils.py:5-26)
 *   orientation features (x, y, deg2rad(angle)) of every keypoint with angle != -1, in input order; 0.5 * size per
 *                        row goes to orient_sizes                    (orientation_features_from_sift, utils.py:29-49)
 * The two arrays become the engine's resident features exactly as after gcr_set_features (scale-only problems keep
 * the scale rows only); gcr_get_features returns them.  n_scale / n_orient (nullable) receive the row counts;
 * orient_sizes (nullable) is a HOST buffer of capacity n.  Results are bit-identical to the Python functions.
 * When a class ends up with fewer rows than a minimal sample the call still succeeds and the rows can be fetched,
 * but the engine holds no features to run on (gcr_run: GCR_ERR_STATE).  Synchronises. */
int gcr_set_keypoints(gcr_handle h, const void* keypoints, int64_t n, int is_f32, int on_device, int64_t* n_scale,
                      int64_t* n_orient, double* orient_sizes);

/* Which of the reference's three estimators the engine runs (default: hybrid).  Call before gcr_set_features.
 *   GCR_PROBLEM_HYBRID               findRectifyingHomographySIFT: 2 scale + 2 orientation samples (two_sift.hpp)
 *   GCR_PROBLEM_SCALE_ONLY           findRectifyingHomographyScaleOnly: 3 scale samples, rows [x, y, s^(1/3)] = 1,
 *                                    residual |log(alpha^3 s w^-3)| (three_sift.hpp:145-193,293-317)
 *   GCR_PROBLEM_SCALE_ONLY_ORIGINAL  findRectifyingHomographyScaleOnlyOriginal: rows [x, y, -s^(-1/3)] = -1, residual
 *                                    |log(s w^-3 / alpha^3)| (three_sift_original.hpp)
 * Scale-only problems: gcr_set_features(h, scale, Ns >= 3, NULL, 0, .), samples are int32[H,4] = (s0, s1, s2, -1), thr[1]
 * is ignored, models are (h7, h8, alpha, 0), every solved model is valid, a score needs >= 3 inliers, masks have
 * ceil(Ns/32) words. */
#define GCR_PROBLEM_HYBRID 0
#define GCR_PROBLEM_SCALE_ONLY 1
#define GCR_PROBLEM_SCALE_ONLY_ORIGINAL 2
int gcr_set_problem(gcr_handle h, int problem);

/* Use `cuda_stream` (a cudaStream_t) for all subsequent work; NULL restores the engine's own stream. */
int gcr_set_stream(gcr_handle h, void* cuda_stream);

/* Order the engine's stream after everything submitted so far to `producer_stream` (a cudaStream_t; NULL = the legacy default stream):
 * an event is recorded there and the engine's stream waits for it.  For callers that hand over device arrays (GCR_FEATURES_ON_DEVICE,
 * GCR_RUN_SAMPLES_ON_DEVICE) written by kernels of another stream -- e.g. a PyTorch tensor produced on torch's current stream -- and do
 * not want to move the engine onto that stream with gcr_set_stream.  Does not synchronise the host. */
int gcr_wait_stream(gcr_handle h, void* producer_stream);

/* Upload (or, with GCR_FEATURES_ON_DEVICE, adopt from device memory) the two feature classes and build the FP32
 * structure-of-arrays working set.  Replaces the cv::Mat construction of gcransac_python.cpp:284-296.  Requires ns >= 2 and
 * no >= 2 (bindings.cpp:199-224).  `where` is a flags word:
 *   0                         host arrays.  Pageable memory is copied before the call returns.  PAGE-LOCKED (device-mapped) arrays are
 *                             read IN PLACE by the preparation kernel after the call has returned: the caller must leave them untouched
 *                             until the next synchronising call on this handle (gcr_run with out != NULL, gcr_best, gcr_fetch, ...),
 *                             as with any cudaMemcpyAsync from pinned memory.  The same holds for pinned `samples` of gcr_run.
 *   GCR_FEATURES_ON_DEVICE    device pointers (16-byte aligned), adopted without a copy: the caller keeps them alive and unchanged for
 *                             as long as the engine runs on them, and orders their producer before this call on the engine's stream
 *                             (gcr_set_stream) or by an event.
 *   GCR_FEATURES_HOST_LIBM    host arrays only: bit-exact replay mode.  The minimal solvers' per-feature transcendental columns
 *                             (pow(s, 1/3): two_sift.hpp:288-299; sin / cos of the orientation: math_utils.hpp:104-109) are evaluated by
 *                             the host's C library -- the functions the reference itself calls -- instead of the device's, so that
 *                             isValidSample's answer is the reference's bit for bit even where it hangs on the last ulp of sin / cos
 *                             (two orientation samples at one position).  Costs ns + no host transcendentals; synchronises. */
#define GCR_FEATURES_ON_DEVICE 1
#define GCR_FEATURES_HOST_LIBM 2
int gcr_set_features(gcr_handle h, const double* scale, int64_t ns, const double* orient, int64_t no, int where);

/* ---- multi-GPU winner exchange over peer memory (NVLink / NVSwitch), one engine per GPU -----------------------------
 * The path's only exchange step is "which rank holds the best model" (the reference is single-process; this is its
 * best-model rule GCRANSAC.h:440-462 applied across shards).  It is fused into the kernel that produces the record:
 * the selecting CTA stores its 64-byte gcr_record into every peer's exchange buffer with system-scope release stores and
 * waits (acquire loads) for the peers' records in its own buffer, so gcr_run's result IS the global winner and no separate
 * collective is launched.  Setup (once):
 *   every rank: gcr_peer_export(h, handle)                      -> 64-byte CUDA IPC handle of its exchange buffer
 *   all-gather the handles by any means (torch.distributed, MPI, a file)
 *   every rank: gcr_peer_attach(h, handles (world x 64 B), world, rank); then a barrier
 * Single-process multi-GPU callers use gcr_peer_buffer + gcr_peer_attach_ptrs with raw device pointers instead
 * (peer access enabled by the caller).  Every rank must then issue the same sequence of GCR_RUN_EXCHANGE runs; a peer
 * that never arrives makes the run fail with GCR_ERR_STATE after 4 s instead of hanging. */
#define GCR_IPC_HANDLE_BYTES 64
int gcr_peer_export(gcr_handle h, void* ipc_handle_64);
int gcr_peer_attach(gcr_handle h, const void* ipc_handles, int world, int rank);
int gcr_peer_buffer(gcr_handle h, void** device_ptr);
int gcr_peer_attach_ptrs(gcr_handle h, void* const* peer_buffers, int world, int rank);
int gcr_peer_detach(gcr_handle h);

/* Copy the resident feature arrays back to HOST buffers f64[Ns,3] / f64[No,3] (either may be NULL).  The host-side
 * local optimisation and refit need the rows of the inliers when the caller handed over device tensors.  Synchronises. */
int gcr_get_features(gcr_handle h, double* scale, double* orient);

/* The hypothesis stage over `n_hyp` hypothesis slots with global indices
 * [first_hyp, first_hyp + n_hyp).
 *   samples == NULL : draw minimal samples on the device, Philox4x32-10 keyed by `seed`, counter =
 *                     global hypothesis index (results do not depend on how slots are sharded);
 *   samples != NULL : replay int32[n_hyp,4] = (scale0, scale1, orient0, orient1) supplied by the caller.
 *   thr             : the two residual thresholds (settings.threshold, gcransac_python.cpp:313-314).
 *   d_record        : optional DEVICE pointer to a gcr_record the final kernel fills (NCCL send buffer).
 * The call is asynchronous with respect to the host unless `out` is non-NULL, in which case it
 * synchronises the stream and fills *out. */
int gcr_run(gcr_handle h, int64_t n_hyp, uint64_t seed, uint64_t first_hyp, const int32_t* samples,
            const double thr[2], uint32_t flags, void* d_record, gcr_result* out);

/* Selected model of the last gcr_run: model, score, counts and packed masks (any pointer may be NULL).
 * Synchronises the stream.  *hyp_index = -1 and zeroed outputs when no model was found.
 * Device-side failures of the run -- replayed sample indices out of range (GCR_ERR_INVALID), guard-band worklist overflow or a winner
 * exchange that timed out (GCR_ERR_STATE) -- are reported here (and by gcr_fetch) also when the run itself was asynchronous
 * (gcr_run with out == NULL), from the counter snapshot the run's last kernel leaves in the result block. */
int gcr_best(gcr_handle h, double model[4], double* score, uint32_t n_inl[2], uint32_t* scale_mask_words,
             uint32_t* orient_mask_words, int64_t* hyp_index);

/* Per-hypothesis outputs of the last gcr_run (needs GCR_RUN_KEEP_PER_HYP / GCR_RUN_EMIT_MASKS), copied
 * to HOST buffers; any pointer may be NULL.  flags u8[H]; models f64[H,4]; score f64[H]; n_inl u32[H,2]
 * (after the MSAC zero-reset rule); masks u32[H, ceil(Ns/32)+ceil(No/32)]; samples i32[H,4]. */
int gcr_fetch(gcr_handle h, uint8_t* flags, double* models, double* score, uint32_t* n_inl, uint32_t* masks,
              int32_t* samples);

/* Score ONE caller-supplied model (h7,h8,alpha,phi) against the resident features with the FP64
 * reference-order kernel: MSACScoringFunction::getScore for a single model.  Used to materialise the
 * global winner's masks on every rank after the multi-GPU combine, and by the host-side local
 * optimisation.  Host output buffers, any may be NULL.  Synchronises. */
int gcr_score_model(gcr_handle h, const double model[4], const double thr[2], double* score, uint32_t n_inl[2],
                    uint32_t* scale_mask_words, uint32_t* orient_mask_words);

/* Batched form of gcr_score_model: `n` (<= 1024) caller-supplied models f64[n,4] scored against the resident
 * features in ONE device pass -- the rescoring step of the local optimisation's inner RANSAC
 * (GCRANSAC.h:963-1046: every trial model goes through getScore) and of the final refit (:628-667).
 * flags: GCR_SCORE_TRUNCATED_SQ compares r^2 with (1.5 t)^2, the relabelling rule of
 * graphCutLocalOptimization (GCRANSAC.h:207-208,931-937), instead of MSAC's (2.25 t) t.
 * Host outputs, any may be NULL: score f64[n], n_inl u32[n,2] (after the zero-reset rule),
 * mask_words u32[n, ceil(Ns/32)+ceil(No/32)].  Synchronises. */
#define GCR_SCORE_TRUNCATED_SQ 1u
#define GCR_SCORE_STRICT 2u /* r^2 < t^2 instead of <=: the one-class relabelling by graph cut over an empty graph (GCRANSAC.h:759-870) */
int gcr_score_models(gcr_handle h, const double* models, int64_t n, const double thr[2], uint32_t flags, double* score,
                     uint32_t* n_inl, uint32_t* mask_words);

/* Non-minimal least-squares fits on the device, `n_fits` (<= 1024) per call: the model-fitting step of the local
 * optimisation's inner RANSAC (GCRANSAC.h:955-1002 -> estimator.estimateModelNonminimal) and of the final refit on all
 * inliers (GCRANSAC.h:628-645).  Replaces RectifyingHomographyEstimator::estimateModelNonminimal
 * (rectifying_homography_estimator.h:164-227) with unit weights, i.e. estimateNonMinimalModel of two_sift.hpp:423-579
 * (rows per scale feature and per PAIR of orientation features, least squares, phi by the 0.5-degree histogram mode),
 * three_sift.hpp:195-254 and three_sift_original.hpp:229-234 for the scale-only problems (gcr_set_problem).
 * Fit k uses the resident scale features scale_idx[scale_off[k] .. scale_off[k+1]) and orientation features
 * orient_idx[orient_off[k] .. orient_off[k+1]) (HOST arrays; offsets start at 0; orient_* NULL for scale-only).
 * Host outputs: models f64[n_fits,4] = (h7, h8, alpha, phi), ok i32[n_fits] (0 where the reference's solver returns no
 * model: too few constraints, non-finite solution, alpha < 1e-9).  Parity: model <= 1e-6 relative.  Synchronises. */
int gcr_fit_models(gcr_handle h, int64_t n_fits, const uint32_t* scale_idx, const int64_t* scale_off,
                   const uint32_t* orient_idx, const int64_t* orient_off, double* models, int32_t* ok);

/* gcr_fit_models followed, in the same launch sequence and with ONE synchronisation, by the scoring of the fitted models against every
 * resident feature: "fit, then getScore" of one inner-RANSAC pass of the local optimisation (GCRANSAC.h:955-1046) and of the final refit
 * (GCRANSAC.h:628-661).  Arguments as gcr_fit_models and gcr_score_models (thr, flags: GCR_SCORE_*); host outputs, any of score / n_inl /
 * mask_words may be NULL: models f64[n_fits,4], ok i32[n_fits], score f64[n_fits], n_inl u32[n_fits,2] (after the zero-reset rule),
 * mask_words u32[n_fits, ceil(Ns/32)+ceil(No/32)]; fits with ok = 0 report score 0 and no inliers.  Results are those of the two separate
 * calls, bit for bit.  Synchronises. */
int gcr_fit_score_models(gcr_handle h, int64_t n_fits, const uint32_t* scale_idx, const int64_t* scale_off, const uint32_t* orient_idx,
                         const int64_t* orient_off, const double thr[2], uint32_t flags, double* models, int32_t* ok, double* score,
                         uint32_t* n_inl, uint32_t* mask_words);

/* One pass of the local optimisation's loop as ONE launch sequence with ONE synchronisation: GCRANSAC::graphCutLocalOptimization,
 * GCRANSAC.h:873-1062 (two classes; one class: :759-870 over the empty neighbourhood graph) --
 *   labels: every resident feature against io->model with the relabelling threshold (label_flags: GCR_SCORE_TRUNCATED_SQ, and
 *     GCR_SCORE_STRICT for the one-class problems)                                                            GCRANSAC.h:931-937
 *   inner RANSAC: n_trials times, min(lim[c], labelled inliers of class c) labelled inliers per class drawn without replacement
 *     (a class no larger than lim[c] is taken whole), non-minimal fit, getScore with `thr`                    GCRANSAC.h:955-1046
 *   update: the first trial that scores strictly higher than io->score and every trial before it              GCRANSAC.h:1034
 * -- what gcr_score_models(labels) + host-side draws + gcr_fit_score_models did in two synchronous calls.  The draws are those of a
 * SplitMix64 generator with state io->rng_state consumed in the order (trial, scale class, orientation class, draw): draw i of a
 * sub-sample swaps positions i and i + floor(u * (n - i) / 2^32) of the ascending list of labelled inliers (u = upper 32 bits of the
 * generator's next output) and takes the new entry at i -- a partial Fisher-Yates shuffle, the role of the reference's
 * get_random_subset (GCRANSAC.h:53-80); the reference's own sampler is seeded from std::random_device and cannot be replayed.
 * In: io->model, io->score (the so-far-best), io->rng_state.  Out: io->sampled = 0 if a class has no more than min_inl[c] labelled
 * inliers (the reference's sampling fails: the optimisation ends; nothing else changes); else io->n_models_scored trials produced a
 * model, io->rng_state is advanced by the draws, and if io->updated the model, score, n_inl (after the zero-reset rule) and mask_words
 * (u32[ceil(Ns/32)+ceil(No/32)], may be NULL) are the winning trial's.  io->phi_mismatch != 0 (never with libstdc++, see
 * gcr_debug_phi_fallbacks): the device's tie order of the orientation histogram differed from this host's std::unordered_map for that
 * many trials; io is otherwise unchanged and the caller runs the round through the two-call path.  lim[c] <= 32, n_trials <= 1024.
 * No counterpart as a single function in the reference: it is the body of its LO loop. */
typedef struct gcr_lo_round_io
{
    double model[4];
    double score;
    uint32_t n_inl[2];
    uint64_t rng_state;
    uint32_t updated, sampled, n_models_scored, phi_mismatch;
    uint32_t n_labelled[2]; /* out: labelled inliers per class */
} gcr_lo_round_io;
int gcr_lo_round(gcr_handle h, gcr_lo_round_io* io, const double thr[2], uint32_t label_flags, int32_t n_trials, const int32_t lim[2],
                 const int32_t min_inl[2], uint32_t* mask_words);

/* Diagnostics: fits of gcr_fit_score_models calls on this handle whose device-side phi (histogram mode under the device's restatement of
 * the reference's unordered_map order) differed from the host's replay of the real map and were scored a second time; 0 when the two agree. */
long long gcr_debug_phi_fallbacks(gcr_handle h);

/* Allocation-free batched model scoring.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same computation as
 * gcr_score_models for the hybrid problem -- MSACScoringFunction::getScore of n caller-supplied models in FP64 reference order -- without an
 * engine handle: the three kernels' scratch (model copies, fixed-point accumulators, counts, packed masks) lives in a file-scope
 * `static __device__` array of 1 MiB that every GPU loading the library holds until exit; the call performs no cudaMalloc / cudaFree /
 * allocating library call and does not synchronise.
 *   device                      CUDA device index (0..63) the pointers live on; the caller's current device is restored
 *   d_scale f64[ns,3], d_orient f64[no,3], d_models f64[n,4] = (h7, h8, alpha, phi)      DEVICE pointers, caller-owned inputs
 *   thr, flags                  as gcr_score_models (GCR_SCORE_TRUNCATED_SQ, GCR_SCORE_STRICT)
 *   d_score f64[n], d_n_inl u32[n,2] (after the zero-reset rule), d_mask_words u32[n, ceil(ns/32)+ceil(no/32)]
 *                               DEVICE pointers, caller-owned outputs, any may be NULL; valid once `cuda_stream` has passed the call
 *   cuda_stream                 a cudaStream_t (NULL = the legacy default stream): the kernels run there, in stream order
 *   err, err_cap                optional buffer for a message when the call fails (this entry point has no handle to ask)
 * Limits (GCR_ERR_INVALID with a message beyond them): 1 <= n <= 1024 models and n x (ceil(ns/32) + ceil(no/32)) <= 242 688 mask words
 * (e.g. 1024 models x 7 500 features, 110 models x 70 000 features); ns, no >= 2.
 * The array is SHARED by every call in the process on that GPU and keeps the last call's contents (the call clears what must start at
 * zero): calls that overlap on the GPU -- issued to different streams without ordering between them -- race on it; serialise them. */
int gcr_score_models_static(int device, const double* d_scale, int64_t ns, const double* d_orient, int64_t no, const double* d_models, int64_t n,
                            const double thr[2], uint32_t flags, double* d_score, uint32_t* d_n_inl, uint32_t* d_mask_words, void* cuda_stream,
                            char* err, size_t err_cap);

/* Device-side helper math, evaluated on the GPU by the same __device__ functions the kernels use.
 * which: 0 clipAngle(a) 1 minAngleDiff(a,b) 2 linesAnglesDiff(a,b)
 *        3 scaleResidual(x=a,y=b,s=c; model) 4 orientationResidual(x=a,y=b,theta=c; model)
 * in: f64[n,3] (a,b,c), model f64[4], out: f64[n].  Host buffers. */
int gcr_debug_eval(gcr_handle h, int which, const double* in, int64_t n, const double model[4], double* out);

/* Library / build identification, e.g. "gcr_b200 0.1 sm_100a". */
const char* gcr_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GCR_B200_H */
