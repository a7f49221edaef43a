// Host-side estimator on top of the C ABI (see gcr_host.hpp).
#include "gcr_host.hpp"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <map>
#include <mutex>
#include <random>
#include <stdexcept>

#include "../../../include/gcr_b200.h"
#include "gcr_refit.hpp"

namespace gcr_host
{
namespace
{

// One engine handle per device, created on first use and kept for the life of the process (device buffers are
// reused between calls).  The reference is re-entrant; so is this: a handle is checked out for one call.
class EnginePool
{
public:
    gcr_handle acquire(int device)
    {
        {
            std::lock_guard<std::mutex> lock(mu_);
            auto& idle = idle_[device];
            if (!idle.empty())
            {
                gcr_handle h = idle.back();
                idle.pop_back();
                return h;
            }
        }
        gcr_handle h = nullptr;
        const int rc = gcr_create(device, &h);
        if (rc != GCR_OK)
        {
            const std::string msg = h ? gcr_last_error(h) : "gcr_create failed";
            if (h) gcr_destroy(h);
            throw DeviceError("gcr_b200: cannot create an engine on device " + std::to_string(device) + ": " + msg);
        }
        return h;
    }
    void release(int device, gcr_handle h)
    {
        std::lock_guard<std::mutex> lock(mu_);
        idle_[device].push_back(h);
    }

private:
    std::mutex mu_;
    std::map<int, std::vector<gcr_handle>> idle_;
};

EnginePool& pool()
{
    static EnginePool* p = new EnginePool(); // leaked on purpose: no CUDA calls during static destruction
    return *p;
}

struct Lease
{
    int device;
    gcr_handle h;
    explicit Lease(int d) : device(d), h(pool().acquire(d)) {}
    ~Lease() { pool().release(device, h); }
    void check(int rc, const char* what) const
    {
        if (rc == GCR_OK) return;
        const std::string msg = std::string("gcr_b200: ") + what + ": " + gcr_last_error(h);
        if (rc == GCR_ERR_INVALID) throw std::invalid_argument(msg);
        throw DeviceError(msg);
    }
};

// splitmix64: the host-side stream for the local optimisation's inner sampler
struct SplitMix
{
    uint64_t s;
    uint64_t next()
    {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    // uniform in [0, n)
    uint32_t below(uint32_t n) { return static_cast<uint32_t>(((next() >> 32) * static_cast<uint64_t>(n)) >> 32); }
};

// k distinct entries of `from`, uniformly (partial Fisher-Yates); the role of get_random_subset (GCRANSAC.h:53-80)
// (`scratch` holds a copy of the list the caller made once; the k swaps are undone before returning, so that the many draws from one
// list -- 50 trials per inner RANSAC, 12 000 entries at C3 -- do not copy it each time)
void drawSubset(size_t k, SplitMix& rng, std::vector<uint32_t>& scratch, std::vector<uint32_t>& out)
{
    static thread_local std::vector<size_t> swapped;
    swapped.resize(k);
    out.clear();
    for (size_t i = 0; i < k; ++i)
    {
        const size_t j = i + rng.below(static_cast<uint32_t>(scratch.size() - i));
        std::swap(scratch[i], scratch[j]);
        swapped[i] = j;
        out.push_back(scratch[i]);
    }
    for (size_t i = k; i-- > 0;) std::swap(scratch[i], scratch[swapped[i]]);
}

void wordsToIndices(const uint32_t* words, size_t n, std::vector<uint32_t>& out)
{
    out.clear();
    for (size_t w = 0; w < (n + 31) / 32; ++w)
    {
        uint32_t bits = words[w];
        while (bits)
        {
            const int b = __builtin_ctz(bits);
            bits &= bits - 1;
            const size_t i = 32 * w + b;
            if (i < n) out.push_back(static_cast<uint32_t>(i));
        }
    }
}

struct Best
{
    bool have = false;
    double score = 0.0;
    double model[4] = {0, 0, 0, 0};
    uint32_t n_inl[2] = {0, 0};
    std::vector<uint32_t> words; // packed masks, scale class then orientation class
};

} // namespace

size_t iterationNumber(const uint32_t n_inl[2], size_t ns, size_t no, double confidence, int problem)
{
    const double r0 = static_cast<double>(n_inl[0]) / static_cast<double>(ns);
    double q = 1.0;
    if (problem == GCR_PROBLEM_HYBRID)
    {
        const double r1 = static_cast<double>(n_inl[1]) / static_cast<double>(no);
        q *= std::pow(r0, 2.0);
        q *= std::pow(r1, 2.0);
    }
    else
        q *= std::pow(r0, 3.0); // one class, sample size 3
    const double lg = std::log(1.0 - q);
    if (std::fabs(lg) < std::numeric_limits<double>::epsilon()) return std::numeric_limits<size_t>::max();
    const double it = std::ceil(std::log(1.0 - confidence) / lg);
    if (!(it < 1.8e19)) return std::numeric_limits<size_t>::max();
    return it <= 0.0 ? 0 : static_cast<size_t>(it);
}

// The estimator for any of the three problems.  problem = GCR_PROBLEM_HYBRID: two classes with minimal samples {2, 2};
// scale-only problems: one class (no == 0, orient == nullptr), minimal sample 3.
static int estimateImpl(int problem, const double* scale, size_t ns, const double* orient, size_t no, double thr_s, double thr_o,
                        size_t min_iters, size_t max_iters, size_t max_lo, std::vector<bool>& scale_inliers,
                        std::vector<bool>& orientation_inliers, std::vector<double>& homography,
                        gcransac::SIFTRectifyingHomography& model, const Options& opt, Stats* stats_out)
{
    const bool hybrid = problem == GCR_PROBLEM_HYBRID;
    const size_t min_s = hybrid ? 2 : 3, min_o = hybrid ? 2 : 0; // sampleSize()
    const size_t lim_s = 7 * min_s, lim_o = 7 * min_o;            // inlierLimit(), rectifying_homography_estimator.h:45-55
    const auto t_begin = std::chrono::steady_clock::now();
    auto lap = [t = t_begin]() mutable {
        const auto now = std::chrono::steady_clock::now();
        const double ms = std::chrono::duration<double, std::milli>(now - t).count();
        t = now;
        return ms;
    };
    Stats st;
    if (!scale || (hybrid && !orient)) throw std::invalid_argument("null feature pointer");
    // GCRANSAC.h:222-228: fewer rows than the minimal sample of a class
    if (ns < min_s || no < min_o) throw std::runtime_error("Data set smaller than minimal sample size for corresponding data type");
    if (!hybrid) thr_o = 0.0;
    if (!(thr_s >= 0.0) || !(thr_o >= 0.0) || !std::isfinite(thr_s) || !std::isfinite(thr_o))
        throw std::invalid_argument("residual thresholds must be finite and non-negative");
    const double thr[2] = {thr_s, thr_o};
    const size_t ws = (ns + 31) / 32, wo = (no + 31) / 32, W = ws + wo;

    st.seed = opt.seed;
    if (!opt.has_seed)
    {
        std::random_device rd; // the reference seeds every draw from random_device (GCRANSAC.h:67-70)
        st.seed = (static_cast<uint64_t>(rd()) << 32) | rd();
    }

    Lease eng(opt.device);
    eng.check(gcr_set_problem(eng.h, problem), "gcr_set_problem");
    // device tensors: the pooled engine runs on its own non-blocking stream; the kernels that wrote the features must come first
    if (opt.features_on_device && opt.wait_producer) eng.check(gcr_wait_stream(eng.h, opt.producer_stream), "gcr_wait_stream");
    eng.check(gcr_set_features(eng.h, scale, static_cast<int64_t>(ns), orient, static_cast<int64_t>(no),
                               opt.features_on_device ? 1 : 0),
              "gcr_set_features");

    // ---- hypothesise / verify: while (iter < min_iters || iter < min(max_iteration, max_iters)), GCRANSAC.h:286-287,
    //      in device batches; every drawn sample is one iteration (S13)
    st.setup_ms = lap();
    Best best;
    best.words.assign(W, 0u);
    size_t done = 0;
    const uint32_t one_each[2] = {1u, 1u}; // GCRANSAC.h:233-236: the initial max_iteration assumes one inlier per class
    size_t target = opt.samples ? static_cast<size_t>(opt.n_samples)
                                : std::max(min_iters, std::min(iterationNumber(one_each, ns, no, opt.confidence, problem), max_iters));
    while (done < target)
    {
        // The reference re-evaluates its stopping rule after every iteration; here it is evaluated between device
        // batches: the first batch covers min_iters (at least `batch_quantum` slots), later ones double.
        size_t n = std::min(target - done, opt.max_batch);
        if (!opt.samples) n = std::min(n, std::max(done == 0 ? min_iters : done, opt.batch_quantum));
        gcr_result res;
        eng.check(gcr_run(eng.h, static_cast<int64_t>(n), st.seed, done, opt.samples ? opt.samples + 4 * done : nullptr, thr,
                          stats_out ? GCR_RUN_TIME_KERNELS : 0u, nullptr, &res), // device timings only when the caller asked for statistics
                  "gcr_run");
        st.device_calls++;
        st.n_sample_valid += res.n_sample_valid;
        st.n_solved += res.n_solved;
        st.n_rechecked += res.n_rechecked;
        st.n_exact_hyp += res.n_exact_hyp;
        st.hypothesis_stage_ms += res.device_ms;
        st.score_kernel_ms += res.score_kernel_ms;
        if (res.best_index >= 0 && res.best_score > best.score) // strict: the earliest batch wins ties (S11)
        {
            best.have = true;
            best.score = res.best_score;
            std::memcpy(best.model, res.best_model, sizeof(best.model));
            best.n_inl[0] = res.best_n_inl[0];
            best.n_inl[1] = res.best_n_inl[1];
            st.best_index = res.best_index;
            eng.check(gcr_best(eng.h, nullptr, nullptr, nullptr, best.words.data(), best.words.data() + ws, nullptr), "gcr_best");
        }
        done += n;
        if (!opt.samples)
        {
            // max_iteration = getIterationNumber(best inlier counts), GCRANSAC.h:480-483
            const size_t adaptive = iterationNumber(best.have ? best.n_inl : one_each, ns, no, opt.confidence, problem);
            target = std::max(min_iters, std::min(adaptive, max_iters));
        }
    }
    st.iteration_number = static_cast<int64_t>(done);
    st.best_raw_score = best.score;
    st.loop_ms = lap();

    scale_inliers.assign(ns, false);
    orientation_inliers.assign(no, false);
    model = gcransac::SIFTRectifyingHomography();
    auto finish = [&](int n_total) {
        const auto Hm = model.getHomography();
        homography.assign(Hm.begin(), Hm.end());
        st.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
        if (stats_out) *stats_out = st;
        return n_total;
    };
    // GCRANSAC.h:535-541: a best model supported by no more than a minimal sample per class is "not found"
    if (!best.have || (best.n_inl[0] <= min_s && best.n_inl[1] <= min_o)) return finish(0);

    // ---- local optimisation on the selected model (graphCutLocalOptimization for two classes, GCRANSAC.h:873-1062)
    SplitMix rng{st.seed ^ 0x6C6F63616C6F7074ull};
    std::vector<uint32_t> lab_words(W), in_s, in_o, smp_s, smp_o, scratch_s, scratch_o;
    std::vector<double> trial_scores;
    std::vector<uint32_t> trial_cnt, trial_words;
    // Non-minimal fits run on the device (gcr_fit_models), all trials of one inner RANSAC in one pass
    std::vector<uint32_t> fit_idx_s, fit_idx_o;
    std::vector<int64_t> fit_off_s, fit_off_o;
    std::vector<double> fit_models;
    std::vector<int32_t> fit_ok;
    auto fitBegin = [&] {
        fit_idx_s.clear();
        fit_idx_o.clear();
        fit_off_s.assign(1, 0);
        fit_off_o.assign(1, 0);
    };
    auto fitAdd = [&](const std::vector<uint32_t>& is, const std::vector<uint32_t>& io) {
        fit_idx_s.insert(fit_idx_s.end(), is.begin(), is.end());
        fit_off_s.push_back(static_cast<int64_t>(fit_idx_s.size()));
        if (hybrid)
        {
            fit_idx_o.insert(fit_idx_o.end(), io.begin(), io.end());
            fit_off_o.push_back(static_cast<int64_t>(fit_idx_o.size()));
        }
    };
    auto fitScoreRun = [&](bool want_masks) {
        const size_t k = fit_off_s.size() - 1;
        fit_models.resize(4 * k);
        fit_ok.resize(k);
        trial_scores.resize(k);
        trial_cnt.resize(2 * k);
        if (want_masks) trial_words.resize(k * W);
        if (fit_idx_s.empty()) fit_idx_s.push_back(0u);
        if (hybrid && fit_idx_o.empty()) fit_idx_o.push_back(0u);
        eng.check(gcr_fit_score_models(eng.h, static_cast<int64_t>(k), fit_idx_s.data(), fit_off_s.data(), hybrid ? fit_idx_o.data() : nullptr,
                                       hybrid ? fit_off_o.data() : nullptr, thr, 0, fit_models.data(), fit_ok.data(), trial_scores.data(),
                                       trial_cnt.data(), want_masks ? trial_words.data() : nullptr),
                  "gcr_fit_score_models");
        return k;
    };
    // two classes: r^2 <= (1.5 t)^2 (GCRANSAC.h:931-937); one class: graph cut over the (empty) neighbourhood graph, which
    // labels a feature an inlier iff r^2 < (1.5 t)^2 for any 0 <= lambda < 1 (GCRANSAC.h:759-870)
    const uint32_t label_flags = hybrid ? GCR_SCORE_TRUNCATED_SQ : (GCR_SCORE_TRUNCATED_SQ | GCR_SCORE_STRICT);

    if (opt.do_local_optimization && max_lo > 0)
    {
        Best lo = best; // max_score / best_model / best_inliers of the reference
        // One launch sequence and one synchronisation per pass (gcr_lo_round: labels, this generator's sub-samples drawn on the device,
        // fits, scores, the strict-< update) where the sub-sample sizes allow; the two-call form below otherwise, and for a pass whose
        // device-side histogram tie order disagreed with this host's std::unordered_map (never with libstdc++).
        const bool device_rounds = lim_s <= 32 && (!hybrid || lim_o <= 32) && max_lo <= 1024;
        while (static_cast<size_t>(++st.graph_cut_number) < opt.max_graph_cut_number)
        {
            bool updated = false;
            if (device_rounds)
            {
                gcr_lo_round_io io;
                std::memset(&io, 0, sizeof(io));
                std::memcpy(io.model, lo.model, sizeof(io.model));
                io.score = lo.score;
                io.rng_state = rng.s;
                const int32_t lim[2] = {static_cast<int32_t>(lim_s), static_cast<int32_t>(lim_o)};
                const int32_t mins[2] = {static_cast<int32_t>(min_s), static_cast<int32_t>(min_o)};
                lab_words.resize(W);
                eng.check(gcr_lo_round(eng.h, &io, thr, label_flags, static_cast<int32_t>(max_lo), lim, mins, lab_words.data()), "gcr_lo_round");
                if (!io.phi_mismatch)
                {
                    if (!io.sampled) break; // sampling fails: neither a draw nor "use all" applies
                    rng.s = io.rng_state;
                    st.lo_models_scored += io.n_models_scored;
                    if (!io.updated) break;
                    lo.score = io.score;
                    std::memcpy(lo.model, io.model, sizeof(lo.model));
                    lo.n_inl[0] = io.n_inl[0];
                    lo.n_inl[1] = io.n_inl[1];
                    lo.words.assign(lab_words.begin(), lab_words.end());
                    continue;
                }
            }
            // relabel every feature with the truncated threshold r^2 <= (1.5 t)^2   (GCRANSAC.h:931-937)
            eng.check(gcr_score_models(eng.h, lo.model, 1, thr, label_flags, nullptr, nullptr, lab_words.data()),
                      "gcr_score_models(label)");
            wordsToIndices(lab_words.data(), ns, in_s);
            wordsToIndices(lab_words.data() + ws, no, in_o);
            // inner RANSAC on the labelled inliers: min(7 x sample size, n) per class  (GCRANSAC.h:955-996)
            const size_t k_s = std::min(lim_s, in_s.size()), k_o = std::min(lim_o, in_o.size());
            if (in_s.size() <= min_s || (hybrid && in_o.size() <= min_o)) break; // sampling fails: neither a draw nor "use all" applies
            fitBegin();
            scratch_s = in_s;
            scratch_o = in_o;
            for (size_t trial = 0; trial < std::min<size_t>(max_lo, 1024); ++trial) // one device pass holds 1024 fits / models
            {
                if (k_s < in_s.size()) drawSubset(k_s, rng, scratch_s, smp_s); else smp_s = in_s;
                if (k_o < in_o.size()) drawSubset(k_o, rng, scratch_o, smp_o); else smp_o = in_o;
                fitAdd(smp_s, smp_o);
            }
            // all trial fits AND their scores in one device pass (gcr_fit_score_models): fit k's model goes through getScore on the device
            // behind its fit, one synchronisation per inner RANSAC
            const size_t n_fit = fitScoreRun(true);
            for (size_t t = 0; t < n_fit; ++t)
            {
                if (!fit_ok[t]) continue;
                st.lo_models_scored += 1;
                if (lo.score < trial_scores[t]) // strict, in trial order (GCRANSAC.h:1034)
                {
                    updated = true;
                    lo.score = trial_scores[t];
                    std::memcpy(lo.model, &fit_models[4 * t], sizeof(lo.model));
                    lo.n_inl[0] = trial_cnt[2 * t];
                    lo.n_inl[1] = trial_cnt[2 * t + 1];
                    lo.words.assign(trial_words.begin() + t * W, trial_words.begin() + (t + 1) * W);
                }
            }
            if (!updated) break;
        }
        if (best.score < lo.score) best = lo;
    }
    st.lo_score = best.score;
    st.lo_ms = lap();

    // ---- final model: one least-squares refit on all inliers, kept if it scores higher (GCRANSAC.h:628-667; the
    //      iterated re-weighted fit before it never produces a model in the reference)
    wordsToIndices(best.words.data(), ns, in_s);
    wordsToIndices(best.words.data() + ws, no, in_o);
    {
        fitBegin();
        fitAdd(in_s, in_o);
        fitScoreRun(true); // the refit and its score in one device pass
        if (fit_ok[0] && best.score < trial_scores[0])
        {
            st.refit_accepted = true;
            best.score = trial_scores[0];
            std::memcpy(best.model, fit_models.data(), sizeof(best.model));
            best.n_inl[0] = trial_cnt[0];
            best.n_inl[1] = trial_cnt[1];
            best.words.assign(trial_words.begin(), trial_words.begin() + W);
        }
    }
    st.final_score = best.score;
    st.refit_ms = lap();

    model.h7 = best.model[0];
    model.h8 = best.model[1];
    model.alpha = best.model[2];
    model.phi = best.model[3];
    wordsToIndices(best.words.data(), ns, in_s);
    wordsToIndices(best.words.data() + ws, no, in_o);
    for (uint32_t i : in_s) scale_inliers[i] = true;
    for (uint32_t i : in_o) orientation_inliers[i] = true;
    return finish(static_cast<int>(in_s.size() + in_o.size()));
}

int findRectifyingHomographySIFT(const double* scale, size_t ns, const double* orient, size_t no, double thr_s, double thr_o,
                                 double /*spatial_coherence_weight: unused by the two-class path, GCRANSAC.h:922-944*/,
                                 size_t min_iters, size_t max_iters, size_t max_lo, std::vector<bool>& scale_inliers,
                                 std::vector<bool>& orientation_inliers, std::vector<double>& homography,
                                 gcransac::SIFTRectifyingHomography& model, const Options& opt, Stats* stats_out)
{
    return estimateImpl(GCR_PROBLEM_HYBRID, scale, ns, orient, no, thr_s, thr_o, min_iters, max_iters, max_lo, scale_inliers,
                        orientation_inliers, homography, model, opt, stats_out);
}

int findRectifyingHomographyScaleOnly(const double* features, size_t n, double thr, double spatial_coherence_weight, size_t min_iters,
                                      size_t max_iters, size_t max_lo, bool original, std::vector<bool>& inliers,
                                      std::vector<double>& homography, gcransac::ScaleBasedRectifyingHomography& model,
                                      const Options& opt, Stats* stats_out)
{
    // the neighbourhood graph of the reference's driver is empty (gcransac_python.cpp:62-70), so the spatial coherence
    // weight only rescales the unary terms of the graph cut; outside [0, 1) the reference's labelling degenerates
    if (!(spatial_coherence_weight >= 0.0 && spatial_coherence_weight < 1.0))
        throw std::invalid_argument("spatial_coherence_weight must be in [0, 1)");
    std::vector<bool> unused;
    gcransac::SIFTRectifyingHomography full;
    const int total = estimateImpl(original ? GCR_PROBLEM_SCALE_ONLY_ORIGINAL : GCR_PROBLEM_SCALE_ONLY, features, n, nullptr, 0, thr,
                                   0.0, min_iters, max_iters, max_lo, inliers, unused, homography, full, opt, stats_out);
    model = gcransac::ScaleBasedRectifyingHomography();
    model.h7 = full.h7;
    model.h8 = full.h8;
    if (total != 0) model.alpha = full.alpha;
    return total;
}

} // namespace gcr_host

namespace
{
int scaleOnlyEntry(const std::vector<double>& features, double thr, double scw, size_t min_iters, size_t max_iters, size_t max_lo,
                   bool original, std::vector<bool>& inliers, std::vector<double>& homography,
                   gcransac::ScaleBasedRectifyingHomography& model, unsigned int verbose_level)
{
    if (features.size() % 3 != 0)
    {
        std::fprintf(stderr, "Feature vector size (%zu) is not a multiple of the feature size (3).\n", features.size());
        return 0;
    }
    const int n = gcr_host::findRectifyingHomographyScaleOnly(features.data(), features.size() / 3, thr, scw, min_iters, max_iters,
                                                              max_lo, original, inliers, homography, model, gcr_host::Options(),
                                                              nullptr);
    if (verbose_level > 0)
        std::printf("\nFinal model:\n\nh7: %g, h8: %g, alpha: %g\n\nNumber of scale-inliers: %d\n", model.h7, model.h8, model.alpha, n);
    return n;
}
} // namespace

// gcransac_python.h:7-31, gcransac_python.cpp:32-254
int findRectifyingHomographyScaleOnly_(const std::vector<double>& features, double scale_residual_thresh,
                                       double spatial_coherence_weight, size_t min_iteration_number, size_t max_iteration_number,
                                       size_t max_local_optimization_number, std::vector<bool>& inliers,
                                       std::vector<double>& homography, gcransac::ScaleBasedRectifyingHomography& model,
                                       unsigned int verbose_level)
{
    return scaleOnlyEntry(features, scale_residual_thresh, spatial_coherence_weight, min_iteration_number, max_iteration_number,
                          max_local_optimization_number, false, inliers, homography, model, verbose_level);
}

int findRectifyingHomographyScaleOnlyOriginal_(const std::vector<double>& features, double scale_residual_thresh,
                                               double spatial_coherence_weight, size_t min_iteration_number,
                                               size_t max_iteration_number, size_t max_local_optimization_number,
                                               std::vector<bool>& inliers, std::vector<double>& homography,
                                               gcransac::ScaleBasedRectifyingHomography& model, unsigned int verbose_level)
{
    return scaleOnlyEntry(features, scale_residual_thresh, spatial_coherence_weight, min_iteration_number, max_iteration_number,
                          max_local_optimization_number, true, inliers, homography, model, verbose_level);
}

int findRectifyingHomographySIFT_(const std::vector<double>& scale_features, const std::vector<double>& orientation_features,
                                  double scale_residual_thresh, double orientation_residual_thresh,
                                  double spatial_coherence_weight, size_t min_iteration_number, size_t max_iteration_number,
                                  size_t max_local_optimization_number, std::vector<bool>& scale_inliers,
                                  std::vector<bool>& orientation_inliers, std::vector<double>& homography,
                                  gcransac::SIFTRectifyingHomography& model, unsigned int verbose_level)
{
    // gcransac_python.cpp:18-30,277-282: a flat vector whose length is not a multiple of 3 is reported and yields 0
    for (const auto* v : {&scale_features, &orientation_features})
        if (v->size() % 3 != 0)
        {
            std::fprintf(stderr, "Feature vector size (%zu) is not a multiple of the feature size (3).\n", v->size());
            return 0;
        }
    gcr_host::Stats st;
    const int n = gcr_host::findRectifyingHomographySIFT(
        scale_features.data(), scale_features.size() / 3, orientation_features.data(), orientation_features.size() / 3,
        scale_residual_thresh, orientation_residual_thresh, spatial_coherence_weight, min_iteration_number,
        max_iteration_number, max_local_optimization_number, scale_inliers, orientation_inliers, homography, model,
        gcr_host::Options(), &st);
    if (verbose_level > 0)
        std::printf("\nFinal model:\n\nh7: %g, h8: %g, alpha: %g\n\nx0: %g, y0: %g, s: %g\n\nRectified image vanishing point "
                    "direction: %g\nNumber of inliers: %d\nIterations: %lld (device %.3f ms)\n",
                    model.h7, model.h8, model.alpha, model.x0, model.y0, model.s, model.phi * 180.0 / 3.14159265358979323846, n,
                    static_cast<long long>(st.iteration_number), st.hypothesis_stage_ms);
    return n;
}
