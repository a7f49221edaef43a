// Host-side estimator of the B200 engine: the C++ surface the reference exposes under its Python module
// (src/pygcransac/include/gcransac_python.h:33-47, src/gcransac_python.cpp:256-406), re-hosted on the C ABI
// of include/gcr_b200.h.  The hypothesise/verify loop (GCRANSAC.h:286-485) is one batched device call per
// block of iterations; local optimisation (GCRANSAC.h:873-1062) and the final refit (GCRANSAC.h:628-667) run
// once on the GPU-selected best model, their model fits on the host (gcr_refit.hpp) and every rescoring
// on the device (gcr_score_models).
#pragma once
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "gcr_model.hpp"

namespace gcr_host
{

struct Options
{
    int device = 0;
    bool has_seed = false;       // false: seed drawn from std::random_device, like the reference's sampler
    uint64_t seed = 0;           // Philox key of the device sampler and of the local-optimisation sampler
    const int32_t* samples = nullptr; // replay mode: int32[n_samples,4] = (scale0, scale1, orient0, orient1), host memory
    int64_t n_samples = 0;            // ... then exactly n_samples iterations are run
    bool features_on_device = false;  // scale/orient pointers are DEVICE pointers on `device` (zero-copy tensors)
    bool wait_producer = false;       // ... written by work on `producer_stream`: the engine's stream is ordered after it (gcr_wait_stream)
    void* producer_stream = nullptr;  //     (a cudaStream_t; nullptr = the legacy default stream)
    double confidence = 0.95;         // settings.h:60
    size_t max_graph_cut_number = 10; // settings.h:57
    size_t max_batch = size_t(1) << 22; // hypothesis slots per device call
    size_t batch_quantum = 8192;        // smallest batch between two evaluations of the adaptive stopping rule
    bool do_local_optimization = true;  // gcransac_python.cpp:315
};

struct Stats
{
    uint64_t seed = 0;
    int64_t iteration_number = 0;   // hypothesis slots evaluated (the reference's statistics.iteration_number)
    int64_t n_sample_valid = 0, n_solved = 0, n_rechecked = 0, n_exact_hyp = 0;
    int64_t best_index = -1;        // global index of the best raw hypothesis
    double best_raw_score = 0.0;    // its MSAC score
    double lo_score = 0.0;          // after local optimisation
    double final_score = 0.0;       // score of the returned model
    int graph_cut_number = 0, lo_models_scored = 0;
    bool refit_accepted = false;
    double hypothesis_stage_ms = 0.0; // CUDA-event time of the device hypothesis stage, summed over batches
    double score_kernel_ms = 0.0;
    double total_ms = 0.0;            // wall time of the whole call
    double setup_ms = 0.0, loop_ms = 0.0, lo_ms = 0.0, refit_ms = 0.0; // wall time of: engine lease + feature upload | hypothesis loop | local optimisation | final refit
    int device_calls = 0;
};

// Thrown for argument errors (the binding lets pybind11 map it to ValueError, as the reference's
// std::invalid_argument, bindings.cpp:199-224) and for device failures (RuntimeError).
struct DeviceError : public std::runtime_error
{
    explicit DeviceError(const std::string& what) : std::runtime_error(what) {}
};

// Pointer form.  scale: f64[ns,3] rows (x, y, s); orient: f64[no,3] rows (x, y, theta).  Outputs as the reference:
// boolean inlier masks, the 9 row-major homography entries, the model.  Returns the total inlier count
// (0 <=> no model, gcransac_python.cpp:404).
int findRectifyingHomographySIFT(const double* scale, size_t ns, const double* orient, size_t no, double scale_residual_thresh,
                                 double orientation_residual_thresh, double spatial_coherence_weight,
                                 size_t min_iteration_number, size_t max_iteration_number,
                                 size_t max_local_optimization_number, std::vector<bool>& scale_inliers,
                                 std::vector<bool>& orientation_inliers, std::vector<double>& homography,
                                 gcransac::SIFTRectifyingHomography& model, const Options& options, Stats* stats);

// Scale-only estimators (gcransac_python.cpp:32-254): features f64[n,3] rows (x, y, s); `original` selects
// findRectifyingHomographyScaleOnlyOriginal.  In replay mode Options::samples rows are (s0, s1, s2, ignored).
int findRectifyingHomographyScaleOnly(const double* features, size_t n, double scale_residual_thresh, double spatial_coherence_weight,
                                      size_t min_iteration_number, size_t max_iteration_number,
                                      size_t max_local_optimization_number, bool original, std::vector<bool>& inliers,
                                      std::vector<double>& homography, gcransac::ScaleBasedRectifyingHomography& model,
                                      const Options& options, Stats* stats);

// GCRANSAC::getIterationNumber (GCRANSAC.h:738-757): sample sizes {2, 2} for the hybrid problem (problem 0), {3} for the
// scale-only problems
size_t iterationNumber(const uint32_t n_inl[2], size_t ns, size_t no, double confidence, int problem = 0);

} // namespace gcr_host

// The reference's own entry points, same signatures and meaning (gcransac_python.h:7-47).
int findRectifyingHomographyScaleOnly_(const std::vector<double>& features, double scale_residual_thresh,
                                       double spatial_coherence_weight, size_t min_iteration_number, size_t max_iteration_number,
                                       size_t max_local_optimization_number, std::vector<bool>& inliers,
                                       std::vector<double>& homography, gcransac::ScaleBasedRectifyingHomography& model,
                                       unsigned int verbose_level = 0);
int findRectifyingHomographyScaleOnlyOriginal_(const std::vector<double>& features, double scale_residual_thresh,
                                               double spatial_coherence_weight, size_t min_iteration_number,
                                               size_t max_iteration_number, size_t max_local_optimization_number,
                                               std::vector<bool>& inliers, std::vector<double>& homography,
                                               gcransac::ScaleBasedRectifyingHomography& model, unsigned int verbose_level = 0);
int findRectifyingHomographySIFT_(const std::vector<double>& scale_features, const std::vector<double>& orientation_features,
                                  double scale_residual_thresh, double orientation_residual_thresh,
                                  double spatial_coherence_weight, size_t min_iteration_number, size_t max_iteration_number,
                                  size_t max_local_optimization_number, std::vector<bool>& scale_inliers,
                                  std::vector<bool>& orientation_inliers, std::vector<double>& homography,
                                  gcransac::SIFTRectifyingHomography& model, unsigned int verbose_level = 0);
