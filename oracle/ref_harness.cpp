// TEST INFRASTRUCTURE ONLY (oracle/).  Never linked into, imported by or executed from the
// product path; only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs use it.
//
// Builds the UNMODIFIED reference (yuvalnis/graph-cut-ransac, sources read where they lie under
// /root/reference) into oracle/_ref/libgcr_ref.so against the storage-only Eigen/OpenCV stand-ins
// in oracle/shim/ (both libraries are un-vendored, absent from this image).
// Everything numerical below is the reference's own code: this file only marshals plain C arrays
// in and out and walks the reference's call sequence of GCRANSAC.h:304-446 per replayed sample.
//
// The reference translation unit is included as a whole because its headers define non-inline
// functions (GCRANSAC.h:53 get_random_subset, two_sift.hpp:141 ...), so it can only live in one TU.
#include "src/gcransac_python.cpp"

#include <cstdint>
#include <chrono>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace
{
using Estimator = gcransac::utils::SIFTBasedRectifyingHomographyEstimator;
using Scoring = gcransac::MSACScoringFunction<Estimator>;
using DataType = Estimator::DataType;
using Inliers = Estimator::InlierContainerType;
using Residual = Estimator::ResidualType;

std::unique_ptr<const cv::Mat> make_mat(const double* src, long n)
{
    auto m = std::make_unique<cv::Mat>(static_cast<int>(n), 3, CV_64F);
    if (n > 0) std::memcpy(m->data, src, sizeof(double) * 3 * static_cast<size_t>(n));
    return std::unique_ptr<const cv::Mat>(std::move(m));
}
} // namespace

namespace
{
constexpr uint8_t kRefSampleValid = 1, kRefSolved = 2, kRefModelValid = 4;

// hypothesis stage of a one-class estimator (three_sift / three_sift_original), see gcr_ref_replay_scale
template <class Est>
int replay_scale(const double* scale, long ns, const int32_t* samples, long H, double thr, uint8_t* flags, double* models,
                 double* score, uint32_t* n_inl, uint32_t* masks, int nthreads)
{
    using Sc = gcransac::MSACScoringFunction<Est>;
    typename Est::DataType data{make_mat(scale, ns)};
    const Est estimator;
    const Sc scoring;
    gcransac::utils::Settings settings(1);
    settings.threshold(0) = thr;
    const typename Est::ResidualType truncated = 1.5 * settings.threshold;
    const long ws = (ns + 31) / 32;
    int used = 1;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
    used = nthreads > 0 ? nthreads : omp_get_max_threads();
#pragma omp parallel
#endif
    {
        typename Est::InlierContainerType sample{}, inl{};
        std::vector<typename Est::Model> found;
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 16)
#endif
        for (long h = 0; h < H; ++h)
        {
            uint8_t f = 0;
            double m[4] = {0, 0, 0, 0}, val = 0.0;
            uint32_t cnt = 0;
            if (masks) std::memset(masks + h * ws, 0, sizeof(uint32_t) * ws);
            sample[0] = {static_cast<size_t>(samples[4 * h]), static_cast<size_t>(samples[4 * h + 1]),
                         static_cast<size_t>(samples[4 * h + 2])};
            found.clear();
            if (estimator.isValidSample(data, sample))
            {
                f |= kRefSampleValid;
                if (estimator.estimateModel(data, sample, found) && !found.empty())
                {
                    f |= kRefSolved;
                    auto& model = found[0];
                    m[0] = model.h7; m[1] = model.h8; m[2] = model.alpha;
                    const auto s = scoring.getScore(data, model, estimator, settings.threshold, inl);
                    val = s.value();
                    cnt = static_cast<uint32_t>(s.num_inliers_by_type(0));
                    bool updated = false;
                    if (estimator.isValidModel(model, data, inl, sample, truncated, updated)) f |= kRefModelValid;
                    if (masks) for (size_t i : inl[0]) masks[h * ws + i / 32] |= 1u << (i % 32);
                }
            }
            flags[h] = f;
            for (int k = 0; k < 4; ++k) models[4 * h + k] = m[k];
            score[h] = val;
            n_inl[2 * h] = cnt;
            n_inl[2 * h + 1] = 0;
        }
    }
    return used;
}
} // namespace


extern "C"
{

// flag bits per hypothesis
enum { GCR_REF_SAMPLE_VALID = 1, GCR_REF_SOLVED = 2, GCR_REF_MODEL_VALID = 4 };

// Hypothesis stage of the reference, per replayed minimal sample, in the order of
// GCRANSAC.h:304-446: isValidSample -> estimateModel -> getScore -> isValidModel.
//   samples  int32[H,4] = (scale0, scale1, orient0, orient1)
//   flags    u8[H]; models f64[H,4] = (h7,h8,alpha,phi); score f64[H]; n_inl u32[H,2]
//   masks    nullable u32[H, ws+wo], ws=ceil(ns/32), wo=ceil(no/32); bit j of word w <-> feature 32w+j,
//            built from the inlier index lists getScore fills (NOT cleared by the zero-reset, S3)
// Returns the number of threads used.
int gcr_ref_replay(const double* scale, long ns, const double* orient, long no,
                   const int32_t* samples, long H, const double* thr,
                   uint8_t* flags, double* models, double* score, uint32_t* n_inl,
                   uint32_t* masks, int nthreads)
{
    DataType data{make_mat(scale, ns), make_mat(orient, no)};
    const Estimator estimator;
    const Scoring scoring;
    gcransac::utils::Settings settings(2);
    settings.threshold(0) = thr[0];
    settings.threshold(1) = thr[1];
    const Residual truncated = 1.5 * settings.threshold; // GCRANSAC.h:207
    const long ws = (ns + 31) / 32, wo = (no + 31) / 32, wt = ws + wo;
    int used = 1;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
    used = nthreads > 0 ? nthreads : omp_get_max_threads();
#pragma omp parallel
#endif
    {
        Inliers sample{}, inl{};
        inl[0].reserve(ns);
        inl[1].reserve(no);
        std::vector<gcransac::SIFTRectifyingHomography> found;
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 16)
#endif
        for (long h = 0; h < H; ++h)
        {
            uint8_t f = 0;
            double m[4] = {0, 0, 0, 0};
            double val = 0.0;
            uint32_t cnt[2] = {0, 0};
            if (masks) std::memset(masks + h * wt, 0, sizeof(uint32_t) * wt);
            sample[0] = {static_cast<size_t>(samples[4 * h + 0]), static_cast<size_t>(samples[4 * h + 1])};
            sample[1] = {static_cast<size_t>(samples[4 * h + 2]), static_cast<size_t>(samples[4 * h + 3])};
            found.clear();
            if (estimator.isValidSample(data, sample))
            {
                f |= GCR_REF_SAMPLE_VALID;
                if (estimator.estimateModel(data, sample, found) && !found.empty())
                {
                    f |= GCR_REF_SOLVED;
                    auto& model = found[0];
                    m[0] = model.h7; m[1] = model.h8; m[2] = model.alpha; m[3] = model.phi;
                    const auto s = scoring.getScore(data, model, estimator, settings.threshold, inl);
                    val = s.value();
                    cnt[0] = static_cast<uint32_t>(s.num_inliers_by_type(0));
                    cnt[1] = static_cast<uint32_t>(s.num_inliers_by_type(1));
                    bool updated = false;
                    if (estimator.isValidModel(model, data, inl, sample, truncated, updated))
                        f |= GCR_REF_MODEL_VALID;
                    if (masks)
                    {
                        for (size_t i : inl[0]) masks[h * wt + i / 32] |= 1u << (i % 32);
                        for (size_t i : inl[1]) masks[h * wt + ws + i / 32] |= 1u << (i % 32);
                    }
                }
            }
            flags[h] = f;
            for (int k = 0; k < 4; ++k) models[4 * h + k] = m[k];
            score[h] = val;
            n_inl[2 * h] = cnt[0];
            n_inl[2 * h + 1] = cnt[1];
        }
    }
    return used;
}

// Score one given model with the reference's MSAC (MSAC_scoring_function.hpp:53-130).
void gcr_ref_score_model(const double* scale, long ns, const double* orient, long no,
                         const double* model4, const double* thr,
                         double* value, uint32_t* n_inl, uint8_t* smask, uint8_t* omask)
{
    DataType data{make_mat(scale, ns), make_mat(orient, no)};
    const Estimator estimator;
    const Scoring scoring;
    gcransac::utils::Settings settings(2);
    settings.threshold(0) = thr[0];
    settings.threshold(1) = thr[1];
    gcransac::SIFTRectifyingHomography model;
    model.h7 = model4[0]; model.h8 = model4[1]; model.alpha = model4[2]; model.phi = model4[3];
    Inliers inl{};
    const auto s = scoring.getScore(data, model, estimator, settings.threshold, inl);
    *value = s.value();
    n_inl[0] = static_cast<uint32_t>(s.num_inliers_by_type(0));
    n_inl[1] = static_cast<uint32_t>(s.num_inliers_by_type(1));
    if (smask) { std::memset(smask, 0, ns); for (size_t i : inl[0]) smask[i] = 1; }
    if (omask) { std::memset(omask, 0, no); for (size_t i : inl[1]) omask[i] = 1; }
}

// Per-feature residuals of the reference solver (two_sift.hpp:621-665).
void gcr_ref_residuals(const double* scale, long ns, const double* orient, long no,
                       const double* model4, double* rs, double* ro)
{
    gcransac::SIFTRectifyingHomography model;
    model.h7 = model4[0]; model.h8 = model4[1]; model.alpha = model4[2]; model.phi = model4[3];
    using Solver = gcransac::estimator::solver::RectifyingHomographyTwoSIFTSolver;
    for (long i = 0; i < ns; ++i)
        rs[i] = Solver::scaleResidual(scale[3 * i], scale[3 * i + 1], scale[3 * i + 2], model);
    for (long i = 0; i < no; ++i)
        ro[i] = Solver::orientationResidual(orient[3 * i], orient[3 * i + 1], orient[3 * i + 2], model);
}

// The reference's full driver (gcransac_python.cpp:256-406), unmodified: its own random_device
// sampler, local optimisation and final refit.  model7 = (x0,y0,s,h7,h8,alpha,phi).
// Returns the total inlier count; *seconds = wall time of the call.
int gcr_ref_find_sift(const double* scale, long ns, const double* orient, long no,
                      double thr_s, double thr_o, double spatial_coherence_weight,
                      unsigned long min_iters, unsigned long max_iters, unsigned long max_lo,
                      uint8_t* smask, uint8_t* omask, double* H9, double* model7, double* seconds)
{
    std::vector<double> sv(scale, scale + 3 * ns), ov(orient, orient + 3 * no);
    std::vector<bool> sm, om;
    std::vector<double> Hh;
    gcransac::SIFTRectifyingHomography model;
    const auto t0 = std::chrono::steady_clock::now();
    const int n = findRectifyingHomographySIFT_(sv, ov, thr_s, thr_o, spatial_coherence_weight,
                                                min_iters, max_iters, max_lo, sm, om, Hh, model, 0);
    const auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    for (long i = 0; i < ns && i < static_cast<long>(sm.size()); ++i) smask[i] = sm[i] ? 1 : 0;
    for (long i = 0; i < no && i < static_cast<long>(om.size()); ++i) omask[i] = om[i] ? 1 : 0;
    for (size_t i = 0; i < 9 && i < Hh.size(); ++i) H9[i] = Hh[i];
    model7[0] = model.x0; model7[1] = model.y0; model7[2] = model.s;
    model7[3] = model.h7; model7[4] = model.h8; model7[5] = model.alpha; model7[6] = model.phi;
    return n;
}

// The reference's non-minimal fit from ALL rows given (estimateModelNonminimal,
// rectifying_homography_estimator.h:164-227 -> two_sift.hpp:423-612); note that its linear solve runs through the
// shim's column-pivoted QR (oracle/shim/Eigen), the row construction and the phi histogram are the reference's own.
// Returns 1 and model7 = (x0,y0,s,h7,h8,alpha,phi) on success.
int gcr_ref_fit_nonminimal(const double* scale, long ns, const double* orient, long no, double* model7)
{
    DataType data{make_mat(scale, ns), make_mat(orient, no)};
    const Estimator estimator;
    Inliers inl{};
    for (long i = 0; i < ns; ++i) inl[0].push_back(static_cast<size_t>(i));
    for (long i = 0; i < no; ++i) inl[1].push_back(static_cast<size_t>(i));
    std::vector<gcransac::SIFTRectifyingHomography> models;
    if (!estimator.estimateModelNonminimal(data, inl, models) || models.empty()) return 0;
    const auto& m = models[0];
    model7[0] = m.x0; model7[1] = m.y0; model7[2] = m.s;
    model7[3] = m.h7; model7[4] = m.h8; model7[5] = m.alpha; model7[6] = m.phi;
    return 1;
}

// ---- scale-only estimators (gcransac_python.cpp:32-254; three_sift.hpp / three_sift_original.hpp) ----
// Hypothesis stage of the scale-only estimators per replayed 3-sample (s0, s1, s2, ignored); outputs as gcr_ref_replay.
int gcr_ref_replay_scale(const double* scale, long ns, const int32_t* samples, long H, double thr, int original,
                         uint8_t* flags, double* models, double* score, uint32_t* n_inl, uint32_t* masks, int nthreads)
{
    if (original)
        return replay_scale<gcransac::utils::OriginalScaleBasedRectifyingHomographyEstimator>(scale, ns, samples, H, thr, flags,
                                                                                               models, score, n_inl, masks, nthreads);
    return replay_scale<gcransac::utils::ScaleBasedRectifyingHomographyEstimator>(scale, ns, samples, H, thr, flags, models, score,
                                                                                  n_inl, masks, nthreads);
}

// Non-minimal fit of the scale-only estimators from all rows given (estimateModelNonminimal -> three_sift*.hpp:195-254).
int gcr_ref_fit_nonminimal_scale(const double* scale, long ns, int original, double* model6)
{
    gcransac::ScaleBasedRectifyingHomography m;
    std::vector<gcransac::ScaleBasedRectifyingHomography> models;
    bool ok;
    if (original)
    {
        using Est = gcransac::utils::OriginalScaleBasedRectifyingHomographyEstimator;
        Est::DataType data{make_mat(scale, ns)};
        Est::InlierContainerType inl{};
        for (long i = 0; i < ns; ++i) inl[0].push_back(static_cast<size_t>(i));
        ok = Est().estimateModelNonminimal(data, inl, models);
    }
    else
    {
        using Est = gcransac::utils::ScaleBasedRectifyingHomographyEstimator;
        Est::DataType data{make_mat(scale, ns)};
        Est::InlierContainerType inl{};
        for (long i = 0; i < ns; ++i) inl[0].push_back(static_cast<size_t>(i));
        ok = Est().estimateModelNonminimal(data, inl, models);
    }
    if (!ok || models.empty()) return 0;
    m = models[0];
    model6[0] = m.x0; model6[1] = m.y0; model6[2] = m.s; model6[3] = m.h7; model6[4] = m.h8; model6[5] = m.alpha;
    return 1;
}

// The reference's full scale-only drivers, unmodified.  model6 = (x0,y0,s,h7,h8,alpha).
int gcr_ref_find_scale_only(const double* scale, long ns, double thr, double spatial_coherence_weight, unsigned long min_iters,
                            unsigned long max_iters, unsigned long max_lo, int original, uint8_t* mask, double* H9, double* model6)
{
    std::vector<double> sv(scale, scale + 3 * ns), Hh;
    std::vector<bool> sm;
    gcransac::ScaleBasedRectifyingHomography model;
    const int n = original ? findRectifyingHomographyScaleOnlyOriginal_(sv, thr, spatial_coherence_weight, min_iters, max_iters,
                                                                        max_lo, sm, Hh, model, 0)
                           : findRectifyingHomographyScaleOnly_(sv, thr, spatial_coherence_weight, min_iters, max_iters, max_lo,
                                                                sm, Hh, model, 0);
    for (long i = 0; i < ns && i < static_cast<long>(sm.size()); ++i) mask[i] = sm[i] ? 1 : 0;
    for (size_t i = 0; i < 9 && i < Hh.size(); ++i) H9[i] = Hh[i];
    model6[0] = model.x0; model6[1] = model.y0; model6[2] = model.s;
    model6[3] = model.h7; model6[4] = model.h8; model6[5] = model.alpha;
    return n;
}

// ---- known-answer probes of the helper math (the functions tests/unit_tests.cpp pins) ----
double gcr_ref_clip_angle(double a) { return gcransac::utils::clipAngle(a); }
double gcr_ref_min_angle_diff(double a, double b) { return gcransac::utils::minAngleDiff(a, b); }
double gcr_ref_lines_angles_diff(double a, double b) { return gcransac::utils::linesAnglesDiff(a, b); }
double gcr_ref_deg2rad(double a) { return gcransac::utils::deg2rad(a); }
double gcr_ref_rad2deg(double a) { return gcransac::utils::rad2deg(a); }
unsigned long gcr_ref_nchoose2(unsigned long n) { return gcransac::utils::nChoose2(n); }
int gcr_ref_are_collinear(double x1, double y1, double x2, double y2, double x3, double y3, double tol)
{
    return gcransac::utils::areCollinear(x1, y1, x2, y2, x3, y3, tol) ? 1 : 0;
}
void gcr_ref_line_from_point_angle(double x, double y, double t, double* l3)
{
    const auto l = gcransac::utils::lineFromPointAndAngle(x, y, t);
    l3[0] = l(0); l3[1] = l(1); l3[2] = l(2);
}
// pts: [n,2] in, hull: [2n,2] out; returns the number of hull vertices
int gcr_ref_convex_hull(const double* pts, int n, double* hull)
{
    std::vector<gcransac::utils::Point2D> p;
    for (int i = 0; i < n; ++i) p.emplace_back(pts[2 * i], pts[2 * i + 1]);
    const auto h = gcransac::utils::computeConvexHull(p);
    for (size_t i = 0; i < h.size(); ++i) { hull[2 * i] = h[i].x(); hull[2 * i + 1] = h[i].y(); }
    return static_cast<int>(h.size());
}
int gcr_ref_point_in_convex_polygon(double x, double y, const double* poly, int n)
{
    std::vector<gcransac::utils::Point2D> p;
    for (int i = 0; i < n; ++i) p.emplace_back(poly[2 * i], poly[2 * i + 1]);
    return gcransac::utils::pointInConvexPolygon({x, y}, p) ? 1 : 0;
}
void gcr_ref_gauss3(const double* a12, double* x3)
{
    Eigen::Matrix<double, 3, 4> m;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 4; ++j) m(i, j) = a12[4 * i + j];
    Eigen::Matrix<double, 3, 1> r;
    gcransac::utils::gaussElimination<3>(m, r);
    x3[0] = r(0); x3[1] = r(1); x3[2] = r(2);
}
// model.h maps: which = 0 rectifiedScale, 1 unrectifiedScale, 2 rectifiedAngle, 3 unrectifiedAngle
double gcr_ref_model_map(int which, double h7, double h8, double x, double y, double v)
{
    gcransac::RectifyingHomography m;
    m.h7 = h7; m.h8 = h8;
    switch (which)
    {
    case 0: return m.rectifiedScale(x, y, v);
    case 1: return m.unrectifiedScale(x, y, v);
    case 2: return m.rectifiedAngle(x, y, v);
    default: return m.unrectifiedAngle(x, y, v);
    }
}
// which = 0 rectifyPoint, 1 unrectifyPoint (the double& overloads, model.h:140-154)
void gcr_ref_model_point(int which, double h7, double h8, double* x, double* y)
{
    gcransac::RectifyingHomography m;
    m.h7 = h7; m.h8 = h8;
    if (which == 0) m.rectifyPoint(*x, *y); else m.unrectifyPoint(*x, *y);
}
void gcr_ref_homography(const double* model7, double* H9)
{
    gcransac::RectifyingHomography m;
    m.x0 = model7[0]; m.y0 = model7[1]; m.s = model7[2]; m.h7 = model7[3]; m.h8 = model7[4];
    const Eigen::Matrix3d H = m.getHomography();
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) H9[3 * i + j] = H(i, j);
}

} // extern "C"
