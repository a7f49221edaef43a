// Host-side non-minimal model fitting for the hybrid rectifying-homography problem: what the reference
// runs on the best model only, off the measured path:
//   constraints   scale rows  [x, y, s^(1/3)] . (h7, h8, alpha) = 1              two_sift.hpp:228-238
//                 one row per PAIR of orientation features, [vp.x, vp.y, 0] . (h7, h8, alpha) = vp.z with
//                 vp = l_i x l_j scaled by its largest |coordinate| when that exceeds 1   two_sift.hpp:240-260
//   solve         least squares over all rows                                   two_sift.hpp:523-537
//   phi           weighted mode of the rectified orientations, 0.5 degree bins  two_sift.hpp:354-394,539-576
// The reference solves with Eigen's column-pivoted Householder QR on the dense N x 3 matrix.  Here the rows are
// streamed through a blocked Householder reduction to a 3 x 3 triangle (N can be millions: all pairs of the
// orientation inliers), then a pivoted back-substitution; parity with the reference is at tolerance level
// (vanishing line <= 1e-6 relative), as SURVEY.md 7.4(2) allows for this off-path step.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <unordered_map>
#include <vector>

#include "gcr_model.hpp"

namespace gcr_host
{

// Least squares min |A x - b| for A of shape N x 3, rows supplied one at a time.
class StreamingLsq3
{
public:
    explicit StreamingLsq3(size_t block_rows = 4096) : cap_(block_rows + 3) { buf_.reserve(cap_ * 4); reset(); }

    void reset()
    {
        buf_.assign(12, 0.0); // rows 0-2: the running triangle [R | Q^T b]
        n_rows_ = 0;
    }
    void addRow(double a0, double a1, double a2, double b)
    {
        buf_.push_back(a0);
        buf_.push_back(a1);
        buf_.push_back(a2);
        buf_.push_back(b);
        ++n_rows_;
        if (buf_.size() >= cap_ * 4) reduce();
    }
    size_t rows() const { return n_rows_; }

    // Rank-revealing solve of the reduced system; free variables of a rank-deficient system are set to zero
    // (the basic solution a column-pivoted QR returns).  False when the solution is not finite.
    bool solve(double x[3])
    {
        reduce();
        double R[3][4];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 4; ++j) R[i][j] = buf_[4 * i + j];
        // column-pivoted Householder QR of the 3 x 3 triangle
        int perm[3] = {0, 1, 2};
        double maxdiag = 0.0;
        int rank = 0;
        for (int k = 0; k < 3; ++k)
        {
            int best = k;
            double best_norm = -1.0;
            for (int j = k; j < 3; ++j)
            {
                double nrm = 0.0;
                for (int i = k; i < 3; ++i) nrm += R[i][j] * R[i][j];
                if (nrm > best_norm) { best_norm = nrm; best = j; }
            }
            if (best != k)
            {
                for (int i = 0; i < 3; ++i) std::swap(R[i][k], R[i][best]);
                std::swap(perm[k], perm[best]);
            }
            householderColumn(&R[0][0], 3, 4, k);
            const double d = std::fabs(R[k][k]);
            if (k == 0) maxdiag = d;
            if (!(d > maxdiag * 3.0 * std::numeric_limits<double>::epsilon())) break;
            ++rank;
        }
        double y[3] = {0.0, 0.0, 0.0};
        for (int i = rank - 1; i >= 0; --i)
        {
            double acc = R[i][3];
            for (int j = i + 1; j < rank; ++j) acc -= R[i][j] * y[j];
            y[i] = acc / R[i][i];
        }
        for (int k = 0; k < 3; ++k) x[perm[k]] = y[k];
        return std::isfinite(x[0]) && std::isfinite(x[1]) && std::isfinite(x[2]);
    }

private:
    // One Householder reflection zeroing column k below the diagonal of the row-major m x n matrix a.
    static void householderColumn(double* a, size_t m, size_t n, size_t k)
    {
        double scale = 0.0;
        for (size_t i = k; i < m; ++i) scale = std::max(scale, std::fabs(a[i * n + k]));
        if (scale == 0.0 || !std::isfinite(scale)) return;
        double sigma = 0.0;
        for (size_t i = k; i < m; ++i)
        {
            const double v = a[i * n + k] / scale;
            sigma += v * v;
        }
        const double akk = a[k * n + k] / scale;
        const double norm = std::sqrt(sigma);
        const double alpha = akk > 0.0 ? -norm : norm;
        // v = x/scale - alpha e_k ; H = I - 2 v v^T / (v^T v)
        const double vk = akk - alpha;
        const double vtv = sigma - akk * akk + vk * vk;
        if (vtv == 0.0) return;
        for (size_t j = k + 1; j < n; ++j)
        {
            double dot = vk * a[k * n + j];
            for (size_t i = k + 1; i < m; ++i) dot += (a[i * n + k] / scale) * a[i * n + j];
            const double f = 2.0 * dot / vtv;
            a[k * n + j] -= f * vk;
            for (size_t i = k + 1; i < m; ++i) a[i * n + j] -= f * (a[i * n + k] / scale);
        }
        a[k * n + k] = alpha * scale;
        for (size_t i = k + 1; i < m; ++i) a[i * n + k] = 0.0;
    }

    void reduce()
    {
        const size_t m = buf_.size() / 4;
        if (m > 3)
        {
            for (size_t k = 0; k < 3; ++k) householderColumn(buf_.data(), m, 4, k);
            buf_.resize(12);
        }
    }

    size_t cap_;
    size_t n_rows_ = 0;
    std::vector<double> buf_;
};

// Weighted mode over bins of width `bin_width`: the weighted mean of the heaviest bin (two_sift.hpp:354-394).
// With unit weights equal-count bins tie EXACTLY, and the reference resolves ties by the iteration order of a
// std::unordered_map<int, double> filled in input order (first strictly heavier bin wins).  The same container,
// filled in the same order, is used here so that ties resolve identically under the same standard library.
inline double weightedMode(const std::vector<double>& angles, const std::vector<double>& weights, double bin_width)
{
    std::unordered_map<int, double> bin_weight, bin_moment;
    for (size_t i = 0; i < angles.size(); ++i)
    {
        const int bin = static_cast<int>(std::round(angles[i] / bin_width));
        bin_weight[bin] += weights[i];
        bin_moment[bin] += angles[i] * weights[i];
    }
    int mode_bin = 0;
    double heaviest = -1.0;
    for (const auto& kv : bin_weight)
        if (kv.second > heaviest)
        {
            heaviest = kv.second;
            mode_bin = kv.first;
        }
    return bin_moment[mode_bin] / bin_weight[mode_bin];
}

// Non-minimal fit from the listed features (rows of f64[N,3] arrays); unit weights.  Returns false when the
// reference's solver would (too few constraints, non-finite solution, alpha < 1e-9).  The model's normalising
// transform stays the identity, as in the reference (two_sift.hpp:843-845).
inline bool fitNonMinimal(const double* scale, const std::vector<uint32_t>& scale_idx, const double* orient,
                          const std::vector<uint32_t>& orient_idx, gcransac::SIFTRectifyingHomography& model)
{
    constexpr double kPi = 3.14159265358979323846;
    const size_t ns = scale_idx.size(), no = orient_idx.size();
    if (ns < 2 || no < 2) return false;
    // exactly a minimal sample goes to the minimal solver in the reference (two_sift.hpp:603-609); neither the local
    // optimisation (needs > 2 per class) nor the final refit (needs a non-minimal inlier set) can get here with one
    if (ns == 2 && no == 2) return false;
    StreamingLsq3 lsq;
    for (uint32_t i : scale_idx)
    {
        const double* f = scale + 3 * static_cast<size_t>(i);
        lsq.addRow(f[0], f[1], std::pow(f[2], 1.0 / 3.0), 1.0);
    }
    std::vector<double> lines(3 * no);
    for (size_t a = 0; a < no; ++a)
    {
        const double* f = orient + 3 * static_cast<size_t>(orient_idx[a]);
        const double c = std::cos(f[2]), s = std::sin(f[2]);
        lines[3 * a] = s;
        lines[3 * a + 1] = -c;
        lines[3 * a + 2] = f[1] * c - f[0] * s;
    }
    for (size_t a = 0; a + 1 < no; ++a)
        for (size_t b = a + 1; b < no; ++b)
        {
            const double* p = &lines[3 * a];
            const double* q = &lines[3 * b];
            double v0 = p[1] * q[2] - p[2] * q[1], v1 = p[2] * q[0] - p[0] * q[2], v2 = p[0] * q[1] - p[1] * q[0];
            const double big = std::max(std::fabs(v0), std::max(std::fabs(v1), std::fabs(v2)));
            if (big > 1.0) { v0 /= big; v1 /= big; v2 /= big; }
            lsq.addRow(v0, v1, 0.0, v2);
        }
    double x[3];
    if (!lsq.solve(x)) return false;
    if (x[2] < 1e-9) return false;
    model = gcransac::SIFTRectifyingHomography();
    model.h7 = x[0];
    model.h8 = x[1];
    model.alpha = x[2];
    std::vector<double> ang(no), wts(no, 1.0 / static_cast<double>(no));
    for (size_t a = 0; a < no; ++a)
    {
        const double* f = orient + 3 * static_cast<size_t>(orient_idx[a]);
        double t = model.rectifiedAngle(f[0], f[1], f[2]);
        if (t > kPi) t -= kPi; // theta and theta + pi are the same line
        ang[a] = t;
    }
    model.phi = weightedMode(ang, wts, 0.5 * kPi / 180.0);
    return true;
}

// Non-minimal fit of the scale-only estimators: rows [x, y, s^(1/3)] = 1 (three_sift.hpp:195-254) or, for the
// "original" variant, [x, y, -s^(-1/3)] = -1 (three_sift_original.hpp:229-234).  Exactly a minimal sample (3 rows) goes
// to the minimal solver in the reference; the local optimisation and the refit never present one (they need > 3 inliers).
inline bool fitNonMinimalScale(const double* scale, const std::vector<uint32_t>& idx, bool original,
                               gcransac::SIFTRectifyingHomography& model)
{
    if (idx.size() <= 3) return false;
    StreamingLsq3 lsq;
    for (uint32_t i : idx)
    {
        const double* f = scale + 3 * static_cast<size_t>(i);
        if (original) lsq.addRow(f[0], f[1], -std::pow(f[2], -1.0 / 3.0), -1.0);
        else lsq.addRow(f[0], f[1], std::pow(f[2], 1.0 / 3.0), 1.0);
    }
    double x[3];
    if (!lsq.solve(x)) return false;
    if (x[2] < 1e-9) return false;
    model = gcransac::SIFTRectifyingHomography();
    model.h7 = x[0];
    model.h8 = x[1];
    model.alpha = x[2];
    return true;
}

} // namespace gcr_host
