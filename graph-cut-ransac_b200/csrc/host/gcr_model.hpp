// Model types of the rectifying-homography problem, host side.
//
// Same names, members and meaning as the reference's model hierarchy (src/pygcransac/include/model.h:42-246)
// so that the Python classes exported by the binding are interchangeable with the reference's:
//   NormalizingTransform {x0, y0, s}                                  model.h:42-126
//   RectifyingHomography : NormalizingTransform {h7, h8} + maps       model.h:128-227
//   ScaleBasedRectifyingHomography {alpha}, OrientationBasedRectifyingHomography {phi}
//   SIFTRectifyingHomography = both                                   model.h:229-246
// Written against plain doubles (no Eigen): the warp is H = [[1,0,0],[0,1,0],[h7,h8,1]], rectification applies
// its inverse, i.e. only the homogeneous coordinate changes.
#pragma once
#include <array>
#include <cmath>

namespace gcransac
{

namespace detail
{
constexpr double kTwoPi = 6.283185307179586476925286766559;

// angle folded into [0, 2 pi)  (math_utils.hpp:78-88)
inline double wrapTwoPi(double a)
{
    a = std::fmod(a, kTwoPi);
    return a < 0.0 ? a + kTwoPi : a;
}
} // namespace detail

struct NormalizingTransform
{
    double x0 = 0.0; // mean of the un-normalised sample positions
    double y0 = 0.0;
    double s = 1.0; // isotropic rescaling factor

    // homogeneous coordinates (x, y, w)
    void normalize(double& x, double& y, double w) const
    {
        x = s * (x - x0 * w);
        y = s * (y - y0 * w);
    }
    void denormalize(double& x, double& y, double w) const
    {
        const double inv = 1.0 / s;
        x = inv * x + x0 * w;
        y = inv * y + y0 * w;
    }
    void normalize(double& x, double& y) const { normalize(x, y, 1.0); }
    void denormalize(double& x, double& y) const { denormalize(x, y, 1.0); }
    void normalizeScale(double& scale) const { scale *= s; }
    void denormalizeScale(double& scale) const { scale /= s; }
    void normalizeFeature(double& x, double& y, double& scale) const
    {
        normalizeScale(scale);
        normalize(x, y, 1.0);
    }
    void denormalizeFeature(double& x, double& y, double& scale) const
    {
        denormalizeScale(scale);
        denormalize(x, y, 1.0);
    }
};

struct RectifyingHomography : public NormalizingTransform
{
    double h7 = 0.0;
    double h8 = 0.0;

    // homogeneous coordinate after rectification / warping of (x, y, w)
    double rectifiedW(double x, double y, double w) const { return -h7 * x - h8 * y + w; }
    double unrectifiedW(double x, double y, double w) const { return h7 * x + h8 * y + w; }

    void rectifyPoint(double& x, double& y) const
    {
        const double w = rectifiedW(x, y, 1.0);
        x /= w;
        y /= w;
    }
    void unrectifyPoint(double& x, double& y) const
    {
        const double w = unrectifiedW(x, y, 1.0);
        x /= w;
        y /= w;
    }

    // direction of a line element at (x, y) with angle `angle` after rectification (model.h:156-165)
    double rectifiedAngle(double x, double y, double angle) const
    {
        const double c = std::cos(angle), sn = std::sin(angle);
        const double num = (-x * sn + y * c) * h7 + sn;
        const double den = (x * sn - y * c) * h8 + c;
        return detail::wrapTwoPi(std::atan2(num, den));
    }
    // ... and after warping a rectified element back (model.h:167-174)
    double unrectifiedAngle(double x, double y, double angle) const
    {
        const double c = std::cos(angle), sn = std::sin(angle);
        const double num = (x * sn - y * c) * h7 + sn;
        const double den = (-x * sn + y * c) * h8 + c;
        return detail::wrapTwoPi(std::atan2(num, den));
    }

    // local area-scale change of the warp at a rectified position, and of its inverse at a warped position
    double localScalePerspectiveWarp(double ux, double uy) const { return std::pow(h7 * ux + h8 * uy + 1.0, -3.0); }
    double localScaleAffineRectification(double dx, double dy) const { return std::pow(-h7 * dx - h8 * dy + 1.0, -3.0); }
    double rectifiedScale(double dx, double dy, double ds) const { return ds * localScaleAffineRectification(dx, dy); }
    double unrectifiedScale(double ux, double uy, double us) const { return us * localScalePerspectiveWarp(ux, uy); }

    // N^-1 H N scaled so that the (2,2) entry is 1, row-major (model.h:211-226); N = [[s,0,-s x0],[0,s,-s y0],[0,0,1]]
    std::array<double, 9> getHomography() const
    {
        // H N: rows 0 and 1 of N, row 2 = h7 N0 + h8 N1 + N2
        const double n0[3] = {s, 0.0, -s * x0}, n1[3] = {0.0, s, -s * y0};
        double r2[3];
        for (int j = 0; j < 3; ++j) r2[j] = h7 * n0[j] + h8 * n1[j] + (j == 2 ? 1.0 : 0.0);
        // N^-1 = [[1/s,0,x0],[0,1/s,y0],[0,0,1]]
        const double inv = 1.0 / s;
        std::array<double, 9> out{};
        for (int j = 0; j < 3; ++j)
        {
            out[j] = inv * n0[j] + x0 * r2[j];
            out[3 + j] = inv * n1[j] + y0 * r2[j];
            out[6 + j] = r2[j];
        }
        const double d = out[8];
        for (double& v : out) v /= d;
        return out;
    }
};

struct ScaleBasedRectifyingHomography : virtual public RectifyingHomography
{
    double alpha{1.0}; // feature-class scale: relative scale of the rectified features
};

struct OrientationBasedRectifyingHomography : virtual public RectifyingHomography
{
    double phi{0.0}; // direction [rad] of the dominant orientation in the rectified image
};

struct SIFTRectifyingHomography : public ScaleBasedRectifyingHomography, OrientationBasedRectifyingHomography
{
};

} // namespace gcransac
