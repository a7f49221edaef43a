// pybind11 module `pygcransac`: the Python surface of the reference (src/bindings.cpp:181-311,315-398) over the
// B200 engine.  Same function name, positional arguments, defaults, return tuple and exported classes, so
// `import pygcransac` keeps working for a user of the reference; the hybrid estimator's hypothesise/verify
// loop runs on the GPU.  Additions (keyword-only, all optional):
//   seed=None      reproducible runs (Philox key); None = non-deterministic like the reference's random_device
//   device=0       CUDA device ordinal
//   samples=None   int32 [H,4] minimal samples to replay instead of drawing (exactly H iterations)
//   return_stats   append a dict of run statistics as a 5th tuple element
// Feature arrays may be NumPy arrays (copied to the device inside the call) or CUDA tensors exposing
// `is_cuda` / `data_ptr()` (torch): those are read in place, zero-copy.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <sstream>

#include "../../../include/gcr_b200.h"
#include "gcr_host.hpp"
#include "gcr_refit.hpp"

namespace py = pybind11;
using namespace gcransac;

namespace
{

struct FeatureArg
{
    const double* ptr = nullptr;
    size_t rows = 0;
    bool on_device = false;
    int device = -1;
    py::object keep; // owner of the memory `ptr` points into
};

// ndim == 2, rows >= 2, cols == 3 (bindings.cpp:199-224), float64, C-contiguous
FeatureArg parseFeatures(const py::object& obj, const char* what, size_t kNumMinFeatures = 2)
{
    constexpr size_t kFeatureSize = 3;
    FeatureArg out;
    py::object src = obj;
    const bool is_cuda = py::hasattr(obj, "is_cuda") && py::hasattr(obj, "data_ptr") && obj.attr("is_cuda").cast<bool>();
    std::vector<size_t> shape;
    if (is_cuda)
    {
        for (auto d : obj.attr("shape")) shape.push_back(d.cast<size_t>());
    }
    else
    {
        auto arr = py::array_t<double, py::array::c_style | py::array::forcecast>::ensure(obj);
        if (!arr) throw std::invalid_argument(std::string(what) + " features must be convertible to a float64 array.");
        src = arr;
        for (py::ssize_t i = 0; i < arr.ndim(); ++i) shape.push_back(static_cast<size_t>(arr.shape(i)));
    }
    if (shape.size() != 2) throw std::invalid_argument("Number of dimensions must be 2.");
    if (shape[0] < kNumMinFeatures || shape[1] != kFeatureSize)
    {
        std::stringstream msg;
        msg << what << (what[0] ? " features" : "Features") << " should be an array with " << kFeatureSize << " columns and at least " << kNumMinFeatures
            << " rows. It has " << shape[1] << " columns and " << shape[0] << " rows.";
        throw std::invalid_argument(msg.str());
    }
    out.rows = shape[0];
    if (is_cuda)
    {
        const std::string dtype = py::str(obj.attr("dtype"));
        if (dtype.find("float64") == std::string::npos)
            throw std::invalid_argument(std::string(what) + " features on the device must be float64 (got " + dtype + ").");
        if (!obj.attr("is_contiguous")().cast<bool>())
            throw std::invalid_argument(std::string(what) + " features on the device must be contiguous.");
        out.ptr = reinterpret_cast<const double*>(obj.attr("data_ptr")().cast<uintptr_t>());
        out.on_device = true;
        py::object idx = obj.attr("device").attr("index");
        out.device = idx.is_none() ? 0 : idx.cast<int>();
    }
    else
    {
        out.ptr = static_cast<const double*>(py::array(src).data());
    }
    out.keep = src;
    return out;
}

// CUDA-tensor inputs: the stream whose work produced them.  `stream` = None: torch's current stream of that device (what a tensor written
// by the preceding torch calls is ordered on); an int: a raw cudaStream_t; an object with .cuda_stream (torch.cuda.Stream).
void setProducerStream(gcr_host::Options& opt, const py::object& stream, int device)
{
    uintptr_t raw = 0;
    if (stream.is_none())
    {
        py::object cur = py::module_::import("torch").attr("cuda").attr("current_stream")(device);
        raw = cur.attr("cuda_stream").cast<uintptr_t>();
    }
    else if (py::hasattr(stream, "cuda_stream")) raw = stream.attr("cuda_stream").cast<uintptr_t>();
    else raw = stream.cast<uintptr_t>();
    opt.wait_producer = true;
    opt.producer_stream = reinterpret_cast<void*>(raw);
}

py::array_t<bool> toMask(const std::vector<bool>& v)
{
    py::array_t<bool> out(static_cast<py::ssize_t>(v.size()));
    bool* p = static_cast<bool*>(out.request().ptr);
    for (size_t i = 0; i < v.size(); ++i) p[i] = v[i];
    return out;
}

py::dict toDict(const gcr_host::Stats& st)
{
    py::dict d;
    d["seed"] = st.seed;
    d["iteration_number"] = st.iteration_number;
    d["n_sample_valid"] = st.n_sample_valid;
    d["n_solved"] = st.n_solved;
    d["n_rechecked"] = st.n_rechecked;
    d["n_exact_hyp"] = st.n_exact_hyp;
    d["best_index"] = st.best_index;
    d["best_raw_score"] = st.best_raw_score;
    d["lo_score"] = st.lo_score;
    d["final_score"] = st.final_score;
    d["graph_cut_number"] = st.graph_cut_number;
    d["lo_models_scored"] = st.lo_models_scored;
    d["refit_accepted"] = st.refit_accepted;
    d["hypothesis_stage_ms"] = st.hypothesis_stage_ms;
    d["score_kernel_ms"] = st.score_kernel_ms;
    d["total_ms"] = st.total_ms;
    d["setup_ms"] = st.setup_ms;
    d["loop_ms"] = st.loop_ms;
    d["lo_ms"] = st.lo_ms;
    d["refit_ms"] = st.refit_ms;
    d["device_calls"] = st.device_calls;
    return d;
}

py::tuple findRectifyingHomographySIFT(py::object scale_features, py::object orientation_features, double scale_residual_thresh,
                                       double orientation_residual_thresh, double spatial_coherence_weight,
                                       size_t min_iteration_number, size_t max_iteration_number,
                                       size_t max_local_optimization_number, py::object seed, int device, py::object samples,
                                       bool return_stats, py::object stream)
{
    FeatureArg fs = parseFeatures(scale_features, "Scale");
    FeatureArg fo = parseFeatures(orientation_features, "Orientation");
    if (fs.on_device != fo.on_device)
        throw std::invalid_argument("Scale and orientation features must both be host arrays or both be CUDA tensors.");
    gcr_host::Options opt;
    opt.device = device;
    if (fs.on_device)
    {
        if (fs.device != fo.device) throw std::invalid_argument("Feature tensors live on different devices.");
        opt.device = fs.device;
        opt.features_on_device = true;
        setProducerStream(opt, stream, fs.device);
    }
    if (!seed.is_none())
    {
        opt.has_seed = true;
        opt.seed = seed.cast<uint64_t>();
    }
    py::array_t<int32_t, py::array::c_style | py::array::forcecast> smp;
    if (!samples.is_none())
    {
        smp = py::array_t<int32_t, py::array::c_style | py::array::forcecast>::ensure(samples);
        if (!smp || smp.ndim() != 2 || smp.shape(1) != 4 || smp.shape(0) < 1)
            throw std::invalid_argument("samples must be an int32 array of shape [H,4] = (scale0, scale1, orient0, orient1).");
        opt.samples = smp.data();
        opt.n_samples = smp.shape(0);
    }

    std::vector<double> homography(9);
    std::vector<bool> scale_inliers, orientation_inliers;
    SIFTRectifyingHomography model;
    gcr_host::Stats st;
    int num_inliers = 0;
    {
        py::gil_scoped_release nogil; // the reference holds the GIL for the whole call; nothing here needs it
        num_inliers = gcr_host::findRectifyingHomographySIFT(
            fs.ptr, fs.rows, fo.ptr, fo.rows, scale_residual_thresh, orientation_residual_thresh, spatial_coherence_weight,
            min_iteration_number, max_iteration_number, max_local_optimization_number, scale_inliers, orientation_inliers,
            homography, model, opt, &st);
    }
    py::array_t<bool> smask = toMask(scale_inliers), omask = toMask(orientation_inliers);
    py::object H = py::none(), M = py::none();
    if (num_inliers != 0) // bindings.cpp:292-310
    {
        py::array_t<double> Harr({3, 3});
        std::copy(homography.begin(), homography.end(), static_cast<double*>(Harr.request().ptr));
        H = Harr;
        M = py::cast(model);
    }
    if (return_stats) return py::make_tuple(H, smask, omask, M, toDict(st));
    return py::make_tuple(H, smask, omask, M);
}

// findRectifyingHomographyScaleOnly / ...Original (bindings.cpp:19-179): (H, inliers, model) or (None, inliers)
py::tuple scaleOnlyImpl(bool original, py::object features, double scale_residual_thresh, double spatial_coherence_weight,
                        size_t min_iteration_number, size_t max_iteration_number, size_t max_local_optimization_number,
                        py::object seed, int device, py::object samples, bool return_stats, py::object stream)
{
    FeatureArg f = parseFeatures(features, "", 3);
    gcr_host::Options opt;
    opt.device = device;
    if (f.on_device)
    {
        opt.device = f.device;
        opt.features_on_device = true;
        setProducerStream(opt, stream, f.device);
    }
    if (!seed.is_none())
    {
        opt.has_seed = true;
        opt.seed = seed.cast<uint64_t>();
    }
    py::array_t<int32_t, py::array::c_style | py::array::forcecast> smp;
    if (!samples.is_none())
    {
        smp = py::array_t<int32_t, py::array::c_style | py::array::forcecast>::ensure(samples);
        if (!smp || smp.ndim() != 2 || smp.shape(1) != 4 || smp.shape(0) < 1)
            throw std::invalid_argument("samples must be an int32 array of shape [H,4] = (s0, s1, s2, ignored).");
        opt.samples = smp.data();
        opt.n_samples = smp.shape(0);
    }
    std::vector<double> homography(9);
    std::vector<bool> inliers;
    ScaleBasedRectifyingHomography model;
    gcr_host::Stats st;
    int num_inliers = 0;
    {
        py::gil_scoped_release nogil;
        num_inliers = gcr_host::findRectifyingHomographyScaleOnly(f.ptr, f.rows, scale_residual_thresh, spatial_coherence_weight,
                                                                  min_iteration_number, max_iteration_number,
                                                                  max_local_optimization_number, original, inliers, homography,
                                                                  model, opt, &st);
    }
    py::array_t<bool> mask = toMask(inliers);
    if (num_inliers == 0)
    {
        if (return_stats) return py::make_tuple(py::none(), mask, toDict(st));
        return py::make_tuple(py::none(), mask);
    }
    py::array_t<double> Harr({3, 3});
    std::copy(homography.begin(), homography.end(), static_cast<double*>(Harr.request().ptr));
    if (return_stats) return py::make_tuple(Harr, mask, py::cast(model), toDict(st));
    return py::make_tuple(Harr, mask, py::cast(model));
}

py::tuple findRectifyingHomographyScaleOnly(py::object features, double thr, double scw, size_t min_it, size_t max_it, size_t max_lo,
                                            py::object seed, int device, py::object samples, bool return_stats, py::object stream)
{
    return scaleOnlyImpl(false, features, thr, scw, min_it, max_it, max_lo, seed, device, samples, return_stats, stream);
}

py::tuple findRectifyingHomographyScaleOnlyOriginal(py::object features, double thr, double scw, size_t min_it, size_t max_it,
                                                    size_t max_lo, py::object seed, int device, py::object samples, bool return_stats,
                                                    py::object stream)
{
    return scaleOnlyImpl(true, features, thr, scw, min_it, max_it, max_lo, seed, device, samples, return_stats, stream);
}

// host-side non-minimal fit, exposed for the parity tests of the refit step (two_sift.hpp:423-579)
py::object fitNonMinimalPy(py::array_t<double, py::array::c_style | py::array::forcecast> scale,
                           py::array_t<double, py::array::c_style | py::array::forcecast> orient)
{
    if (scale.ndim() != 2 || orient.ndim() != 2 || scale.shape(1) != 3 || orient.shape(1) != 3)
        throw std::invalid_argument("features must be [N,3] arrays");
    std::vector<uint32_t> si(scale.shape(0)), oi(orient.shape(0));
    for (size_t i = 0; i < si.size(); ++i) si[i] = static_cast<uint32_t>(i);
    for (size_t i = 0; i < oi.size(); ++i) oi[i] = static_cast<uint32_t>(i);
    SIFTRectifyingHomography m;
    if (!gcr_host::fitNonMinimal(scale.data(), si, orient.data(), oi, m)) return py::none();
    return py::cast(m);
}

} // namespace

PYBIND11_MODULE(pygcransac, m)
{
    m.doc() = "pygcransac on the B200 hybrid-RANSAC engine (gcr_b200): drop-in for yuvalnis/graph-cut-ransac's module";
    m.attr("__backend__") = gcr_version();

    py::class_<NormalizingTransform>(m, "NormalizingTransform")
        .def(py::init<>())
        .def_readwrite("x0", &NormalizingTransform::x0)
        .def_readwrite("y0", &NormalizingTransform::y0)
        .def_readwrite("s", &NormalizingTransform::s);

    py::class_<RectifyingHomography, NormalizingTransform>(m, "RectifyingHomography")
        .def(py::init<>())
        .def_readwrite("h7", &RectifyingHomography::h7)
        .def_readwrite("h8", &RectifyingHomography::h8)
        .def("rectifiedScale", &RectifyingHomography::rectifiedScale)
        .def("unrectifiedScale", &RectifyingHomography::unrectifiedScale)
        .def("rectifiedAngle", &RectifyingHomography::rectifiedAngle)
        .def("unrectifiedAngle", &RectifyingHomography::unrectifiedAngle)
        .def("rectifiedPoint",
             [](const RectifyingHomography& self, double x, double y) {
                 self.rectifyPoint(x, y);
                 return py::make_tuple(x, y);
             },
             "Rectify point")
        .def("unrectifiedPoint",
             [](const RectifyingHomography& self, double x, double y) {
                 self.unrectifyPoint(x, y);
                 return py::make_tuple(x, y);
             },
             "Unrectify point")
        .def("getHomography",
             [](const RectifyingHomography& self) {
                 const auto h = self.getHomography();
                 py::array_t<double> out({3, 3});
                 std::copy(h.begin(), h.end(), static_cast<double*>(out.request().ptr));
                 return out;
             },
             "Get homography matrix");

    py::class_<ScaleBasedRectifyingHomography, RectifyingHomography>(m, "ScaleBasedRectifyingHomography")
        .def(py::init<>())
        .def_readwrite("alpha", &ScaleBasedRectifyingHomography::alpha);

    py::class_<OrientationBasedRectifyingHomography, RectifyingHomography>(m, "OrientationBasedRectifyingHomography")
        .def(py::init<>())
        .def_readwrite("phi", &OrientationBasedRectifyingHomography::phi);

    py::class_<SIFTRectifyingHomography, ScaleBasedRectifyingHomography, OrientationBasedRectifyingHomography>(
        m, "SIFTRectifyingHomography")
        .def(py::init<>());

    m.def("findRectifyingHomographyScaleOnly", &findRectifyingHomographyScaleOnly,
          "Scale-only rectifying-homography RANSAC (3 scale features per sample); hypothesis stage on the GPU.",
          py::arg("features"), py::arg("scale_residual_thresh"), py::arg("spatial_coherence_weight") = 0.0,
          py::arg("min_iteration_number") = 10000, py::arg("max_iteration_number") = 10000,
          py::arg("max_local_optimization_number") = 50, py::kw_only(), py::arg("seed") = py::none(), py::arg("device") = 0,
          py::arg("samples") = py::none(), py::arg("return_stats") = false, py::arg("stream") = py::none());

    m.def("findRectifyingHomographyScaleOnlyOriginal", &findRectifyingHomographyScaleOnlyOriginal,
          "The 'original' scale-only formulation (rows [x, y, -s^(-1/3)] = -1); hypothesis stage on the GPU.",
          py::arg("features"), py::arg("scale_residual_thresh"), py::arg("spatial_coherence_weight") = 0.0,
          py::arg("min_iteration_number") = 10000, py::arg("max_iteration_number") = 10000,
          py::arg("max_local_optimization_number") = 50, py::kw_only(), py::arg("seed") = py::none(), py::arg("device") = 0,
          py::arg("samples") = py::none(), py::arg("return_stats") = false, py::arg("stream") = py::none());

    m.def("findRectifyingHomographySIFT", &findRectifyingHomographySIFT,
          "Hybrid (scale + orientation) rectifying-homography RANSAC; hypothesis stage on the GPU.",
          py::arg("scale_features"), py::arg("orientation_features"), py::arg("scale_residual_thresh"),
          py::arg("orientation_residual_thresh"), py::arg("spatial_coherence_weight") = 0.0,
          py::arg("min_iteration_number") = 10000, py::arg("max_iteration_number") = 10000,
          py::arg("max_local_optimization_number") = 50, py::kw_only(), py::arg("seed") = py::none(),
          py::arg("device") = 0, py::arg("samples") = py::none(), py::arg("return_stats") = false, py::arg("stream") = py::none());

    m.def(
        "_fit_non_minimal_scale",
        [](py::array_t<double, py::array::c_style | py::array::forcecast> scale, bool original) -> py::object {
            if (scale.ndim() != 2 || scale.shape(1) != 3) throw std::invalid_argument("features must be an [N,3] array");
            std::vector<uint32_t> idx(scale.shape(0));
            for (size_t i = 0; i < idx.size(); ++i) idx[i] = static_cast<uint32_t>(i);
            SIFTRectifyingHomography m;
            if (!gcr_host::fitNonMinimalScale(scale.data(), idx, original, m)) return py::none();
            return py::cast(static_cast<ScaleBasedRectifyingHomography&>(m));
        },
        "Host non-minimal fit of the scale-only estimators from all rows given (test hook).", py::arg("features"),
        py::arg("original") = false);
    m.def("_fit_non_minimal", &fitNonMinimalPy, "Host non-minimal fit from all rows given (test hook).", py::arg("scale_features"),
          py::arg("orientation_features"));
    m.def(
        "_iteration_number",
        [](uint32_t n_scale_inliers, uint32_t n_orient_inliers, size_t ns, size_t no, double confidence, int problem) {
            const uint32_t n[2] = {n_scale_inliers, n_orient_inliers};
            return gcr_host::iterationNumber(n, ns, no, confidence, problem);
        },
        "GCRANSAC::getIterationNumber for sample sizes {2,2} (problem 0) or {3} (test hook).", py::arg("n_scale_inliers"),
        py::arg("n_orient_inliers"), py::arg("ns"), py::arg("no"), py::arg("confidence"), py::arg("problem") = 0);
}
