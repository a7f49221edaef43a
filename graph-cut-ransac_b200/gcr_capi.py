"""ctypes view of the C ABI in include/gcr_b200.h (libgcr_b200.so).

This is what tests/ (``-m gpu``) and bench.py call: plain pointers and sizes straight into the shared
library, no torch types.  The product has no CPU fallback: if the library is missing or was not built
for this GPU, construction fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GCR_B200_LIB", os.path.join(PKG_DIR, "libgcr_b200.so"))  # override: kernel-variant sweeps

FLAG_SAMPLE_VALID, FLAG_SOLVED, FLAG_MODEL_VALID, FLAG_EXACT = 1, 2, 4, 8
RUN_EMIT_MASKS, RUN_SAMPLES_ON_DEVICE, RUN_KEEP_PER_HYP, RUN_ALL_EXACT, RUN_EXCHANGE, RUN_TIME_KERNELS, RUN_SAMPLES_U16 = 1, 2, 4, 8, 16, 32, 64
FEATURES_ON_DEVICE, FEATURES_HOST_LIBM = 1, 2
IPC_HANDLE_BYTES = 64
PROBLEM_HYBRID, PROBLEM_SCALE_ONLY, PROBLEM_SCALE_ONLY_ORIGINAL = 0, 1, 2
ERR_INVALID, ERR_CUDA, ERR_STATE, ERR_NOMEM = -1, -2, -3, -4
SCORE_TRUNCATED_SQ, SCORE_STRICT = 1, 2

EXPORTS = [
    "gcr_create", "gcr_destroy", "gcr_last_error", "gcr_set_stream", "gcr_wait_stream", "gcr_set_problem", "gcr_set_features", "gcr_get_features", "gcr_run",
    "gcr_best",
    "gcr_fetch", "gcr_score_model", "gcr_score_models", "gcr_fit_models", "gcr_set_keypoints", "gcr_debug_eval", "gcr_version", "gcr_peer_export",
    "gcr_peer_attach", "gcr_peer_buffer", "gcr_peer_attach_ptrs", "gcr_peer_detach", "gcr_score_models_static", "gcr_fit_score_models", "gcr_debug_phi_fallbacks",
    "gcr_lo_round", "gcr_set_sm_share",
]


class GcrResult(C.Structure):
    _fields_ = [
        ("best_index", C.c_int64), ("best_score", C.c_double), ("best_model", C.c_double * 4),
        ("best_n_inl", C.c_uint32 * 2), ("pad_", C.c_uint32 * 2), ("n_hyp", C.c_int64),
        ("n_sample_valid", C.c_int64), ("n_solved", C.c_int64), ("n_rechecked", C.c_int64),
        ("n_exact_hyp", C.c_int64), ("n_candidates", C.c_int64), ("device_ms", C.c_float),
        ("score_kernel_ms", C.c_float), ("winner_rank", C.c_int32), ("score_kernel", C.c_int32),
        ("tail_phase_us", C.c_float * 6),
        ("stage_phase_us", C.c_float * 2),
        ("n_refined", C.c_int64),
        ("n_scored", C.c_int64),
    ]


class GcrLoRoundIo(C.Structure):  # gcr_lo_round_io
    _fields_ = [
        ("model", C.c_double * 4), ("score", C.c_double), ("n_inl", C.c_uint32 * 2), ("rng_state", C.c_uint64),
        ("updated", C.c_uint32), ("sampled", C.c_uint32), ("n_models_scored", C.c_uint32), ("phi_mismatch", C.c_uint32),
        ("n_labelled", C.c_uint32 * 2),
    ]


RECORD_DTYPE = np.dtype([("score", "<f8"), ("index", "<i8"), ("model", "<f8", 4), ("n_inl", "<u4", 2),
                         ("reserved", "<u8")])
assert RECORD_DTYPE.itemsize == 64


class GcrError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"gcr_b200 error {code}: {msg}")
        self.code = code


def load_library() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i64, u64, u32 = C.c_void_p, C.c_int64, C.c_uint64, C.c_uint32
    lib.gcr_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.gcr_destroy.argtypes = [vp]
    lib.gcr_last_error.argtypes = [vp]
    lib.gcr_last_error.restype = C.c_char_p
    lib.gcr_set_stream.argtypes = [vp, vp]
    lib.gcr_wait_stream.argtypes = [vp, vp]
    lib.gcr_set_problem.argtypes = [vp, C.c_int]
    lib.gcr_set_sm_share.argtypes = [vp, C.c_int]
    lib.gcr_set_features.argtypes = [vp, vp, i64, vp, i64, C.c_int]
    lib.gcr_get_features.argtypes = [vp, vp, vp]
    lib.gcr_run.argtypes = [vp, i64, u64, u64, vp, C.POINTER(C.c_double), u32, vp, C.POINTER(GcrResult)]
    lib.gcr_best.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    lib.gcr_fetch.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    lib.gcr_score_model.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    lib.gcr_score_models.argtypes = [vp, vp, i64, vp, u32, vp, vp, vp]
    lib.gcr_fit_models.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp]
    lib.gcr_fit_score_models.argtypes = [vp, i64, vp, vp, vp, vp, vp, u32, vp, vp, vp, vp, vp]
    lib.gcr_lo_round.argtypes = [vp, C.POINTER(GcrLoRoundIo), vp, u32, C.c_int32, vp, vp, vp]
    lib.gcr_set_keypoints.argtypes = [vp, vp, i64, C.c_int, C.c_int, vp, vp, vp]
    lib.gcr_peer_export.argtypes = [vp, vp]
    lib.gcr_peer_attach.argtypes = [vp, vp, C.c_int, C.c_int]
    lib.gcr_peer_buffer.argtypes = [vp, C.POINTER(vp)]
    lib.gcr_peer_attach_ptrs.argtypes = [vp, C.POINTER(vp), C.c_int, C.c_int]
    lib.gcr_peer_detach.argtypes = [vp]
    lib.gcr_debug_eval.argtypes = [vp, C.c_int, vp, i64, vp, vp]
    lib.gcr_score_models_static.argtypes = [C.c_int, vp, i64, vp, i64, vp, i64, C.POINTER(C.c_double), u32, vp, vp, vp, vp, C.c_char_p, C.c_size_t]
    lib.gcr_version.restype = C.c_char_p
    for name in EXPORTS:
        fn = getattr(lib, name)
        if name == "gcr_debug_phi_fallbacks":
            fn.restype, fn.argtypes = C.c_longlong, [vp]
            continue
        if name not in ("gcr_last_error", "gcr_version"):
            fn.restype = C.c_int
    return lib


def unpack_mask(words: np.ndarray, n: int) -> np.ndarray:
    bits = np.unpackbits(np.ascontiguousarray(words, np.uint32).view(np.uint8), bitorder="little")
    return bits[:n].astype(bool)


class Engine:
    """Thin RAII wrapper of a gcr_handle."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        self.h = C.c_void_p()
        rc = self.lib.gcr_create(device, C.byref(self.h))
        if rc != 0:
            msg = self.lib.gcr_last_error(self.h).decode() if self.h else "gcr_create failed"
            raise GcrError(rc, msg)
        self.ns = self.no = 0
        self.n_hyp = 0
        self.problem = PROBLEM_HYBRID

    def close(self):
        if getattr(self, "h", None):
            self.lib.gcr_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise GcrError(rc, self.lib.gcr_last_error(self.h).decode())

    def set_stream(self, cuda_stream: int):
        self._check(self.lib.gcr_set_stream(self.h, C.c_void_p(cuda_stream)))

    def wait_stream(self, cuda_stream: int):
        """The engine's stream waits for the work submitted so far to `cuda_stream` (0 = legacy default stream)."""
        self._check(self.lib.gcr_wait_stream(self.h, C.c_void_p(cuda_stream)))

    def set_sm_share(self, share: int):
        """Size every launch for 1/share of the SMs: `share` engines on `share` streams then run side by side (image batches)."""
        self._check(self.lib.gcr_set_sm_share(self.h, int(share)))

    def set_problem(self, problem: int):
        """PROBLEM_HYBRID (default), PROBLEM_SCALE_ONLY or PROBLEM_SCALE_ONLY_ORIGINAL; call before set_features."""
        self._check(self.lib.gcr_set_problem(self.h, problem))
        self.problem = problem
        self.ns = self.no = 0

    def set_features(self, scale: np.ndarray, orient: np.ndarray = None, host_libm: bool = False):
        """Host float64 [N,3] arrays (copied to the device inside the call); orient=None for scale-only problems.
        host_libm: GCR_FEATURES_HOST_LIBM, the bit-exact replay mode (solver columns evaluated by the host's C library)."""
        scale = np.ascontiguousarray(scale, np.float64)
        if orient is None:
            orient = np.zeros((0, 3))
        orient = np.ascontiguousarray(orient, np.float64)
        if scale.ndim != 2 or orient.ndim != 2 or scale.shape[1] != 3 or orient.shape[1] != 3:
            raise ValueError("features must be [N,3] arrays")
        self._keep = (scale, orient)
        self.ns, self.no = len(scale), len(orient)
        self.fetch_ns, self.fetch_no = self.ns, self.no
        self._check(self.lib.gcr_set_features(self.h, scale.ctypes.data, self.ns, orient.ctypes.data if self.no else None,
                                              self.no, FEATURES_HOST_LIBM if host_libm else 0))

    def set_keypoints(self, keypoints=None, device_ptr: int = None, n: int = None, is_f32: bool = None, want_sizes=True):
        """Keypoint rows (x, y, size, angle_deg) -> resident scale / orientation features (examples/utils.py:5-49 on the
        device).  Either a host array [n,4] (float32 or float64) or a device pointer with n and is_f32.
        Returns (n_scale, n_orient, sizes f64[n_orient] or None)."""
        if device_ptr is None:
            keypoints = np.asarray(keypoints)
            if keypoints.dtype != np.float32:
                keypoints = keypoints.astype(np.float64, copy=False)
            keypoints = np.ascontiguousarray(keypoints)
            if keypoints.ndim != 2 or keypoints.shape[1] != 4:
                raise ValueError("keypoints must be an [N,4] array of (x, y, size, angle in degrees)")
            n, is_f32, ptr, on_dev = len(keypoints), keypoints.dtype == np.float32, keypoints.ctypes.data, 0
        else:
            ptr, on_dev = device_ptr, 1
        ns, no = C.c_int64(0), C.c_int64(0)
        sizes = np.zeros(max(n, 1)) if want_sizes else None
        rc = self.lib.gcr_set_keypoints(self.h, C.c_void_p(ptr), n, 1 if is_f32 else 0, on_dev, C.byref(ns), C.byref(no),
                                        sizes.ctypes.data if want_sizes else None)
        self._check(rc)
        self.fetch_ns, self.fetch_no = int(ns.value), int(no.value)
        enough = self.fetch_ns >= (2 if self.problem == 0 else 3) and (self.problem != 0 or self.fetch_no >= 2)
        self.ns, self.no = (self.fetch_ns, self.fetch_no if self.problem == 0 else 0) if enough else (0, 0)
        return int(ns.value), int(no.value), (sizes[:no.value] if want_sizes else None)

    def get_features(self):
        """The resident feature rows as host arrays (scale f64[ns,3], orient f64[no,3])."""
        ns, no = getattr(self, "fetch_ns", self.ns), getattr(self, "fetch_no", self.no)
        scale, orient = np.zeros((ns, 3)), np.zeros((no, 3))
        self._check(self.lib.gcr_get_features(self.h, scale.ctypes.data, orient.ctypes.data if no else None))
        return scale, orient

    def set_features_device(self, scale_ptr: int, ns: int, orient_ptr: int, no: int):
        """Device pointers to float64 [N,3] row-major arrays (zero-copy; caller keeps them alive)."""
        self.ns, self.no = ns, no
        self.fetch_ns, self.fetch_no = ns, no
        self._check(self.lib.gcr_set_features(self.h, C.c_void_p(scale_ptr), ns, C.c_void_p(orient_ptr), no, 1))

    def run(self, n_hyp, thr, seed=0, first_hyp=0, samples=None, flags=0, record_ptr=None, sync=True,
            samples_ptr=None):
        thr_c = (C.c_double * 2)(float(thr[0]), float(thr[1]))
        sp = None
        if samples is not None:
            if np.asarray(samples).dtype == np.uint16:  # GCR_RUN_SAMPLES_U16: half the bytes of a replayed host array
                samples = np.ascontiguousarray(samples, np.uint16)
                flags |= RUN_SAMPLES_U16
            else:
                samples = np.ascontiguousarray(samples, np.int32)
            if samples.shape != (n_hyp, 4):
                raise ValueError("samples must be int32 or uint16 [n_hyp,4]")
            self._samples_keep = samples
            sp = C.c_void_p(samples.ctypes.data)
        elif samples_ptr is not None:
            sp = C.c_void_p(samples_ptr)
            flags |= RUN_SAMPLES_ON_DEVICE
        res = GcrResult()
        self.n_hyp = n_hyp
        rc = self.lib.gcr_run(self.h, n_hyp, seed, first_hyp, sp, thr_c, flags,
                              C.c_void_p(record_ptr) if record_ptr else None, C.byref(res) if sync else None)
        self._check(rc)
        return res if sync else None

    def peer_export(self) -> bytes:
        buf = C.create_string_buffer(IPC_HANDLE_BYTES)
        self._check(self.lib.gcr_peer_export(self.h, buf))
        return buf.raw

    def peer_attach(self, handles: bytes, world: int, rank: int):
        assert len(handles) == world * IPC_HANDLE_BYTES
        self._check(self.lib.gcr_peer_attach(self.h, handles, world, rank))

    def peer_detach(self):
        self._check(self.lib.gcr_peer_detach(self.h))

    def best(self):
        model = np.zeros(4)
        score = C.c_double()
        n_inl = np.zeros(2, np.uint32)
        sm = np.zeros((self.ns + 31) // 32, np.uint32)
        om = np.zeros((self.no + 31) // 32, np.uint32)
        idx = C.c_int64()
        self._check(self.lib.gcr_best(self.h, model.ctypes.data, C.byref(score), n_inl.ctypes.data, sm.ctypes.data,
                                      om.ctypes.data, C.byref(idx)))
        return dict(model=model, score=score.value, n_inl=n_inl, scale_mask=unpack_mask(sm, self.ns),
                    orient_mask=unpack_mask(om, self.no), index=idx.value, scale_words=sm, orient_words=om)

    def fetch(self, masks=False, samples=False):
        H = self.n_hyp
        W = (self.ns + 31) // 32 + (self.no + 31) // 32
        out = dict(flags=np.zeros(H, np.uint8), models=np.zeros((H, 4)), score=np.zeros(H),
                   n_inl=np.zeros((H, 2), np.uint32))
        out["masks"] = np.zeros((H, W), np.uint32) if masks else None
        out["samples"] = np.zeros((H, 4), np.int32) if samples else None
        self._check(self.lib.gcr_fetch(self.h, out["flags"].ctypes.data, out["models"].ctypes.data,
                                       out["score"].ctypes.data, out["n_inl"].ctypes.data,
                                       out["masks"].ctypes.data if masks else None,
                                       out["samples"].ctypes.data if samples else None))
        return out

    def score_model(self, model, thr):
        model = np.ascontiguousarray(model, np.float64)
        thr = np.ascontiguousarray(thr, np.float64)
        score = C.c_double()
        n_inl = np.zeros(2, np.uint32)
        sm = np.zeros((self.ns + 31) // 32, np.uint32)
        om = np.zeros((self.no + 31) // 32, np.uint32)
        self._check(self.lib.gcr_score_model(self.h, model.ctypes.data, thr.ctypes.data, C.byref(score),
                                             n_inl.ctypes.data, sm.ctypes.data, om.ctypes.data))
        return score.value, n_inl, unpack_mask(sm, self.ns), unpack_mask(om, self.no)

    def score_models(self, models, thr, truncated_sq=False, masks=False, strict=False):
        """Batched exact scoring of caller-supplied models f64[n,4] (n <= 1024)."""
        models = np.ascontiguousarray(models, np.float64).reshape(-1, 4)
        thr = np.ascontiguousarray(thr, np.float64)
        n = len(models)
        W = (self.ns + 31) // 32 + (self.no + 31) // 32
        score = np.zeros(n)
        n_inl = np.zeros((n, 2), np.uint32)
        words = np.zeros((n, W), np.uint32) if masks else None
        self._check(self.lib.gcr_score_models(self.h, models.ctypes.data, n, thr.ctypes.data, (1 if truncated_sq else 0) | (2 if strict else 0),
                                              score.ctypes.data, n_inl.ctypes.data, words.ctypes.data if masks else None))
        return score, n_inl, words

    def fit_models(self, scale_lists, orient_lists=None):
        """Device non-minimal fits from index lists (one list per fit and class); -> (models f64[K,4], ok bool[K])."""
        K = len(scale_lists)
        def pack(lists):
            off = np.zeros(K + 1, np.int64)
            off[1:] = np.cumsum([len(l) for l in lists])
            idx = np.ascontiguousarray(np.concatenate([np.asarray(l, np.uint32).ravel() for l in lists]) if off[-1] else np.zeros(1, np.uint32), np.uint32)
            return idx, off
        si, so = pack(scale_lists)
        if orient_lists is not None:
            assert len(orient_lists) == K
            oi, oo = pack(orient_lists)
        models = np.zeros((K, 4))
        ok = np.zeros(K, np.int32)
        self._check(self.lib.gcr_fit_models(self.h, K, si.ctypes.data, so.ctypes.data,
                                            oi.ctypes.data if orient_lists is not None else None,
                                            oo.ctypes.data if orient_lists is not None else None, models.ctypes.data, ok.ctypes.data))
        return models, ok.astype(bool)

    def fit_score_models(self, scale_lists, orient_lists, thr, truncated_sq=False, strict=False, masks=True):
        """gcr_fit_score_models: fits from index lists and their scores in one device pass -> (models, ok, score, n_inl, mask words)."""
        K = len(scale_lists)
        def pack(lists):
            off = np.zeros(K + 1, np.int64)
            off[1:] = np.cumsum([len(l) for l in lists])
            idx = np.ascontiguousarray(np.concatenate([np.asarray(l, np.uint32).ravel() for l in lists]) if off[-1] else np.zeros(1, np.uint32), np.uint32)
            return idx, off
        si, so = pack(scale_lists)
        if orient_lists is not None:
            oi, oo = pack(orient_lists)
        thr = np.ascontiguousarray(thr, np.float64)
        W = (self.ns + 31) // 32 + (self.no + 31) // 32
        models, ok, score = np.zeros((K, 4)), np.zeros(K, np.int32), np.zeros(K)
        n_inl = np.zeros((K, 2), np.uint32)
        words = np.zeros((K, W), np.uint32) if masks else None
        self._check(self.lib.gcr_fit_score_models(self.h, K, si.ctypes.data, so.ctypes.data, oi.ctypes.data if orient_lists is not None else None,
                                                  oo.ctypes.data if orient_lists is not None else None, thr.ctypes.data,
                                                  (1 if truncated_sq else 0) | (2 if strict else 0), models.ctypes.data, ok.ctypes.data,
                                                  score.ctypes.data, n_inl.ctypes.data, words.ctypes.data if masks else None))
        return models, ok.astype(bool), score, n_inl, words

    def lo_round(self, model, score, rng_state, thr, n_trials, lim, min_inl, label_flags=SCORE_TRUNCATED_SQ):
        """gcr_lo_round: one pass of the local optimisation's loop on the device -> (io struct, mask words of the new best or None)."""
        io = GcrLoRoundIo()
        for c in range(4):
            io.model[c] = float(model[c])
        io.score, io.rng_state = float(score), int(rng_state) & 0xFFFFFFFFFFFFFFFF
        thr = np.ascontiguousarray(thr, np.float64)
        lim_a, min_a = np.ascontiguousarray(lim, np.int32), np.ascontiguousarray(min_inl, np.int32)
        words = np.zeros((self.ns + 31) // 32 + (self.no + 31) // 32, np.uint32)
        self._check(self.lib.gcr_lo_round(self.h, C.byref(io), thr.ctypes.data, label_flags, int(n_trials), lim_a.ctypes.data, min_a.ctypes.data,
                                          words.ctypes.data))
        return io, (words if io.updated else None)

    def debug_eval(self, which, abc, model=None):
        abc = np.ascontiguousarray(abc, np.float64).reshape(-1, 3)
        out = np.zeros(len(abc))
        m = np.ascontiguousarray(model if model is not None else [0, 0, 1, 0], np.float64)
        self._check(self.lib.gcr_debug_eval(self.h, which, abc.ctypes.data, len(abc), m.ctypes.data, out.ctypes.data))
        return out
