"""ctypes bindings of the CPU checker libraries (TEST INFRASTRUCTURE; see oracle/gcr_oracle.c).

``Oracle("port")``      -> oracle/libgcr_oracle.so   (plain-C restatement, symbols gcr_o_*)
``Oracle("reference")`` -> oracle/_ref/libgcr_ref.so (the unmodified reference built on shims, gcr_ref_*)

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
PORT_SO = os.path.join(ORACLE_DIR, "libgcr_oracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libgcr_ref.so")

FLAG_SAMPLE_VALID, FLAG_SOLVED, FLAG_MODEL_VALID = 1, 2, 4

_dp = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u8 = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_u32 = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")


def build_oracle(force: bool = False) -> None:
    """Compile the checker libraries (building the checker is not using it)."""
    src = os.path.join(ORACLE_DIR, "gcr_oracle.c")
    if force or not os.path.exists(PORT_SO) or os.path.getmtime(PORT_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "libgcr_oracle.so"], stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference/src/pygcransac/include"):
        hs = os.path.join(ORACLE_DIR, "ref_harness.cpp")
        if force or not os.path.exists(REF_SO) or os.path.getmtime(REF_SO) < os.path.getmtime(hs):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "ref"], stdout=subprocess.DEVNULL)


def have_reference() -> bool:
    return os.path.exists(REF_SO)


class Oracle:
    def __init__(self, kind: str = "port"):
        assert kind in ("port", "reference")
        self.kind = kind
        if kind == "port":
            build_oracle()
            self.lib = C.CDLL(PORT_SO)
            self.p = "gcr_o_"
        else:
            if not have_reference():
                build_oracle()
            self.lib = C.CDLL(REF_SO)
            self.p = "gcr_ref_"
        L = self.lib
        rp = getattr(L, self.p + "replay")
        rp.restype = C.c_int
        rp.argtypes = [_dp, C.c_long, _dp, C.c_long, _ip, C.c_long, _dp, _u8, _dp, _dp, _u32, C.c_void_p, C.c_int]
        self._replay = rp

    # ------------------------------------------------------------------ hypothesis stage
    def replay(self, scale, orient, samples, thr, want_masks=True, nthreads=1):
        scale = np.ascontiguousarray(scale, np.float64)
        orient = np.ascontiguousarray(orient, np.float64)
        samples = np.ascontiguousarray(samples, np.int32)
        thr = np.ascontiguousarray(thr, np.float64)
        ns, no, H = len(scale), len(orient), len(samples)
        ws, wo = (ns + 31) // 32, (no + 31) // 32
        flags = np.zeros(H, np.uint8)
        models = np.zeros((H, 4), np.float64)
        score = np.zeros(H, np.float64)
        n_inl = np.zeros((H, 2), np.uint32)
        masks = np.zeros((H, ws + wo), np.uint32) if want_masks else None
        used = self._replay(scale, ns, orient, no, samples, H, thr, flags, models, score, n_inl,
                            masks.ctypes.data if want_masks else None, nthreads)
        return dict(flags=flags, models=models, score=score, n_inl=n_inl, masks=masks, threads=used)

    def select_best(self, score, flags) -> int:
        """GCRANSAC.h:440-446 selection rule (same few lines in either library)."""
        if self.kind == "port":
            f = self.lib.gcr_o_select_best
            f.restype = C.c_long
            f.argtypes = [_dp, _u8, C.c_long]
            return int(f(np.ascontiguousarray(score), np.ascontiguousarray(flags), len(score)))
        best, idx = 0.0, -1
        for h in range(len(score)):
            if flags[h] & FLAG_SOLVED and best < score[h] and flags[h] & FLAG_MODEL_VALID:
                best, idx = score[h], h
        return idx

    def score_model(self, scale, orient, model, thr):
        scale = np.ascontiguousarray(scale, np.float64)
        orient = np.ascontiguousarray(orient, np.float64)
        model = np.ascontiguousarray(model, np.float64)
        thr = np.ascontiguousarray(thr, np.float64)
        ns, no = len(scale), len(orient)
        n_inl = np.zeros(2, np.uint32)
        if self.kind == "port":
            f = self.lib.gcr_o_score
            f.restype = C.c_double
            f.argtypes = [_dp, C.c_long, _dp, C.c_long, _dp, _dp, _u32, _u32, _u32]
            sm = np.zeros((ns + 31) // 32, np.uint32)
            om = np.zeros((no + 31) // 32, np.uint32)
            val = f(scale, ns, orient, no, model, thr, n_inl, sm, om)
            smask = unpack_mask(sm, ns)
            omask = unpack_mask(om, no)
        else:
            f = self.lib.gcr_ref_score_model
            f.restype = None
            f.argtypes = [_dp, C.c_long, _dp, C.c_long, _dp, _dp, _dp, _u32, _u8, _u8]
            v = np.zeros(1)
            sm = np.zeros(ns, np.uint8)
            om = np.zeros(no, np.uint8)
            f(scale, ns, orient, no, model, thr, v, n_inl, sm, om)
            val, smask, omask = float(v[0]), sm.astype(bool), om.astype(bool)
        return float(val), n_inl, smask, omask

    def residuals(self, scale, orient, model):
        scale = np.ascontiguousarray(scale, np.float64)
        orient = np.ascontiguousarray(orient, np.float64)
        model = np.ascontiguousarray(model, np.float64)
        rs, ro = np.zeros(len(scale)), np.zeros(len(orient))
        if self.kind == "reference":
            f = self.lib.gcr_ref_residuals
            f.restype = None
            f.argtypes = [_dp, C.c_long, _dp, C.c_long, _dp, _dp, _dp]
            f(scale, len(scale), orient, len(orient), model, rs, ro)
        else:
            fs, fo = self.lib.gcr_o_scale_residual, self.lib.gcr_o_orientation_residual
            for f in (fs, fo):
                f.restype = C.c_double
                f.argtypes = [C.c_double] * 3 + [_dp]
            for i, r in enumerate(scale):
                rs[i] = fs(r[0], r[1], r[2], model)
            for i, r in enumerate(orient):
                ro[i] = fo(r[0], r[1], r[2], model)
        return rs, ro

    # ------------------------------------------------------------------ helper math (known answers)
    def _f(self, name, restype, argtypes):
        f = getattr(self.lib, self.p + name)
        f.restype, f.argtypes = restype, argtypes
        return f

    def clip_angle(self, a):
        return self._f("clip_angle", C.c_double, [C.c_double])(a)

    def min_angle_diff(self, a, b):
        return self._f("min_angle_diff", C.c_double, [C.c_double] * 2)(a, b)

    def lines_angles_diff(self, a, b):
        return self._f("lines_angles_diff", C.c_double, [C.c_double] * 2)(a, b)

    def deg2rad(self, a):
        return self._f("deg2rad", C.c_double, [C.c_double])(a)

    def rad2deg(self, a):
        return self._f("rad2deg", C.c_double, [C.c_double])(a)

    def nchoose2(self, n):
        return int(self._f("nchoose2", C.c_uint64, [C.c_uint64])(n))

    def are_collinear(self, p1, p2, p3, tol):
        return bool(self._f("are_collinear", C.c_int, [C.c_double] * 7)(*p1, *p2, *p3, tol))

    def line_from_point_angle(self, x, y, t):
        out = np.zeros(3)
        self._f("line_from_point_angle", None, [C.c_double] * 3 + [_dp])(x, y, t, out)
        return out

    def convex_hull(self, pts):
        pts = np.ascontiguousarray(pts, np.float64)
        out = np.zeros((2 * len(pts), 2))
        n = self._f("convex_hull", C.c_int, [_dp, C.c_int, _dp])(pts, len(pts), out)
        return out[:n]

    def point_in_convex_polygon(self, x, y, poly):
        poly = np.ascontiguousarray(poly, np.float64)
        return bool(self._f("point_in_convex_polygon", C.c_int, [C.c_double, C.c_double, _dp, C.c_int])(x, y, poly, len(poly)))

    def gauss3(self, a):
        a = np.ascontiguousarray(a, np.float64).reshape(3, 4).copy()
        x = np.zeros(3)
        self._f("gauss3", None, [_dp, _dp])(a, x)
        return x

    def model_map(self, which, h7, h8, x, y, v):
        if self.kind == "reference":
            return self._f("model_map", C.c_double, [C.c_int] + [C.c_double] * 5)(which, h7, h8, x, y, v)
        name = ["rectified_scale", "unrectified_scale", "rectified_angle", "unrectified_angle"][which]
        return self._f(name, C.c_double, [C.c_double] * 5)(h7, h8, x, y, v)

    def model_point(self, which, h7, h8, x, y):
        cx, cy = C.c_double(x), C.c_double(y)
        if self.kind == "reference":
            self._f("model_point", None, [C.c_int, C.c_double, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)])(
                which, h7, h8, C.byref(cx), C.byref(cy))
        else:
            name = "rectify_point" if which == 0 else "unrectify_point"
            self._f(name, None, [C.c_double, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)])(
                h7, h8, C.byref(cx), C.byref(cy))
        return cx.value, cy.value

    # ------------------------------------------------------------------ port-only / reference-only extras
    def sample_philox(self, seed, first_hyp, H, ns, no):
        assert self.kind == "port"
        out = np.zeros((H, 4), np.int32)
        f = self.lib.gcr_o_sample_philox
        f.restype = None
        f.argtypes = [C.c_uint64, C.c_uint64, C.c_long, C.c_long, C.c_long, _ip]
        f(seed, first_hyp, H, ns, no, out)
        return out

    def find_sift(self, scale, orient, thr_s, thr_o, min_iters, max_iters, max_lo=50, scw=0.0):
        """The reference's full driver (its own random_device sampler, LO, refit). reference only."""
        assert self.kind == "reference"
        scale = np.ascontiguousarray(scale, np.float64)
        orient = np.ascontiguousarray(orient, np.float64)
        ns, no = len(scale), len(orient)
        sm, om = np.zeros(ns, np.uint8), np.zeros(no, np.uint8)
        Hm, model, sec = np.zeros(9), np.zeros(7), np.zeros(1)
        f = self.lib.gcr_ref_find_sift
        f.restype = C.c_int
        f.argtypes = [_dp, C.c_long, _dp, C.c_long, C.c_double, C.c_double, C.c_double,
                      C.c_ulong, C.c_ulong, C.c_ulong, _u8, _u8, _dp, _dp, _dp]
        n = f(scale, ns, orient, no, thr_s, thr_o, scw, min_iters, max_iters, max_lo, sm, om, Hm, model, sec)
        return dict(n_inliers=n, scale_mask=sm.astype(bool), orient_mask=om.astype(bool),
                    H=Hm.reshape(3, 3), model=model, seconds=float(sec[0]))


    # ------------------------------------------------------------------ scale-only estimators
    def replay_scale(self, scale, samples, thr, original=False, want_masks=True, nthreads=1):
        """Hypothesis stage of findRectifyingHomographyScaleOnly[Original] per replayed 3-sample (s0, s1, s2, ignored)."""
        scale = np.ascontiguousarray(scale, np.float64)
        samples = np.ascontiguousarray(samples, np.int32)
        ns, H = len(scale), len(samples)
        ws = (ns + 31) // 32
        flags, models, score = np.zeros(H, np.uint8), np.zeros((H, 4)), np.zeros(H)
        n_inl = np.zeros((H, 2), np.uint32)
        masks = np.zeros((H, ws), np.uint32) if want_masks else None
        f = getattr(self.lib, self.p + "replay_scale")
        f.restype = C.c_int
        f.argtypes = [_dp, C.c_long, _ip, C.c_long, C.c_double, C.c_int, _u8, _dp, _dp, _u32, C.c_void_p, C.c_int]
        used = f(scale, ns, samples, H, float(thr), int(original), flags, models, score, n_inl,
                 masks.ctypes.data if want_masks else None, nthreads)
        return dict(flags=flags, models=models, score=score, n_inl=n_inl, masks=masks, threads=used)

    def sample_philox_scale3(self, seed, first_hyp, H, ns):
        assert self.kind == "port"
        out = np.zeros((H, 4), np.int32)
        f = self.lib.gcr_o_sample_philox_scale3
        f.restype = None
        f.argtypes = [C.c_uint64, C.c_uint64, C.c_long, C.c_long, _ip]
        f(seed, first_hyp, H, ns, out)
        return out

    def score_scale(self, scale, model, thr, original=False):
        """One-class MSAC of one model (port only): value, count, bool mask."""
        assert self.kind == "port"
        scale = np.ascontiguousarray(scale, np.float64)
        model = np.ascontiguousarray(model, np.float64)
        ns = len(scale)
        n = np.zeros(1, np.uint32)
        sm = np.zeros((ns + 31) // 32, np.uint32)
        f = self.lib.gcr_o_score_scale
        f.restype = C.c_double
        f.argtypes = [_dp, C.c_long, _dp, C.c_double, C.c_int, _u32, _u32]
        val = f(scale, ns, model, float(thr), int(original), n, sm)
        return float(val), int(n[0]), unpack_mask(sm, ns)

    def find_scale_only(self, scale, thr, min_iters, max_iters, max_lo=50, scw=0.0, original=False):
        """The reference's full scale-only driver (reference only)."""
        assert self.kind == "reference"
        scale = np.ascontiguousarray(scale, np.float64)
        ns = len(scale)
        sm, Hm, model = np.zeros(ns, np.uint8), np.zeros(9), np.zeros(6)
        f = self.lib.gcr_ref_find_scale_only
        f.restype = C.c_int
        f.argtypes = [_dp, C.c_long, C.c_double, C.c_double, C.c_ulong, C.c_ulong, C.c_ulong, C.c_int, _u8, _dp, _dp]
        n = f(scale, ns, thr, scw, min_iters, max_iters, max_lo, int(original), sm, Hm, model)
        return dict(n_inliers=n, mask=sm.astype(bool), H=Hm.reshape(3, 3), model=model)

    def fit_nonminimal_scale(self, scale, original=False):
        assert self.kind == "reference"
        scale = np.ascontiguousarray(scale, np.float64)
        model = np.zeros(6)
        f = self.lib.gcr_ref_fit_nonminimal_scale
        f.restype = C.c_int
        f.argtypes = [_dp, C.c_long, C.c_int, _dp]
        return model if f(scale, len(scale), int(original), model) else None

    def fit_nonminimal(self, scale, orient):
        """estimateModelNonminimal on all rows given (reference only); None when the fit fails."""
        assert self.kind == "reference"
        scale = np.ascontiguousarray(scale, np.float64)
        orient = np.ascontiguousarray(orient, np.float64)
        model = np.zeros(7)
        f = self.lib.gcr_ref_fit_nonminimal
        f.restype = C.c_int
        f.argtypes = [_dp, C.c_long, _dp, C.c_long, _dp]
        return model if f(scale, len(scale), orient, len(orient), model) else None

    def homography(self, model7):
        assert self.kind == "reference"
        out = np.zeros(9)
        self._f("homography", None, [_dp, _dp])(np.ascontiguousarray(model7, np.float64), out)
        return out.reshape(3, 3)


def unpack_mask(words: np.ndarray, n: int) -> np.ndarray:
    """u32 words (bit j of word w <-> feature 32w+j) -> bool[n]."""
    bits = np.unpackbits(np.ascontiguousarray(words, np.uint32).view(np.uint8), bitorder="little")
    return bits[:n].astype(bool)
