"""Seeded synthetic planar scenes for the hybrid rectifying-homography RANSAC.

The reference ships no data (its example notebook is missing), so this recipe *is* the workload
definition used by the tests and by bench.py.  Feature formats follow the reference's front end
(examples/utils.py:25,47): scale features are rows (x, y, scale), orientation features are rows
(x, y, theta[rad]); both float64 [N,3], C-contiguous.

Ground truth is a warping homography [[1,0,0],[0,1,0],[h7,h8,1]] with class scale ``alpha`` and
principal direction ``phi``.  Inliers are built from the exact zero-residual identities of the
reference's residuals (two_sift.hpp:621-665) plus Gaussian noise of sigma = threshold/3:

* scale:        s = ((1 - h7 x - h8 y) / alpha)^3 * exp(eps)           ->  residual = |eps|
* orientation:  theta = unrectifiedAngle(u, phi + m*pi/2 + eps) at the rectified point u of (x,y)
                (model.h:167-174; the round trip of tests/unit_tests.cpp:249-282)  ->  residual = |eps|

Outliers: uniform position, s = exp(U(ln 1.5, ln 60)), theta ~ U(0, 2 pi).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

IMAGE_W = 4000.0
IMAGE_H = 3000.0
DEFAULT_THR_SCALE = 0.2
DEFAULT_THR_ORIENT = math.radians(3.0)


@dataclass
class Scene:
    scale: np.ndarray  # float64 [Ns,3]  (x, y, s)
    orient: np.ndarray  # float64 [No,3]  (x, y, theta)
    scale_is_inlier: np.ndarray  # bool [Ns]
    orient_is_inlier: np.ndarray  # bool [No]
    model: np.ndarray  # float64 [4] (h7, h8, alpha, phi) ground truth
    thr: np.ndarray  # float64 [2] (scale threshold, orientation threshold)


def _draw_plane(rng: np.random.Generator):
    """h7,h8 ~ U(-4e-4, 4e-4) with 1 - h7 x - h8 y > 0.2 on the whole image (horizon well outside)."""
    while True:
        h7, h8 = rng.uniform(-4e-4, 4e-4, size=2)
        corners = np.array([[0.0, 0.0], [IMAGE_W, 0.0], [0.0, IMAGE_H], [IMAGE_W, IMAGE_H]])
        w = 1.0 - h7 * corners[:, 0] - h8 * corners[:, 1]
        if w.min() > 0.2:
            return float(h7), float(h8)


def _unrectified_angle(h7, h8, ux, uy, angle):
    # model.h:167-174 with clipAngle (math_utils.hpp:78-88)
    ct, st = np.cos(angle), np.sin(angle)
    numer = (ux * st - uy * ct) * h7 + st
    denom = (-ux * st + uy * ct) * h8 + ct
    a = np.fmod(np.arctan2(numer, denom), 2.0 * math.pi)
    return np.where(a < 0.0, a + 2.0 * math.pi, a)


def make_scene(
    n_scale: int,
    n_orient: int,
    inlier_ratio: float,
    seed: int,
    thr_scale: float = DEFAULT_THR_SCALE,
    thr_orient: float = DEFAULT_THR_ORIENT,
    alpha: float = 0.5,
) -> Scene:
    rng = np.random.Generator(np.random.PCG64(seed))
    h7, h8 = _draw_plane(rng)
    phi = float(rng.uniform(0.0, 0.5 * math.pi))

    # ---- scale class
    n_in_s = int(round(inlier_ratio * n_scale))
    xy = np.column_stack([rng.uniform(0.0, IMAGE_W, n_scale), rng.uniform(0.0, IMAGE_H, n_scale)])
    s = np.exp(rng.uniform(math.log(1.5), math.log(60.0), n_scale))
    w = 1.0 - h7 * xy[:n_in_s, 0] - h8 * xy[:n_in_s, 1]
    s[:n_in_s] = (w / alpha) ** 3 * np.exp(rng.normal(0.0, thr_scale / 3.0, n_in_s))
    scale = np.column_stack([xy, s])
    s_inl = np.zeros(n_scale, dtype=bool)
    s_inl[:n_in_s] = True

    # ---- orientation class
    n_in_o = int(round(inlier_ratio * n_orient))
    xy = np.column_stack([rng.uniform(0.0, IMAGE_W, n_orient), rng.uniform(0.0, IMAGE_H, n_orient)])
    theta = rng.uniform(0.0, 2.0 * math.pi, n_orient)
    x, y = xy[:n_in_o, 0], xy[:n_in_o, 1]
    w = 1.0 - h7 * x - h8 * y  # rectifyPoint, model.h:128-131
    ux, uy = x / w, y / w
    theta_r = phi + rng.integers(0, 4, n_in_o) * (0.5 * math.pi) + rng.normal(0.0, thr_orient / 3.0, n_in_o)
    theta[:n_in_o] = _unrectified_angle(h7, h8, ux, uy, theta_r)
    wu = h7 * ux + h8 * uy + 1.0  # unrectifyPoint, model.h:133-138 (returns to (x, y) up to round-off)
    xy[:n_in_o, 0], xy[:n_in_o, 1] = ux / wu, uy / wu
    orient = np.column_stack([xy, theta])
    o_inl = np.zeros(n_orient, dtype=bool)
    o_inl[:n_in_o] = True

    ps, po = rng.permutation(n_scale), rng.permutation(n_orient)
    return Scene(
        scale=np.ascontiguousarray(scale[ps]),
        orient=np.ascontiguousarray(orient[po]),
        scale_is_inlier=s_inl[ps],
        orient_is_inlier=o_inl[po],
        model=np.array([h7, h8, alpha, phi]),
        thr=np.array([thr_scale, thr_orient]),
    )


def draw_samples(n_hyp: int, n_scale: int, n_orient: int, seed: int) -> np.ndarray:
    """Seeded stand-in for the reference's per-class draw (GCRANSAC.h:53-80: head of a full shuffle =
    uniform ORDERED pair without replacement).  Returns int32 [H,4] = (s0, s1, o0, o1)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    out = np.empty((n_hyp, 4), dtype=np.int32)
    for col, n in ((0, n_scale), (2, n_orient)):
        a = rng.integers(0, n, n_hyp)
        b = rng.integers(0, n - 1, n_hyp)
        b += b >= a
        out[:, col] = a
        out[:, col + 1] = b
    return out


def draw_samples_scale3(n_hyp: int, n_scale: int, seed: int) -> np.ndarray:
    """Seeded ordered triples without replacement for the scale-only estimators: int32 [H,4] = (s0, s1, s2, -1)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    a = rng.integers(0, n_scale, n_hyp)
    b = rng.integers(0, n_scale - 1, n_hyp)
    b += b >= a
    c = rng.integers(0, n_scale - 2, n_hyp)
    lo, hi = np.minimum(a, b), np.maximum(a, b)
    c += c >= lo
    c += c >= hi
    return np.column_stack([a, b, c, np.full(n_hyp, -1)]).astype(np.int32)


# The named BASELINE.json configurations: sizes, inlier ratio, scene seed.
CONFIGS = {
    "C1": dict(n_scale=2000, n_orient=500, inlier_ratio=0.30, seed=1234, n_hyp=5000),
    "C2": dict(n_scale=2000, n_orient=500, inlier_ratio=0.30, seed=1234, n_hyp=100_000),
    "C3": dict(n_scale=50_000, n_orient=20_000, inlier_ratio=0.10, seed=2025, n_hyp=1_000_000),
    "C4": dict(n_scale=4000, n_orient=1000, inlier_ratio=0.30, seed=10_000, n_hyp=100_000, n_images=256),
    "C5": dict(n_scale=2000, n_orient=500, inlier_ratio=0.30, seed=777, n_hyp=1_000_000),
}
SAMPLE_SEED = 4321


def config_scene(name: str, image: int = 0, **over) -> Scene:
    c = dict(CONFIGS[name])
    c.update(over)
    return make_scene(c["n_scale"], c["n_orient"], c["inlier_ratio"], c["seed"] + image,
                      **{k: v for k, v in c.items() if k in ("thr_scale", "thr_orient")})
