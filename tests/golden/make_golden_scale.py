"""Golden vectors for the scale-only estimators (findRectifyingHomographyScaleOnly / ...Original), produced by the
REFERENCE ITSELF (oracle/_ref/libgcr_ref.so: unmodified three_sift.hpp / three_sift_original.hpp through
ScaleBasedRectifyingHomographyEstimator and MSACScoringFunction).  Build container only:

    python tests/golden/make_golden_scale.py
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "graph-cut-ransac_b200"))

import gcr_scenes as gs  # noqa: E402
from oracle_lib import Oracle  # noqa: E402


def main():
    ref = Oracle("reference")
    out = {}
    sc = gs.make_scene(331, 40, 0.4, 4242)
    scale = sc.scale.copy()
    scale[3] = scale[2]          # duplicate feature
    scale[5, 2] = 0.0            # zero scale
    scale[6, 2] = -1.5           # negative scale -> NaN rows
    scale[7, :2] = [2e6, -3e6]   # far beyond the horizon of most models
    smp = gs.draw_samples_scale3(700, len(scale), 21)
    smp[:6, :3] = [[2, 3, 9], [3, 2, 9], [5, 1, 0], [0, 6, 1], [7, 8, 9], [10, 11, 12]]
    # three nearly collinear positions (|signed distance| around the 1.0 tolerance)
    scale[20, :2], scale[21, :2], scale[22, :2] = [100.0, 100.0], [900.0, 500.0], [500.0, 300.9]
    smp[6, :3], smp[7, :3] = [20, 21, 22], [21, 20, 22]
    out.update(scale=scale, samples=smp, thr=np.float64(sc.thr[0]))
    for name, original in (("std", False), ("orig", True)):
        r = ref.replay_scale(scale, smp, sc.thr[0], original=original)
        for k in ("flags", "models", "score", "n_inl", "masks"):
            out[f"{name}_{k}"] = r[k]
        out[f"{name}_solved"] = np.int64(((r["flags"] & 2) > 0).sum())
    # full drivers on a clean scene (statistical comparison only: random_device sampler)
    clean = gs.make_scene(600, 10, 0.5, 555)
    out["run_scale"], out["run_truth"], out["run_thr"] = clean.scale, clean.model, np.float64(clean.thr[0])
    for name, original in (("std", False), ("orig", True)):
        r = ref.find_scale_only(clean.scale, clean.thr[0], 3000, 3000, original=original)
        out[f"run_{name}_n"], out[f"run_{name}_mask"], out[f"run_{name}_model6"], out[f"run_{name}_H"] = (
            np.int64(r["n_inliers"]), r["mask"], r["model"], r["H"])
    # non-minimal fits on subsets of the clean scene's inliers
    rng = np.random.Generator(np.random.PCG64(8))
    inl = np.flatnonzero(clean.scale_is_inlier)
    for k, n in enumerate((4, 21, 120, len(inl))):
        rows = clean.scale[rng.choice(inl, n, replace=False)]
        out[f"fit{k}_rows"] = rows
        for name, original in (("std", False), ("orig", True)):
            m = ref.fit_nonminimal_scale(rows, original)
            out[f"fit{k}_{name}_ok"] = np.int64(m is not None)
            out[f"fit{k}_{name}_model6"] = m if m is not None else np.zeros(6)
    out["n_fits"] = np.int64(4)
    np.savez_compressed(os.path.join(HERE, "scale_only.npz"), **out)
    print({k: (int(v) if np.ndim(v) == 0 and k.endswith(("solved", "_n")) else None) for k, v in out.items() if k.endswith(("solved", "_n"))})
    print("std model", out["run_std_model6"], "orig model", out["run_orig_model6"], "truth", clean.model)


if __name__ == "__main__":
    main()
