"""Golden vectors for the host side (local optimisation / refit helpers and the full driver), produced by the
REFERENCE ITSELF through oracle/_ref/libgcr_ref.so.  Build container only:

    python tests/golden/make_golden_host.py

* non-minimal fits (estimateModelNonminimal, rectifying_homography_estimator.h:164-227) of several row subsets;
  the reference's linear solve runs through oracle/shim's QR, so these pin row construction, the phi histogram
  and the solution to tolerance level;
* getHomography (model.h:211-226) for models with a non-trivial normalising transform;
* the unmodified driver findRectifyingHomographySIFT_ (gcransac_python.cpp:256-406) on small seeded scenes: its
  sampler is random_device-seeded, so only the recovered model / inlier sets are kept, for statistical comparison.
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
    rng = np.random.Generator(np.random.PCG64(2024))
    # 1. non-minimal fits
    sc = gs.make_scene(400, 160, 0.5, 17)
    si_all = np.flatnonzero(sc.scale_is_inlier)
    oi_all = np.flatnonzero(sc.orient_is_inlier)
    cases = [(14, 14), (3, 3), (2, 5), (40, 9), (len(si_all), len(oi_all)), (25, 2)]
    for k, (a, b) in enumerate(cases):
        s_rows = sc.scale[rng.choice(si_all, a, replace=False)]
        o_rows = sc.orient[rng.choice(oi_all, b, replace=False)]
        if k == 3:  # mix in outliers
            s_rows = np.vstack([s_rows, sc.scale[~sc.scale_is_inlier][:6]])
            o_rows = np.vstack([o_rows, sc.orient[~sc.orient_is_inlier][:3]])
        m = ref.fit_nonminimal(s_rows, o_rows)
        # the reference's linear solve ran through oracle/shim's QR: cross-check every golden fit against LAPACK (tests/lsq_independent.py)
        import lsq_independent
        x = lsq_independent.fit(s_rows, o_rows)
        assert (x is None) == (m is None) and (x is None or np.allclose(x, m[3:6], rtol=1e-10, atol=0.0)), (k, x, m)
        out[f"fit{k}_scale"] = s_rows
        out[f"fit{k}_orient"] = o_rows
        out[f"fit{k}_ok"] = np.int64(m is not None)
        out[f"fit{k}_model7"] = m if m is not None else np.zeros(7)
    out["n_fits"] = np.int64(len(cases))
    # 2. homographies
    m7 = np.array([[0, 0, 1, 3e-4, -2e-4, 0.5, 1.0], [1500.0, 900.0, 1e-3, 0.2, -0.1, 0.7, 0.3],
                   [-20.0, 33.0, 0.05, -4e-3, 1e-3, 1.2, 2.0]])
    out["hom_model7"] = m7
    out["hom_H"] = np.stack([ref.homography(m) for m in m7])
    # 3. full driver on small scenes
    runs = []
    for i, (ns, no, ratio, seed) in enumerate([(500, 150, 0.4, 301), (300, 120, 0.3, 302), (800, 200, 0.5, 303)]):
        s = gs.make_scene(ns, no, ratio, seed)
        r = ref.find_sift(s.scale, s.orient, s.thr[0], s.thr[1], 3000, 3000)
        out[f"run{i}_scale"], out[f"run{i}_orient"], out[f"run{i}_thr"], out[f"run{i}_truth"] = s.scale, s.orient, s.thr, s.model
        out[f"run{i}_model7"], out[f"run{i}_H"] = r["model"], r["H"]
        out[f"run{i}_smask"], out[f"run{i}_omask"] = r["scale_mask"], r["orient_mask"]
        out[f"run{i}_n"] = np.int64(r["n_inliers"])
        runs.append((r["n_inliers"], r["model"][3:5], s.model[:2]))
    out["n_runs"] = np.int64(3)
    np.savez_compressed(os.path.join(HERE, "host_refit.npz"), **out)
    print("host golden written:", runs)


if __name__ == "__main__":
    main()
