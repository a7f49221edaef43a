"""Generates the golden vectors under tests/golden/ by running the REFERENCE ITSELF.

Run in the build container only (needs /root/reference to build oracle/_ref/libgcr_ref.so, the
unmodified reference compiled against oracle/shim/):

    python tests/golden/make_golden.py

The reference has no fixtures or tests for its hypothesis stage, so these files --
outputs of the reference's own isValidSample/estimateModel/getScore/isValidModel on seeded inputs --
are what pins oracle/gcr_oracle.c (tests/test_oracle_golden.py) and, through it, the CUDA path.
Inputs are stored in the fixtures (not regenerated), so numpy RNG drift cannot invalidate them.
"""
from __future__ import annotations

import hashlib
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "graph-cut-ransac_b200"))

import gcr_scenes as gs  # noqa: E402
from oracle_lib import Oracle  # noqa: E402


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def edge_scene():
    """Hand-built awkward inputs: duplicates, parallel/identical orientation samples, features on and
    beyond the horizon of typical hypotheses, non-positive scales, ragged sizes (not multiples of 32)."""
    rng = np.random.Generator(np.random.PCG64(99))
    sc = gs.make_scene(75, 45, 0.4, 31337)
    scale, orient = sc.scale.copy(), sc.orient.copy()
    scale[3] = scale[2]  # duplicate scale feature
    scale[5, 2] = 0.0  # zero scale -> pow(0,1/3)=0 row / rectified scale 0 -> DBL_MAX residual
    scale[6, 2] = -2.0  # negative scale -> pow() NaN in the solver, negative rectified scale
    scale[7, :2] = [1e6, -1e6]  # far outside the image: beyond the horizon of most models
    scale[8, 2] = 1e-12
    scale[9, 2] = 1e12
    orient[4] = orient[1]  # identical orientation features -> zero vanishing point
    orient[6, 2] = orient[5, 2]  # parallel lines -> vanishing point at infinity
    orient[7, :2] = orient[5, :2]  # same position, different angle -> vp at the feature itself
    orient[8, 2] = orient[5, 2] + math.pi  # same line direction modulo pi
    orient[9, 2] = 7.5 * math.pi  # angle outside [0, 2 pi)
    orient[10, 2] = -3.25 * math.pi
    H = 600
    smp = gs.draw_samples(H, len(scale), len(orient), 5)
    # force the awkward rows into many samples
    forced = np.array([
        [2, 3, 0, 1], [3, 2, 0, 1], [0, 1, 1, 4], [0, 1, 4, 1], [0, 1, 5, 6], [0, 1, 6, 5],
        [0, 1, 5, 7], [0, 1, 5, 8], [5, 0, 2, 3], [0, 5, 2, 3], [6, 0, 2, 3], [0, 6, 2, 3],
        [7, 0, 2, 3], [8, 9, 2, 3], [9, 8, 3, 2], [0, 1, 9, 10], [0, 1, 10, 9], [7, 8, 9, 10],
    ], dtype=np.int32)
    smp[: len(forced)] = forced
    extra = rng.integers(0, 11, size=(100, 4)).astype(np.int32)
    for r in extra:
        if r[0] == r[1]:
            r[1] = (r[1] + 1) % 11
        if r[2] == r[3]:
            r[3] = (r[3] + 1) % 11
    smp[len(forced): len(forced) + 100] = extra
    return scale, orient, smp, sc.thr


def main():
    ref = Oracle("reference")
    out = {}

    # 1. small generic scene, everything stored (masks included)
    sc = gs.make_scene(203, 61, 0.35, 7)
    smp = gs.draw_samples(512, 203, 61, 11)
    r = ref.replay(sc.scale, sc.orient, smp, sc.thr)
    np.savez_compressed(os.path.join(HERE, "hyp_stage_small.npz"), scale=sc.scale, orient=sc.orient,
                        samples=smp, thr=sc.thr, flags=r["flags"], models=r["models"], score=r["score"],
                        n_inl=r["n_inl"], masks=r["masks"], best=np.int64(ref.select_best(r["score"], r["flags"])))
    out["small"] = (r["flags"] & 2).astype(bool).sum()

    # 2. edge cases
    scale, orient, smp, thr = edge_scene()
    r = ref.replay(scale, orient, smp, thr)
    np.savez_compressed(os.path.join(HERE, "hyp_stage_edge.npz"), scale=scale, orient=orient, samples=smp,
                        thr=thr, flags=r["flags"], models=r["models"], score=r["score"], n_inl=r["n_inl"],
                        masks=r["masks"], best=np.int64(ref.select_best(r["score"], r["flags"])))
    out["edge"] = (r["flags"] & 2).astype(bool).sum()

    # 3. threshold sweep on the small scene (C5 style: level p -> thr_o = p deg, thr_s = 0.05 p)
    sweep = {}
    for p in (1, 2, 4, 8, 20):
        thr = np.array([0.05 * p, math.radians(p)])
        rr = ref.replay(sc.scale, sc.orient, gs.draw_samples(128, 203, 61, 100 + p), thr)
        sweep[f"p{p}_thr"] = thr
        sweep[f"p{p}_samples"] = gs.draw_samples(128, 203, 61, 100 + p)
        for k in ("flags", "models", "score", "n_inl", "masks"):
            sweep[f"p{p}_{k}"] = rr[k]
    np.savez_compressed(os.path.join(HERE, "hyp_stage_sweep.npz"), scale=sc.scale, orient=sc.orient, **sweep)

    # 4. BASELINE config C1 (2000 + 500 features, 5000 replayed samples): inputs stored, per-hypothesis
    #    outputs stored, masks as a digest + per-hypothesis popcounts (full masks would be 1.6 MB)
    c1 = gs.config_scene("C1")
    smp = gs.draw_samples(5000, 2000, 500, gs.SAMPLE_SEED)
    r = ref.replay(c1.scale, c1.orient, smp, c1.thr)
    ws = (2000 + 31) // 32
    pop = np.array([[int(np.unpackbits(m[:ws].view(np.uint8)).sum()), int(np.unpackbits(m[ws:].view(np.uint8)).sum())]
                    for m in r["masks"]], dtype=np.uint32)
    best = ref.select_best(r["score"], r["flags"])
    np.savez_compressed(os.path.join(HERE, "hyp_stage_c1.npz"), scale=c1.scale, orient=c1.orient, samples=smp,
                        thr=c1.thr, truth=c1.model, flags=r["flags"], models=r["models"], score=r["score"],
                        n_inl=r["n_inl"], mask_popcounts=pop, masks_sha256=np.array(sha(r["masks"])),
                        best=np.int64(best), best_masks=r["masks"][best])
    out["c1"] = (r["flags"] & 2).astype(bool).sum()

    # 5. residual tables of the reference for a handful of models (pins the per-feature arithmetic)
    models = np.vstack([c1.model, r["models"][best], r["models"][(r["flags"] & 2) > 0][:6]])
    rs = np.zeros((len(models), 2000))
    ro = np.zeros((len(models), 500))
    for i, m in enumerate(models):
        rs[i], ro[i] = ref.residuals(c1.scale, c1.orient, m)
    np.savez_compressed(os.path.join(HERE, "residuals_c1.npz"), models=models, rs=rs, ro=ro)
    print("golden vectors written; solved hypotheses per fixture:", out)


if __name__ == "__main__":
    main()
