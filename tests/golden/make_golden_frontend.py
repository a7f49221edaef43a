"""Golden vectors of the feature front end, produced by the REFERENCE's own examples/utils.py on cv2.KeyPoint lists.
Build container only (needs /root/reference and cv2):

    python tests/golden/make_golden_frontend.py
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference/examples")
import utils as ref_utils  # noqa: E402


def make_rows(n, seed, grid, dup_frac, no_angle_frac):
    """SIFT-like keypoints: float32 fields, several keypoints per coordinate (one per dominant orientation), some
    sharing only the integer pixel, a share without orientation (angle == -1)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    base = rng.uniform(-2.0, grid, size=(n, 2))
    size = rng.uniform(1.5, 40.0, n)
    ang = rng.uniform(0.0, 360.0, n)
    src = rng.integers(0, n, n)
    dup = rng.random(n) < dup_frac           # exact repeat of an earlier / later keypoint's position (other angle)
    base[dup] = base[src[dup]]
    size[dup] = size[src[dup]]
    ang[rng.random(n) < no_angle_frac] = -1.0
    return np.stack([base[:, 0], base[:, 1], size, ang], 1).astype(np.float32)


def main():
    out = {}
    cases = [make_rows(1, 1, 100, 0.0, 0.0), make_rows(7, 2, 3, 0.5, 0.3), make_rows(300, 3, 40, 0.3, 0.1),
             make_rows(6000, 4, 900, 0.35, 0.05), make_rows(4000, 5, 30, 0.2, 0.5)]
    cases[1][:, :2] = np.float32([[0.5, 0.5], [-0.5, -0.9], [0.2, -0.2], [1.0, 0.99], [1.5, 0.0], [0.5, 0.5], [-1.0, 0.3]])
    cases[1][2, 3] = -1.0
    for k, rows in enumerate(cases):
        kps = [cv2.KeyPoint(float(r[0]), float(r[1]), float(r[2]), float(r[3])) for r in rows]
        back = np.array([[kp.pt[0], kp.pt[1], kp.size, kp.angle] for kp in kps], np.float32)
        assert np.array_equal(back, rows)  # cv2 keeps the float32 fields unchanged
        s = ref_utils.scale_features_from_sift(kps)
        o, sizes = ref_utils.orientation_features_from_sift(kps)
        out[f"case{k}_rows"] = rows
        out[f"case{k}_scale"] = np.asarray(s, np.float64).reshape(-1, 3)
        out[f"case{k}_orient"] = np.asarray(o, np.float64).reshape(-1, 3)
        out[f"case{k}_sizes"] = np.asarray(sizes, np.float64).reshape(-1)
    out["n_cases"] = np.int64(len(cases))
    np.savez_compressed(os.path.join(HERE, "frontend.npz"), **out)
    print({k: v.shape for k, v in out.items() if k.endswith(("scale", "orient"))})


if __name__ == "__main__":
    main()
