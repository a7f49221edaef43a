"""Feature front end (reference examples/utils.py:5-49).

CPU: the numpy restatement (oracle/frontend_oracle.py) against the reference's own outputs (tests/golden/frontend.npz).
GPU: gcr_set_keypoints through the C ABI against those vectors (bit-exact: integer / copy work plus one FP64 product),
against the restatement at sizes the golden file does not hold, from device tensors, and into the estimator.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import frontend_oracle as fo  # noqa: E402


@pytest.fixture(scope="module")
def g(golden_dir):
    return np.load(os.path.join(golden_dir, "frontend.npz"))


def _sift_like(n, seed, grid=2000.0, dup=0.4):
    rng = np.random.Generator(np.random.PCG64(seed))
    xy = rng.uniform(-1.0, grid, (n, 2))
    src = rng.integers(0, n, n)
    d = rng.random(n) < dup
    xy[d] = xy[src[d]]
    ang = rng.uniform(0, 360, n)
    ang[rng.random(n) < 0.1] = -1.0
    return np.stack([xy[:, 0], xy[:, 1], rng.uniform(1.5, 40, n), ang], 1)


def test_oracle_restatement_matches_reference_golden(g):
    for k in range(int(g["n_cases"])):
        rows = g[f"case{k}_rows"]
        assert np.array_equal(fo.scale_features_from_rows(rows), g[f"case{k}_scale"]), k
        o, s = fo.orientation_features_from_rows(rows)
        assert np.array_equal(o, g[f"case{k}_orient"]) and np.array_equal(s, g[f"case{k}_sizes"]), k


def test_oracle_truncation_and_first_wins():
    rows = np.array([[0.5, 0.5, 1, 10], [-0.5, -0.9, 2, 20], [0.9, 0.1, 3, -1], [1.2, 0.0, 4, 30], [1.9, 0.7, 5, 40]], np.float64)
    s = fo.scale_features_from_rows(rows)
    assert np.array_equal(s, rows[[0, 3]][:, :3])  # int(-0.5) == int(0.5) == 0: one coordinate; first keypoint wins
    o, sz = fo.orientation_features_from_rows(rows)
    assert len(o) == 4 and np.array_equal(sz, [0.5, 1.0, 2.0, 2.5]) and o[0, 2] == np.deg2rad(10.0)


@pytest.mark.gpu
def test_device_front_end_matches_reference_golden(g):
    import gcr_capi
    e = gcr_capi.Engine(0)
    for k in range(int(g["n_cases"])):
        rows = g[f"case{k}_rows"]  # float32, as cv2.KeyPoint holds them
        for arr in (rows, rows.astype(np.float64)):
            ns, no, sizes = e.set_keypoints(arr)
            s, o = e.get_features()
            assert (ns, no) == (len(g[f"case{k}_scale"]), len(g[f"case{k}_orient"])), k
            assert np.array_equal(s, g[f"case{k}_scale"]), k
            assert np.array_equal(o, g[f"case{k}_orient"]), k
            assert np.array_equal(sizes, g[f"case{k}_sizes"]), k
    e.close()


@pytest.mark.gpu
def test_device_front_end_large_and_device_tensors():
    import torch
    import gcr_capi
    import gcr_frontend as gf
    e = gcr_capi.Engine(0)
    for n, seed, grid in ((200_000, 1, 3000.0), (1_000_000, 2, 600.0), (257, 3, 4.0)):
        rows = _sift_like(n, seed, grid)
        want_s = fo.scale_features_from_rows(rows)
        want_o, want_sz = fo.orientation_features_from_rows(rows)
        t = torch.from_numpy(rows).cuda()
        ns, no, sizes = gf.keypoints_to_engine(e, t)  # read in place from device memory
        s, o = e.get_features()
        assert np.array_equal(s, want_s) and np.array_equal(o, want_o) and np.array_equal(sizes, want_sz), n
        rows32 = rows.astype(np.float32)
        ns32, _, _ = gf.keypoints_to_engine(e, torch.from_numpy(rows32).cuda())
        assert np.array_equal(e.get_features()[0], fo.scale_features_from_rows(rows32))
    e.close()
    # the reference's two functions, same names and return values
    rows = _sift_like(5000, 9, 500.0)
    assert np.array_equal(gf.scale_features_from_sift(rows), fo.scale_features_from_rows(rows))
    o, sz = gf.orientation_features_from_sift(rows)
    assert np.array_equal(o, fo.orientation_features_from_rows(rows)[0]) and np.array_equal(sz, 0.5 * rows[rows[:, 3] != -1, 2])
    one = np.array([[3.0, 4.0, 2.0, -1.0]])
    assert gf.scale_features_from_sift(one).shape == (1, 3) and gf.orientation_features_from_sift(one)[0].shape == (0, 3)


@pytest.mark.gpu
def test_keypoints_straight_into_the_hypothesis_stage(port):
    """Keypoints built from a synthetic scene (duplicates added) -> gcr_set_keypoints -> gcr_run: same result as
    uploading the de-duplicated arrays with gcr_set_features."""
    import gcr_capi
    import gcr_scenes as gs
    sc = gs.config_scene("C1")
    rng = np.random.default_rng(3)
    n_s, n_o = len(sc.scale), len(sc.orient)
    # one keypoint per scale feature (no orientation) followed by one per orientation feature (size arbitrary), then
    # repeats of scale keypoints that must be dropped; coordinates of the scene are distinct pixels with probability ~1
    kp_s = np.column_stack([sc.scale, np.full(n_s, -1.0)])
    kp_o = np.column_stack([sc.orient[:, :2], rng.uniform(2, 9, n_o), np.rad2deg(sc.orient[:, 2])])
    rows = np.vstack([kp_s, kp_o, kp_s[rng.integers(0, n_s, 300)]])
    want_s = fo.scale_features_from_rows(rows)
    want_o, _ = fo.orientation_features_from_rows(rows)
    a, b = gcr_capi.Engine(0), gcr_capi.Engine(0)
    ns, no, _ = a.set_keypoints(rows)
    assert (ns, no) == (len(want_s), len(want_o))
    b.set_features(want_s, want_o)
    ra = a.run(20000, sc.thr, seed=5)
    rb = b.run(20000, sc.thr, seed=5)
    assert ra.best_index == rb.best_index and ra.best_score == rb.best_score
    assert tuple(ra.best_n_inl) == tuple(rb.best_n_inl) and ra.best_index >= 0
    a.close()
    b.close()
    e = gcr_capi.Engine(0)
    assert e.set_keypoints(rows[:1])[:2] == (1, 0)
    with pytest.raises(gcr_capi.GcrError):
        e.run(100, sc.thr)  # fewer rows than a minimal sample: nothing to run on
    e.close()
