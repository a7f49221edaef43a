"""The reference-facing host side: the `pygcransac` module built from graph-cut-ransac_b200/csrc/host/.

CPU part (no device): module surface equals the reference's (src/bindings.cpp:315-398), model-class maps
against the oracle's known answers, host non-minimal fit / homography / iteration count against golden vectors
produced by the reference itself (tests/golden/make_golden_host.py), argument errors, and that a run
without a GPU fails loudly instead of falling back.
GPU part (-m gpu): the full estimator through the module against the oracle and the reference's own results.
"""
import math
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def pg():
    import __graft_entry__ as g
    g.build()
    import pygcransac
    return pygcransac


@pytest.fixture(scope="module")
def host_golden(golden_dir):
    return np.load(os.path.join(golden_dir, "host_refit.npz"))


def test_module_surface_matches_reference(pg):
    for name in ("NormalizingTransform", "RectifyingHomography", "ScaleBasedRectifyingHomography",
                 "OrientationBasedRectifyingHomography", "SIFTRectifyingHomography", "findRectifyingHomographySIFT"):
        assert hasattr(pg, name), name
    m = pg.SIFTRectifyingHomography()
    assert (m.x0, m.y0, m.s, m.h7, m.h8, m.alpha, m.phi) == (0.0, 0.0, 1.0, 0.0, 0.0, 1.0, 0.0)  # model.h defaults
    assert isinstance(m, pg.ScaleBasedRectifyingHomography) and isinstance(m, pg.OrientationBasedRectifyingHomography)
    assert isinstance(m, pg.RectifyingHomography) and isinstance(m, pg.NormalizingTransform)
    for meth in ("rectifiedScale", "unrectifiedScale", "rectifiedAngle", "unrectifiedAngle", "rectifiedPoint",
                 "unrectifiedPoint", "getHomography"):
        assert hasattr(m, meth)
    assert b"sm_100a" in pg.__backend__.encode()


def test_model_maps_match_oracle(pg, port):
    rng = np.random.default_rng(5)
    m = pg.RectifyingHomography()
    for _ in range(200):
        m.h7, m.h8 = rng.uniform(-5e-4, 5e-4, 2)
        x, y = rng.uniform(0, 4000), rng.uniform(0, 3000)
        v = rng.uniform(0.5, 50.0)
        a = rng.uniform(-7.0, 7.0)
        assert m.rectifiedScale(x, y, v) == port.model_map(0, m.h7, m.h8, x, y, v)
        assert m.unrectifiedScale(x, y, v) == port.model_map(1, m.h7, m.h8, x, y, v)
        assert m.rectifiedAngle(x, y, a) == port.model_map(2, m.h7, m.h8, x, y, a)
        assert m.unrectifiedAngle(x, y, a) == port.model_map(3, m.h7, m.h8, x, y, a)
        assert m.rectifiedPoint(x, y) == port.model_point(0, m.h7, m.h8, x, y)
        assert m.unrectifiedPoint(x, y) == port.model_point(1, m.h7, m.h8, x, y)
    # rectify / unrectify round trip of the reference's own unit test (tests/unit_tests.cpp:249-282)
    m.h7, m.h8 = 3e-4, -2e-4
    ux, uy = m.rectifiedPoint(1234.0, 777.0)
    bx, by = m.unrectifiedPoint(ux, uy)
    assert abs(bx - 1234.0) < 1e-9 and abs(by - 777.0) < 1e-9
    t = m.unrectifiedAngle(ux, uy, 0.7)
    assert abs(m.rectifiedAngle(bx, by, t) - 0.7) < 1e-9


def test_homography_matches_reference_golden(pg, host_golden):
    for m7, H in zip(host_golden["hom_model7"], host_golden["hom_H"]):
        m = pg.RectifyingHomography()
        m.x0, m.y0, m.s, m.h7, m.h8 = m7[:5]
        assert np.allclose(m.getHomography(), H, rtol=1e-12, atol=1e-15)


def test_non_minimal_fit_matches_reference_golden(pg, host_golden):
    g = host_golden
    for k in range(int(g["n_fits"])):
        m = pg._fit_non_minimal(g[f"fit{k}_scale"], g[f"fit{k}_orient"])
        if not int(g[f"fit{k}_ok"]):
            assert m is None
            continue
        ref = g[f"fit{k}_model7"]
        assert (m.x0, m.y0, m.s) == (0.0, 0.0, 1.0)  # normalizePoints resets to the identity (two_sift.hpp:843-845)
        # vanishing line (h7, h8, 1) <= 1e-6 relative (north star); alpha, phi likewise
        assert np.allclose([m.h7, m.h8, m.alpha], ref[3:6], rtol=1e-6, atol=1e-12), (k, m.h7, m.h8, ref)
        assert abs(m.phi - ref[6]) < 1e-9


def test_iteration_number_matches_oracle(pg, port):
    import ctypes as C
    f = port.lib.gcr_o_iteration_number
    f.restype = C.c_uint64
    f.argtypes = [C.POINTER(C.c_uint32), C.c_long, C.c_long, C.c_double]
    for n0, n1, ns, no in [(1, 1, 2000, 500), (600, 150, 2000, 500), (2000, 500, 2000, 500), (2, 2, 75, 45), (40, 9, 203, 61)]:
        arr = (C.c_uint32 * 2)(n0, n1)
        assert pg._iteration_number(n0, n1, ns, no, 0.95) == f(arr, ns, no, 0.95)


def test_argument_errors_are_value_errors(pg):
    ok = np.zeros((4, 3))
    with pytest.raises(ValueError):
        pg.findRectifyingHomographySIFT(np.zeros(12), ok, 0.2, 0.05)  # ndim != 2 (bindings.cpp:199-202)
    with pytest.raises(ValueError):
        pg.findRectifyingHomographySIFT(np.zeros((1, 3)), ok, 0.2, 0.05)  # < 2 rows (:206-214)
    with pytest.raises(ValueError):
        pg.findRectifyingHomographySIFT(ok, np.zeros((4, 2)), 0.2, 0.05)  # != 3 columns (:215-224)
    with pytest.raises(ValueError):
        pg.findRectifyingHomographySIFT(ok, ok, 0.2, 0.05, samples=np.zeros((3, 3), np.int32))


def test_no_gpu_fails_loudly(pg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    sc = np.random.default_rng(0).uniform(1, 100, (10, 3))
    with pytest.raises(RuntimeError, match="gcr_b200"):
        pg.findRectifyingHomographySIFT(sc, sc, 0.2, 0.05, min_iteration_number=10, max_iteration_number=10)


# ---------------------------------------------------------------------------------------------- GPU
def _line_err(model, truth):
    return float(np.hypot(model.h7 - truth[0], model.h8 - truth[1]) / np.hypot(truth[0], truth[1]))


@pytest.mark.gpu
def test_full_estimator_replay_against_oracle(pg, port, golden_dir):
    """Replayed C1 samples: the raw best hypothesis is the reference's (golden), LO/refit only ever raise the
    score, and the returned masks are exactly the oracle's inlier sets of the returned model."""
    g = np.load(os.path.join(golden_dir, "hyp_stage_c1.npz"))
    H, sm, om, model, st = pg.findRectifyingHomographySIFT(g["scale"], g["orient"], float(g["thr"][0]), float(g["thr"][1]),
                                                           samples=g["samples"], seed=9, return_stats=True)
    assert st["iteration_number"] == 5000 and st["best_index"] == int(g["best"])
    assert st["best_raw_score"] == pytest.approx(float(g["score"][int(g["best"])]), rel=1e-12)
    assert st["final_score"] >= st["lo_score"] >= st["best_raw_score"]
    m4 = np.array([model.h7, model.h8, model.alpha, model.phi])
    val, n_inl, smask, omask = port.score_model(g["scale"], g["orient"], m4, g["thr"])
    assert np.array_equal(sm, smask) and np.array_equal(om, omask)
    assert val == pytest.approx(st["final_score"], rel=1e-9)
    assert sm.dtype == np.bool_ and sm.shape == (2000,) and om.shape == (500,)
    assert np.allclose(H, [[1, 0, 0], [0, 1, 0], [model.h7, model.h8, 1]])
    assert _line_err(model, g["truth"]) < 0.02
    # deterministic under a fixed seed
    H2, sm2, om2, model2 = pg.findRectifyingHomographySIFT(g["scale"], g["orient"], float(g["thr"][0]), float(g["thr"][1]),
                                                           samples=g["samples"], seed=9)
    assert np.array_equal(H, H2) and np.array_equal(sm, sm2) and np.array_equal(om, om2)


@pytest.mark.gpu
def test_full_estimator_against_reference_driver(pg, host_golden):
    """Device sampler + LO + refit vs the UNMODIFIED reference driver's stored results on the same scenes: the
    samplers differ (the reference's is random_device-seeded), so agreement is statistical: the same plane."""
    g = host_golden
    for i in range(int(g["n_runs"])):
        thr = g[f"run{i}_thr"]
        H, sm, om, model = pg.findRectifyingHomographySIFT(g[f"run{i}_scale"], g[f"run{i}_orient"], float(thr[0]), float(thr[1]),
                                                           min_iteration_number=3000, max_iteration_number=3000, seed=100 + i)
        ref7 = g[f"run{i}_model7"]
        truth = g[f"run{i}_truth"]
        ours, theirs = _line_err(model, truth), float(np.hypot(*(ref7[3:5] - truth[:2])) / np.hypot(*truth[:2]))
        assert ours < max(0.05, 2.0 * theirs), (i, ours, theirs)
        n_ref = int(g[f"run{i}_n"])
        assert sm.sum() + om.sum() >= 0.9 * n_ref
        agree = (sm == g[f"run{i}_smask"]).mean()
        assert agree > 0.9, (i, agree)


@pytest.mark.gpu
def test_device_tensors_zero_copy_and_none_result(pg):
    import torch
    import gcr_scenes as gs
    sc = gs.make_scene(700, 250, 0.4, 77)
    ts, to = torch.from_numpy(sc.scale).cuda(), torch.from_numpy(sc.orient).cuda()
    a = pg.findRectifyingHomographySIFT(sc.scale, sc.orient, sc.thr[0], sc.thr[1], min_iteration_number=2000,
                                        max_iteration_number=2000, seed=5)
    b = pg.findRectifyingHomographySIFT(ts, to, sc.thr[0], sc.thr[1], min_iteration_number=2000,
                                        max_iteration_number=2000, seed=5)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    with pytest.raises(ValueError):
        pg.findRectifyingHomographySIFT(ts.float(), to.float(), 0.2, 0.05)
    # pure noise: no model with more than a minimal inlier set -> (None, masks, masks, None), bindings.cpp:292-300
    rng = np.random.default_rng(3)
    ns = np.column_stack([rng.uniform(0, 4000, 40), rng.uniform(0, 3000, 40), np.exp(rng.uniform(0, 8, 40))])
    no = np.column_stack([rng.uniform(0, 4000, 40), rng.uniform(0, 3000, 40), rng.uniform(0, 2 * math.pi, 40)])
    H, sm, om, model = pg.findRectifyingHomographySIFT(ns, no, 1e-7, 1e-7, min_iteration_number=500, max_iteration_number=500, seed=1)
    assert H is None and model is None and not sm.any() and not om.any()


@pytest.mark.gpu
def test_device_tensors_are_ordered_after_their_producer_stream(pg):
    """Feature tensors written by an asynchronous torch kernel on a side stream (behind a long-running kernel, so that they do not
    exist yet when the call is made): the pooled engine runs on its own non-blocking stream and must wait for the producer
    (gcr_wait_stream; `stream=` defaults to torch's current stream, may be given explicitly, 0 = the legacy default stream)."""
    import torch
    import gcr_scenes as gs
    sc = gs.make_scene(900, 300, 0.4, 78)
    want = pg.findRectifyingHomographySIFT(sc.scale, sc.orient, sc.thr[0], sc.thr[1], min_iteration_number=3000,
                                           max_iteration_number=3000, seed=9)
    hs, ho = torch.from_numpy(sc.scale).pin_memory(), torch.from_numpy(sc.orient).pin_memory()
    big = torch.empty(256 << 20, dtype=torch.float32, device="cuda")
    side = torch.cuda.Stream()
    for explicit in (False, True, "legacy"):
        ts = torch.full((900, 3), float("nan"), dtype=torch.float64, device="cuda")
        to = torch.full((300, 3), float("nan"), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        if explicit == "legacy":
            for _ in range(20):
                big.mul_(1.0001)  # ~ms of work on the default stream in front of the copies
            ts.copy_(hs, non_blocking=True)
            to.copy_(ho, non_blocking=True)
            got = pg.findRectifyingHomographySIFT(ts, to, sc.thr[0], sc.thr[1], min_iteration_number=3000, max_iteration_number=3000,
                                                  seed=9, stream=0)
        else:
            with torch.cuda.stream(side):
                for _ in range(20):
                    big.mul_(1.0001)
                ts.copy_(hs, non_blocking=True)
                to.copy_(ho, non_blocking=True)
                kw = {"stream": side} if explicit else {}
                got = pg.findRectifyingHomographySIFT(ts, to, sc.thr[0], sc.thr[1], min_iteration_number=3000,
                                                      max_iteration_number=3000, seed=9, **kw)
        assert got[0] is not None and np.array_equal(want[0], got[0]) and np.array_equal(want[1], got[1]) and np.array_equal(want[2], got[2])
    torch.cuda.synchronize()


@pytest.mark.gpu
def test_adaptive_iteration_count(pg):
    """min < max: the loop stops at max(min, getIterationNumber(best inlier counts)) (GCRANSAC.h:286-287,480-483)."""
    import gcr_scenes as gs
    sc = gs.make_scene(1000, 300, 0.6, 11)
    *_, st = pg.findRectifyingHomographySIFT(sc.scale, sc.orient, sc.thr[0], sc.thr[1], min_iteration_number=256,
                                             max_iteration_number=200000, seed=3, return_stats=True)
    assert 256 <= st["iteration_number"] < 200000
    assert st["device_calls"] >= 1
