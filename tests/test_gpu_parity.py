"""GPU parity tests (run on the B200 box with ``-m gpu``): the CUDA hypothesis stage, called through the
C ABI of libgcr_b200.so, against the CPU oracle on identical replayed minimal samples.

Bars (BASELINE.json north_star): inlier masks and counts bit-exact, MSAC scores to 1e-5 relative,
model / vanishing line (h7, h8, 1) to 1e-6 relative; flags and the selected hypothesis identical.
"""
import hashlib
import math
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SCORE_RTOL = 1e-5
MODEL_RTOL = 1e-6


@pytest.fixture(scope="module")
def eng():
    import gcr_capi
    e = gcr_capi.Engine(0)
    yield e
    e.close()


def _compare(dev, ora, ns, no, check_masks=True, label=""):
    f_d, f_o = dev["flags"] & 7, ora["flags"] & 7
    assert np.array_equal(f_d, f_o), f"{label}: flags differ at {np.nonzero(f_d != f_o)[0][:10]}"
    solved = (f_o & 2) > 0
    md, mo = dev["models"][solved], ora["models"][solved]
    assert np.allclose(md, mo, rtol=MODEL_RTOL, atol=1e-300), f"{label}: models differ, max rel {np.max(np.abs(md - mo) / np.abs(mo))}"
    assert np.array_equal(dev["n_inl"], ora["n_inl"]), (
        f"{label}: inlier counts differ at {np.nonzero((dev['n_inl'] != ora['n_inl']).any(1))[0][:10]}")
    assert np.allclose(dev["score"], ora["score"], rtol=SCORE_RTOL, atol=1e-9, equal_nan=True), (
        f"{label}: scores differ, max rel {np.max(np.abs(dev['score'] - ora['score']) / np.maximum(ora['score'], 1e-9))}")
    if check_masks:
        assert np.array_equal(dev["masks"], ora["masks"]), (
            f"{label}: masks differ in hypotheses {np.nonzero((dev['masks'] != ora['masks']).any(1))[0][:10]}")


def _run_replay(eng, scale, orient, samples, thr, masks=True, extra_flags=0, host_libm=False):
    import gcr_capi
    eng.set_features(scale, orient, host_libm=host_libm)
    flags = (gcr_capi.RUN_EMIT_MASKS if masks else gcr_capi.RUN_KEEP_PER_HYP) | extra_flags
    res = eng.run(len(samples), thr, samples=samples, flags=flags)
    dev = eng.fetch(masks=masks)
    best = eng.best()
    # the winner-only mode (no per-hypothesis outputs: what bench.py and the estimator run) selects the same record
    res0 = eng.run(len(samples), thr, samples=samples, flags=extra_flags)
    assert (res0.best_index, res0.best_score, tuple(res0.best_n_inl)) == (res.best_index, res.best_score, tuple(res.best_n_inl))
    b0 = eng.best()
    assert np.array_equal(b0["scale_words"], best["scale_words"]) and np.array_equal(b0["orient_words"], best["orient_words"])
    assert (res0.n_solved, res0.n_sample_valid) == (res.n_solved, res.n_sample_valid)
    with pytest.raises(Exception):
        eng.fetch()  # per-hypothesis scores were not kept by that run
    return res, dev


def test_device_helper_math(eng, port):
    """The 16 known-answer cases of tests/unit_tests.cpp that concern angles, evaluated by the device functions."""
    pi = math.pi
    cases = [(-pi / 2, 1.5 * pi), (0.7861, 0.7861), (0.0, 0.0), (2.7 * pi, 0.7 * pi), (-pi, pi)]
    out = eng.debug_eval(0, [(a, 0, 0) for a, _ in cases])
    assert np.allclose(out, [e for _, e in cases], atol=1e-9)
    cases = [(0.0, 0.0, 0.0), (-pi / 2, 0.0, pi / 2), (-3.48 * pi, 7.41 * pi, 0.89 * pi), (0.25 * pi, 1.25 * pi, pi)]
    out = eng.debug_eval(1, [(a, b, 0) for a, b, _ in cases])
    assert np.allclose(out, [e for _, _, e in cases], atol=1e-9)
    cases = [(0.0, 0.0, 0.0), (-pi / 2, 0.0, pi / 2), (0.52 * pi, 1.49 * pi, 0.03 * pi), (1.49 * pi, 0.52 * pi, 0.03 * pi),
             (-3.48 * pi, 5.60 * pi, 0.08 * pi), (0.25 * pi, 1.25 * pi, 0.0)]
    out = eng.debug_eval(2, [(a, b, 0) for a, b, _ in cases])
    assert np.allclose(out, [e for _, _, e in cases], atol=1e-9)
    # random: the angle helpers use only exact operations (fmod, +, -, fabs, fmin): identical doubles
    rng = np.random.default_rng(0)
    ab = rng.uniform(-40, 40, size=(2000, 3))
    for which, fn in ((0, lambda r: port.clip_angle(r[0])), (1, lambda r: port.min_angle_diff(r[0], r[1])),
                      (2, lambda r: port.lines_angles_diff(r[0], r[1]))):
        assert np.array_equal(eng.debug_eval(which, ab), np.array([fn(r) for r in ab]))


def test_device_residuals_vs_golden_reference(eng, golden_dir):
    """FP64 device residuals vs the reference's own residual tables (differences = libm last ulps)."""
    g = np.load(os.path.join(golden_dir, "residuals_c1.npz"))
    c1 = np.load(os.path.join(golden_dir, "hyp_stage_c1.npz"))
    for m, rs, ro in zip(g["models"], g["rs"], g["ro"]):
        ds = eng.debug_eval(3, c1["scale"], m)
        do = eng.debug_eval(4, c1["orient"], m)
        big = rs > 1e300  # DBL_MAX sentinel
        assert np.array_equal(ds > 1e300, big)
        assert np.allclose(ds[~big], rs[~big], rtol=1e-12, atol=1e-14)
        assert np.allclose(do, ro, rtol=1e-12, atol=1e-14)


@pytest.mark.parametrize("name", ["hyp_stage_small.npz", "hyp_stage_edge.npz"])
def test_replay_matches_reference_golden(eng, golden_dir, name):
    g = np.load(os.path.join(golden_dir, name))
    res, dev = _run_replay(eng, g["scale"], g["orient"], g["samples"], g["thr"])
    ora = {k: g[k] for k in ("flags", "models", "score", "n_inl", "masks")}
    # Knife edge by construction (edge fixture only): two orientation samples at the SAME position put the
    # vanishing point exactly on a vertex of the sample hull, so isValidSample's inside/outside answer is the
    # last ulp of sin/cos (glibc vs CUDA libm).  Those hypothesis slots are compared on nothing.
    o = g["orient"][g["samples"][:, 2:]]
    tie = (o[:, 0, 0] == o[:, 1, 0]) & (o[:, 0, 1] == o[:, 1, 1])
    assert tie.sum() <= 8
    keep = ~tie
    dev = {k: v[keep] for k, v in dev.items() if v is not None}
    ora = {k: v[keep] for k, v in ora.items()}
    _compare(dev, ora, len(g["scale"]), len(g["orient"]), label=name)
    if not tie[int(g["best"])]:
        assert res.best_index == int(g["best"])


def test_edge_golden_bit_exact_replay_mode_no_skipped_slots(eng, golden_dir, port):
    """GCR_FEATURES_HOST_LIBM (include/gcr_b200.h): the solvers' per-feature columns sin / cos (math_utils.hpp:104-109) and
    pow(s, 1/3) (two_sift.hpp:288-299) come from the host's C library, so isValidSample (two_sift.hpp:141-215) runs on the reference's
    own inputs: EVERY slot of the edge fixture is compared, including the ones whose two orientation samples share a position
    (vanishing point on a vertex of the sample hull), and so is the selected hypothesis."""
    g = np.load(os.path.join(golden_dir, "hyp_stage_edge.npz"))
    res, dev = _run_replay(eng, g["scale"], g["orient"], g["samples"], g["thr"], host_libm=True)
    ora = {k: g[k] for k in ("flags", "models", "score", "n_inl", "masks")}
    o = g["orient"][g["samples"][:, 2:]]
    tie = (o[:, 0, 0] == o[:, 1, 0]) & (o[:, 0, 1] == o[:, 1, 1])
    assert tie.sum() >= 1  # the fixture does hold such slots
    _compare(dev, ora, len(g["scale"]), len(g["orient"]), label="edge, host-libm columns")
    assert res.best_index == int(g["best"])
    # exactly-zero thresholds under the same mode: flags of every slot, and the models they yield
    import gcr_scenes as gs
    sc = gs.make_scene(100, 50, 0.5, 1)
    smp = gs.draw_samples(64, 100, 50, 2)
    smp[5, 3] = smp[5, 2] ^ 1  # plus a few slots with the knife edge planted: the second orientation sample moved onto the first one's position
    orient = sc.orient.copy()
    orient[smp[5, 3], :2] = orient[smp[5, 2], :2]
    thr = np.array([0.0, 0.0])
    ora = port.replay(sc.scale, orient, smp, thr)
    _, dev = _run_replay(eng, sc.scale, orient, smp, thr, host_libm=True)
    assert np.array_equal(dev["flags"] & 7, ora["flags"])
    ok = (ora["flags"] & 2) > 0
    assert np.allclose(dev["models"][ok], ora["models"][ok], rtol=MODEL_RTOL, atol=1e-12)


def test_threshold_sweep_matches_reference_golden(eng, golden_dir):
    g = np.load(os.path.join(golden_dir, "hyp_stage_sweep.npz"))
    for p in (1, 2, 4, 8, 20):
        ora = {k: g[f"p{p}_{k}"] for k in ("flags", "models", "score", "n_inl", "masks")}
        _, dev = _run_replay(eng, g["scale"], g["orient"], g[f"p{p}_samples"], g[f"p{p}_thr"])
        _compare(dev, ora, len(g["scale"]), len(g["orient"]), label=f"sweep p={p}")


def test_config_c1_matches_reference_golden(eng, golden_dir):
    """BASELINE config C1: 2000 + 500 features, 5000 replayed samples; mask digest of the reference."""
    g = np.load(os.path.join(golden_dir, "hyp_stage_c1.npz"))
    res, dev = _run_replay(eng, g["scale"], g["orient"], g["samples"], g["thr"])
    _compare(dev, g, 2000, 500, check_masks=False, label="C1")
    assert hashlib.sha256(np.ascontiguousarray(dev["masks"]).tobytes()).hexdigest() == str(g["masks_sha256"])
    assert res.best_index == int(g["best"])
    b = eng.best()
    ws = (2000 + 31) // 32
    assert np.array_equal(np.concatenate([b["scale_words"], b["orient_words"]]), g["best_masks"])
    assert np.array_equal(b["n_inl"], g["n_inl"][int(g["best"])])
    assert np.allclose(b["model"], g["models"][int(g["best"])], rtol=MODEL_RTOL)
    assert abs(b["score"] - g["score"][int(g["best"])]) <= 1e-9 * g["score"][int(g["best"])]
    # vanishing line (h7, h8, 1) of the winner vs the generating plane
    assert np.allclose(b["model"][:2], g["truth"][:2], atol=5e-5)
    assert res.n_solved == int(((g["flags"] & 2) > 0).sum())
    assert ws == len(b["scale_words"])


def test_config_c2_replay_vs_oracle(eng, port):
    """BASELINE config C2: the C1 scene, 100k replayed hypotheses, all masks bit-exact vs the oracle."""
    import gcr_scenes as gs
    sc = gs.config_scene("C2")
    smp = gs.draw_samples(100_000, 2000, 500, gs.SAMPLE_SEED)
    ora = port.replay(sc.scale, sc.orient, smp, sc.thr, nthreads=0)
    res, dev = _run_replay(eng, sc.scale, sc.orient, smp, sc.thr)
    _compare(dev, ora, 2000, 500, label="C2")
    assert res.best_index == port.select_best(ora["score"], ora["flags"])
    # the fast path did the work: only a small share of pairs/hypotheses took the FP64 routes
    assert res.n_exact_hyp < 0.06 * res.n_solved
    assert res.n_rechecked < 1e-3 * res.n_solved * 2500


def test_all_exact_mode_equals_fast_mode(eng, port):
    import gcr_capi
    import gcr_scenes as gs
    sc = gs.make_scene(700, 333, 0.3, 21)
    smp = gs.draw_samples(3000, 700, 333, 22)
    _, fast = _run_replay(eng, sc.scale, sc.orient, smp, sc.thr)
    _, exact = _run_replay(eng, sc.scale, sc.orient, smp, sc.thr, extra_flags=gcr_capi.RUN_ALL_EXACT)
    assert np.array_equal(fast["masks"], exact["masks"])
    assert np.array_equal(fast["n_inl"], exact["n_inl"])
    assert np.allclose(fast["score"], exact["score"], rtol=SCORE_RTOL)
    ora = port.replay(sc.scale, sc.orient, smp, sc.thr)
    _compare(exact, ora, 700, 333, label="all-exact")
    assert np.allclose(exact["score"], ora["score"], rtol=1e-11)


def test_philox_sampler_matches_oracle_and_sharding_invariance(eng, port):
    import gcr_capi
    import gcr_scenes as gs
    sc = gs.make_scene(1500, 400, 0.3, 5)
    eng.set_features(sc.scale, sc.orient)
    H, seed = 20_000, 0xC0FFEE1234
    res = eng.run(H, sc.thr, seed=seed, flags=gcr_capi.RUN_EMIT_MASKS)
    dev = eng.fetch(masks=True, samples=True)
    smp = port.sample_philox(seed, 0, H, 1500, 400)
    assert np.array_equal(dev["samples"], smp)
    ora = port.replay(sc.scale, sc.orient, smp, sc.thr, nthreads=0)
    _compare(dev, ora, 1500, 400, label="philox")
    assert res.best_index == port.select_best(ora["score"], ora["flags"])
    whole = eng.best()
    # two shards with global counters reproduce the same hypotheses; the better shard winner is the global one
    recs = []
    for first in (0, H // 2):
        r = eng.run(H // 2, sc.thr, seed=seed, first_hyp=first, flags=gcr_capi.RUN_KEEP_PER_HYP)
        d = eng.fetch()
        assert np.array_equal(d["flags"], dev["flags"][first:first + H // 2])
        assert np.array_equal(d["n_inl"], dev["n_inl"][first:first + H // 2])
        recs.append((r.best_score, -r.best_index, r.best_index))
    assert max(recs)[2] == whole["index"]


def test_large_scene_slice_and_properties(eng, port):
    """C3-sized feature set (50k + 20k, 10 % inliers): oracle parity on a slice, then size-independent
    properties on a bigger batch: winner masks == single-model scoring, popcount == counts, idempotence."""
    import gcr_capi
    import gcr_scenes as gs
    sc = gs.config_scene("C3")
    smp = gs.draw_samples(3000, 50_000, 20_000, 77)
    ora = port.replay(sc.scale, sc.orient, smp, sc.thr, nthreads=0)
    res, dev = _run_replay(eng, sc.scale, sc.orient, smp, sc.thr)
    _compare(dev, ora, 50_000, 20_000, label="C3 slice")
    assert res.best_index == port.select_best(ora["score"], ora["flags"])
    # bigger batch, Philox, no per-hypothesis masks
    r1 = eng.run(60_000, sc.thr, seed=9)
    b1 = eng.best()
    r2 = eng.run(60_000, sc.thr, seed=9)
    b2 = eng.best()
    assert r1.best_index == r2.best_index and b1["score"] == b2["score"]
    assert np.array_equal(b1["scale_words"], b2["scale_words"])
    s, n, sm, om = eng.score_model(b1["model"], sc.thr)
    assert s == b1["score"] and np.array_equal(n, b1["n_inl"])
    assert np.array_equal(sm, b1["scale_mask"]) and np.array_equal(om, b1["orient_mask"])
    assert (sm.sum(), om.sum()) == tuple(b1["n_inl"])
    # the winner explains the planted plane
    assert np.allclose(b1["model"][:2], sc.model[:2], atol=5e-5)
    assert (sm & sc.scale_is_inlier).sum() > 0.9 * sc.scale_is_inlier.sum()
    # and matches the oracle's scoring of the same model bit-for-bit on masks
    so, no_, smo, omo = port.score_model(sc.scale, sc.orient, b1["model"], sc.thr)
    assert np.array_equal(sm, smo) and np.array_equal(om, omo) and abs(s - so) <= 1e-10 * so


def test_minimal_and_ragged_inputs(eng, port):
    import gcr_scenes as gs
    rng = np.random.default_rng(3)
    for ns, no in ((2, 2), (3, 5), (33, 31), (513, 511), (1025, 64)):
        sc = gs.make_scene(ns, no, 0.5, 100 + ns)
        smp = gs.draw_samples(257, ns, no, ns)
        ora = port.replay(sc.scale, sc.orient, smp, sc.thr)
        _, dev = _run_replay(eng, sc.scale, sc.orient, smp, sc.thr)
        _compare(dev, ora, ns, no, label=f"ragged {ns}/{no}")
    # vanishing thresholds: only the minimal sample itself (residual ~1e-16) can be an inlier; every hypothesis
    # leaves the FP32 path's proven range and is scored by the FP64 kernel
    sc = gs.make_scene(100, 50, 0.5, 1)
    smp = gs.draw_samples(64, 100, 50, 2)
    thr = np.array([1e-9, 1e-9])
    ora = port.replay(sc.scale, sc.orient, smp, thr)
    _, dev = _run_replay(eng, sc.scale, sc.orient, smp, thr)
    _compare(dev, ora, 100, 50, label="tiny thr")
    # exactly zero thresholds: r^2 <= 0 holds only for residuals that are exactly 0.0, i.e. the last ulp of
    # pow/log/atan2 (glibc vs CUDA) decides -> only flags and models are comparable; 0/0 scores are NaN on both
    thr = np.array([0.0, 0.0])
    ora = port.replay(sc.scale, sc.orient, smp, thr)
    _, dev = _run_replay(eng, sc.scale, sc.orient, smp, thr)
    assert np.array_equal(dev["flags"] & 7, ora["flags"])
    # very loose thresholds route the orientation class to the exact path
    thr = np.array([2.0, 0.6])
    ora = port.replay(sc.scale, sc.orient, smp, thr)
    _, dev = _run_replay(eng, sc.scale, sc.orient, smp, thr)
    _compare(dev, ora, 100, 50, label="loose thr")


def test_asynchronous_runs_report_device_side_failures(port):
    """gcr_run with out == NULL (asynchronous) cannot report what its kernels find; gcr_best / gcr_fetch must (GCRANSAC.h:222-228: the
    reference throws rather than continue): a guard-band worklist overflow (forced with a one-slot worklist) and replayed sample
    indices out of range."""
    import gcr_capi
    import gcr_scenes as gs
    os.environ["GCR_WORKLIST_CAP"] = "1"
    try:
        e = gcr_capi.Engine(0)
    finally:
        del os.environ["GCR_WORKLIST_CAP"]
    sc = gs.make_scene(2000, 500, 0.3, 3)
    smp = gs.draw_samples(20_000, 2000, 500, 4)
    e.set_features(sc.scale, sc.orient)
    e.run(len(smp), sc.thr, samples=smp, flags=gcr_capi.RUN_EMIT_MASKS, sync=False)
    with pytest.raises(gcr_capi.GcrError) as ei:
        e.best()
    assert ei.value.code == gcr_capi.ERR_STATE and "overflow" in str(ei.value)
    e.run(len(smp), sc.thr, samples=smp, flags=gcr_capi.RUN_EMIT_MASKS, sync=False)
    with pytest.raises(gcr_capi.GcrError):
        e.fetch(masks=True)
    with pytest.raises(gcr_capi.GcrError):  # the synchronous form reports it itself
        e.run(len(smp), sc.thr, samples=smp, flags=gcr_capi.RUN_EMIT_MASKS)
    # a winner-only run has no guard bands, hence no worklist to overflow
    res = e.run(len(smp), sc.thr, samples=smp)
    ora = port.replay(sc.scale, sc.orient, smp, sc.thr, want_masks=False, nthreads=0)
    assert res.best_index == port.select_best(ora["score"], ora["flags"])
    e.close()
    e = gcr_capi.Engine(0)
    e.set_features(sc.scale, sc.orient)
    bad = smp[:1000].copy()
    bad[17, 2] = 500
    e.run(len(bad), sc.thr, samples=bad, sync=False)
    with pytest.raises(gcr_capi.GcrError) as ei:
        e.best()
    assert ei.value.code == gcr_capi.ERR_INVALID
    e.run(1000, sc.thr, samples=smp[:1000], sync=False)  # and the next run is clean again
    assert e.best()["index"] >= 0
    e.close()


def test_entry_points_leave_the_callers_current_device_alone(port):
    """Every ABI entry point works on the engine's device and restores the caller's (a torch process reads the thread's current
    device with cudaGetDevice: an engine on device 1 must not move later allocations there).  Needs two devices."""
    import torch
    import gcr_capi
    import gcr_scenes as gs
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    torch.cuda.set_device(0)
    sc = gs.make_scene(500, 200, 0.4, 8)
    smp = gs.draw_samples(2000, 500, 200, 9)
    e = gcr_capi.Engine(1)
    assert torch.cuda.current_device() == 0
    e.set_features(sc.scale, sc.orient)
    res = e.run(len(smp), sc.thr, samples=smp)
    b = e.best()
    e.score_model(b["model"], sc.thr)
    assert torch.cuda.current_device() == 0
    assert torch.zeros(1, device="cuda").device.index == 0
    ora = port.replay(sc.scale, sc.orient, smp, sc.thr, want_masks=False, nthreads=0)
    assert res.best_index == port.select_best(ora["score"], ora["flags"])
    e.close()
    assert torch.cuda.current_device() == 0


def test_error_conventions(eng):
    import gcr_capi
    with pytest.raises(gcr_capi.GcrError):
        eng.set_features(np.zeros((1, 3)), np.zeros((5, 3)))  # < 2 rows (bindings.cpp:206-224)
    import gcr_scenes as gs
    sc = gs.make_scene(50, 20, 0.5, 1)
    eng.set_features(sc.scale, sc.orient)
    bad = np.array([[0, 1, 0, 99]], np.int32)
    with pytest.raises(gcr_capi.GcrError):
        eng.run(1, sc.thr, samples=bad, flags=gcr_capi.RUN_KEEP_PER_HYP)
    with pytest.raises(gcr_capi.GcrError):
        eng.run(10, [float("nan"), 0.1])


def test_batched_model_scoring_vs_oracle(eng, port):
    """gcr_score_models: the rescoring step of the local optimisation (GCRANSAC.h:1004-1046) for a batch of caller
    models: scores, counts and masks of every model against the oracle; the truncated-threshold relabelling
    (GCRANSAC.h:931-937) against the oracle's per-feature residuals."""
    import gcr_scenes as gs
    sc = gs.make_scene(1500, 333, 0.35, 909)
    smp = gs.draw_samples(400, 1500, 333, 17)
    ora = port.replay(sc.scale, sc.orient, smp, sc.thr)
    models = np.vstack([sc.model, ora["models"][(ora["flags"] & 2) > 0][:97]])
    eng.set_features(sc.scale, sc.orient)
    score, n_inl, words = eng.score_models(models, sc.thr, masks=True)
    ws = (1500 + 31) // 32
    for k, m in enumerate(models):
        val, cnt, smask, omask = port.score_model(sc.scale, sc.orient, m, sc.thr)
        assert np.array_equal(n_inl[k], cnt), k
        assert score[k] == pytest.approx(val, rel=1e-9, abs=1e-12), k
        from gcr_capi import unpack_mask
        assert np.array_equal(unpack_mask(words[k, :ws], 1500), smask) and np.array_equal(unpack_mask(words[k, ws:], 333), omask)
    # single-model entry point is the n = 1 case
    v1, c1, sm1, om1 = eng.score_model(models[3], sc.thr)
    assert v1 == score[3] and np.array_equal(c1, n_inl[3])
    # relabelling with (1.5 t)^2
    _, _, lab = eng.score_models(models[:5], sc.thr, truncated_sq=True, masks=True)
    tsq = (1.5 * sc.thr) ** 2
    for k in range(5):
        rs, ro = port.residuals(sc.scale, sc.orient, models[k])
        from gcr_capi import unpack_mask
        assert np.array_equal(unpack_mask(lab[k, :ws], 1500), rs * rs <= tsq[0])
        assert np.array_equal(unpack_mask(lab[k, ws:], 333), ro * ro <= tsq[1])
    with pytest.raises(Exception):
        eng.score_models(np.zeros((1025, 4)), sc.thr)


def test_randomised_scenes_vs_oracle(eng, port):
    """40 seeded scenes of ragged sizes, thresholds from tight to loose, different inlier ratios, with awkward rows
    injected (duplicates, zero / negative / tiny / huge scales, repeated and out-of-range angles, far-away points):
    every per-hypothesis output against the CPU oracle on the same replayed samples."""
    import gcr_scenes as gs
    rng = np.random.Generator(np.random.PCG64(20261104))
    n_solved = 0
    for case in range(40):
        ns, no = int(rng.integers(2, 700)), int(rng.integers(2, 400))
        ratio = float(rng.choice([0.05, 0.2, 0.5, 0.9]))
        lvl = float(rng.choice([0.25, 1.0, 3.0, 8.0, 25.0]))
        sc = gs.make_scene(ns, no, ratio, 5000 + case, thr_scale=0.05 * lvl, thr_orient=math.radians(lvl))
        scale, orient = sc.scale.copy(), sc.orient.copy()
        if case % 3 == 0 and ns > 12 and no > 12:
            scale[3] = scale[2]
            scale[5, 2] = 0.0
            scale[6, 2] = -1.5
            scale[7, :2] = [3e5, -2e5]
            scale[8, 2] = 1e-11
            scale[9, 2] = 1e11
            orient[4] = orient[1]
            orient[6, 2] = orient[5, 2]
            # (two orientation features at the SAME position put the vanishing point on a hull vertex, where the sign of an
            #  exactly-zero cross product is decided by the last ulp of sin/cos: see the edge fixture, which skips those slots)
            orient[8, 2] = orient[5, 2] + math.pi
            orient[9, 2] = 9.5 * math.pi
            orient[10, 2] = -2.75 * math.pi
        H = int(rng.integers(1, 1500))
        smp = gs.draw_samples(H, ns, no, 900 + case)
        if case % 3 == 0 and ns > 12 and no > 12 and H > 64:
            smp[:64, :2] = rng.integers(0, 11, (64, 2))
            smp[:64, 2:] = rng.integers(0, 11, (64, 2))
            same = smp[:64, 0] == smp[:64, 1]
            smp[:64, 1][same] = (smp[:64, 1][same] + 1) % 11
            same = smp[:64, 2] == smp[:64, 3]
            smp[:64, 3][same] = (smp[:64, 3][same] + 1) % 11
        ora = port.replay(scale, orient, smp, sc.thr, nthreads=0)
        res, dev = _run_replay(eng, scale, orient, smp, sc.thr)
        _compare(dev, ora, ns, no, label=f"random case {case} ({ns}+{no} features, level {lvl}, {H} hypotheses)")
        want = port.select_best(ora["score"], ora["flags"])
        tie = np.isclose(ora["score"], ora["score"][want], rtol=1e-12) & ((ora["flags"] & 4) > 0) if want >= 0 else None
        if want < 0 or tie.sum() == 1:
            assert res.best_index == want, case
        n_solved += int(((ora["flags"] & 2) > 0).sum())
    assert n_solved > 2000


@pytest.mark.parametrize("H, min_tied", [(6000, 2000), (1400, 600), (120, 50)])
def test_thousands_of_tied_hypotheses_select_the_earliest(eng, port, H, min_tied):
    """Many exactly tied near-best hypotheses: a scene whose minimal samples all come from a handful of exactly consistent
    features; the reference keeps the first (S11).  6000 slots: more ties than the re-scoring pass holds at once (1024
    candidate slots), i.e. several collect | re-score | select rounds of k_tail; 1400 and 120: hundreds / tens of candidates
    in one round."""
    import gcr_scenes as gs
    sc = gs.make_scene(40, 30, 1.0, 77, thr_scale=0.2, thr_orient=math.radians(3.0))
    # noise-free inliers: regenerate the scale / orientation values exactly from the ground-truth model
    scale, orient = sc.scale.copy(), sc.orient.copy()
    h7, h8, alpha, phi = sc.model
    w = 1.0 - h7 * scale[:, 0] - h8 * scale[:, 1]
    scale[:, 2] = (w / alpha) ** 3
    smp = gs.draw_samples(6000, 40, 30, 3)[:H].copy()
    smp[:, :2] = smp[0, :2]  # every hypothesis uses the same two scale features ...
    smp[:, 2:] = smp[np.arange(H) % 2, 2:]  # ... and one of 2 orientation pairs: 2 distinct models, H / 2 copies each
    ora = port.replay(scale, orient, smp, sc.thr, nthreads=0)
    want = port.select_best(ora["score"], ora["flags"])
    assert want >= 0
    n_tied = int(np.sum(ora["score"] == ora["score"][want]))
    assert n_tied > min_tied
    res, dev = _run_replay(eng, scale, orient, smp, sc.thr)
    _compare(dev, ora, 40, 30, label="ties")
    assert res.best_index == want, (res.best_index, want, n_tied, res.n_candidates)
    assert res.best_score == pytest.approx(ora["score"][want], rel=1e-12)
    assert res.n_candidates >= n_tied


def test_pinned_host_inputs_are_read_in_place_with_identical_results(eng, port):
    """Page-locked host inputs take the zero-copy route (k_prep_features / k_sample_solve read them over PCIe, the tail
    writes the result block into pinned staging): every output equals the pageable-memory route's."""
    import ctypes as C
    import torch
    import gcr_capi
    import gcr_scenes as gs
    sc = gs.make_scene(777, 333, 0.3, 12)
    smp = gs.draw_samples(5000, 777, 333, 13)
    eng.set_features(sc.scale, sc.orient)  # pageable numpy arrays: staged copies
    ref_res = eng.run(len(smp), sc.thr, samples=smp, flags=gcr_capi.RUN_EMIT_MASKS)
    ref = eng.fetch(masks=True, samples=True)
    ref_best = eng.best()
    e = gcr_capi.Engine(0)
    p_scale, p_orient = torch.from_numpy(sc.scale).pin_memory(), torch.from_numpy(sc.orient).pin_memory()
    p_smp = torch.from_numpy(smp).pin_memory()
    assert e.lib.gcr_set_features(e.h, C.c_void_p(p_scale.data_ptr()), 777, C.c_void_p(p_orient.data_ptr()), 333, 0) == 0
    e.ns, e.no, e.fetch_ns, e.fetch_no = 777, 333, 777, 333
    thr_c = (C.c_double * 2)(*sc.thr)
    out = gcr_capi.GcrResult()
    assert e.lib.gcr_run(e.h, len(smp), 0, 0, C.c_void_p(p_smp.data_ptr()), thr_c, gcr_capi.RUN_EMIT_MASKS, None, C.byref(out)) == 0
    e.n_hyp = len(smp)
    got = e.fetch(masks=True, samples=True)
    for k in ("flags", "models", "n_inl", "masks", "samples"):
        assert np.array_equal(got[k], ref[k]), k
    # FP32-path sums of a hypothesis split over several work items are combined with FP64 atomics (last-ulp run-to-run
    # freedom), and in the resident kernel the FP32 partial sums are grouped by the piece cuts, which follow the position of
    # the hypothesis in the compacted list (atomic order): reporting-mode scores repeat to FP32 summation accuracy (a few 1e-8
    # observed, bar 2e-6: still 5x inside the 1e-5 parity bar), the winner (re-scored in FP64 reference order) bit for bit
    assert np.allclose(got["score"], ref["score"], rtol=2e-6, atol=0.0)
    assert (out.best_index, out.best_score, tuple(out.best_n_inl)) == (ref_res.best_index, ref_res.best_score, tuple(ref_res.best_n_inl))
    b = e.best()
    assert np.array_equal(b["scale_words"], ref_best["scale_words"]) and np.array_equal(b["orient_words"], ref_best["orient_words"])
    s, o = e.get_features()
    assert np.array_equal(s, sc.scale) and np.array_equal(o, sc.orient)  # the device copy written by the preparation kernel
    e.close()


def test_uint16_replay_samples_equal_int32_replay_samples(eng):
    """GCR_RUN_SAMPLES_U16: the replayed minimal samples as uint16[n_hyp,4] (half the bytes over PCIe) -- pageable, page-locked (read in
    place) and device arrays -- give what the int32 array gives: flags, models, counts, masks bit for bit, the samples read back widened;
    the scale-only problems' unused fourth index travels as 0xFFFF; a class of more than 65 535 features is refused."""
    import ctypes as C
    import torch
    import gcr_capi
    import gcr_scenes as gs
    sc = gs.make_scene(1500, 400, 0.4, 21)
    smp = gs.draw_samples(20000, 1500, 400, 22)
    smp[7] = [3, 3, 5, 9]  # a degenerate sample and one out of range stay what they are
    eng.set_features(sc.scale, sc.orient)
    ref_res = eng.run(len(smp), sc.thr, samples=smp, flags=gcr_capi.RUN_EMIT_MASKS)
    ref = eng.fetch(masks=True, samples=True)
    narrow = smp.astype(np.uint16)
    thr_c = (C.c_double * 2)(*sc.thr)
    pinned, on_device = torch.from_numpy(narrow).pin_memory(), torch.from_numpy(narrow).cuda()
    for how in ("pageable", "pinned", "device"):
        if how == "pageable":
            res = eng.run(len(smp), sc.thr, samples=narrow, flags=gcr_capi.RUN_EMIT_MASKS)
        else:
            res = gcr_capi.GcrResult()
            ptr = pinned.data_ptr() if how == "pinned" else on_device.data_ptr()
            fl = gcr_capi.RUN_EMIT_MASKS | gcr_capi.RUN_SAMPLES_U16 | (gcr_capi.RUN_SAMPLES_ON_DEVICE if how == "device" else 0)
            assert eng.lib.gcr_run(eng.h, len(smp), 0, 0, C.c_void_p(ptr), thr_c, fl, None, C.byref(res)) == 0, how
        got = eng.fetch(masks=True, samples=True)
        for k in ("flags", "models", "n_inl", "masks", "samples"):
            assert np.array_equal(got[k], ref[k]), (how, k)
        assert (res.best_index, res.best_score) == (ref_res.best_index, ref_res.best_score), how
    # winner-only run (no per-hypothesis outputs)
    res = eng.run(len(smp), sc.thr, samples=narrow)
    assert (res.best_index, res.best_score) == (ref_res.best_index, ref_res.best_score)
    bad = narrow.copy()
    bad[11, 2] = 400  # one past the orientation class
    with pytest.raises(gcr_capi.GcrError):
        eng.run(len(bad), sc.thr, samples=bad)
    # scale-only: (s0, s1, s2, -1) <-> (s0, s1, s2, 0xFFFF)
    e = gcr_capi.Engine(0)
    e.set_problem(gcr_capi.PROBLEM_SCALE_ONLY)
    e.set_features(sc.scale)
    s3 = gs.draw_samples_scale3(5000, 1500, 23)
    r32 = e.run(len(s3), [sc.thr[0], 0.0], samples=s3, flags=gcr_capi.RUN_KEEP_PER_HYP)
    f32 = e.fetch(samples=True)
    r16 = e.run(len(s3), [sc.thr[0], 0.0], samples=s3.astype(np.uint16), flags=gcr_capi.RUN_KEEP_PER_HYP)
    f16 = e.fetch(samples=True)
    assert (r16.best_index, r16.best_score) == (r32.best_index, r32.best_score) and r32.best_index >= 0
    for k in ("flags", "models", "n_inl", "samples"):
        assert np.array_equal(f16[k], f32[k]), k
    e.close()
    # more features than 16 bits index
    big = gs.make_scene(66000, 300, 0.2, 24)
    e = gcr_capi.Engine(0)
    e.set_features(big.scale, big.orient)
    with pytest.raises(gcr_capi.GcrError):
        e.run(100, big.thr, samples=np.zeros((100, 4), np.uint16))
    e.close()


def test_sm_share_does_not_change_results(eng):
    """gcr_set_sm_share: an engine sized for 1/k of the SMs (k engines side by side on k streams: image batches) returns what the full-size
    engine returns -- reporting mode: flags, models, counts, masks bit for bit, scores to FP32 summation accuracy (the partial sums are
    grouped by the work cuts, which follow the grid); winner-only mode: the winner, its FP64 score, counts and masks bit for bit -- for the
    resident kernel (small scene) and the ring kernel (larger scene), replayed and Philox samples."""
    import gcr_capi
    import gcr_scenes as gs
    for sc, n_hyp in ((gs.make_scene(2000, 500, 0.3, 31), 30000), (gs.make_scene(9000, 4000, 0.2, 32), 40000)):
        smp = gs.draw_samples(n_hyp, len(sc.scale), len(sc.orient), 33)
        eng.set_sm_share(1)
        eng.set_features(sc.scale, sc.orient)
        eng.run(n_hyp, sc.thr, samples=smp, flags=gcr_capi.RUN_EMIT_MASKS)
        ref = eng.fetch(masks=True)
        ref_w = eng.run(n_hyp, sc.thr, seed=77)
        ref_b = eng.best()
        for share in (2, 8, 37):
            e = gcr_capi.Engine(0)
            e.set_sm_share(share)
            e.set_features(sc.scale, sc.orient)
            e.run(n_hyp, sc.thr, samples=smp, flags=gcr_capi.RUN_EMIT_MASKS)
            got = e.fetch(masks=True)
            for k in ("flags", "models", "n_inl", "masks"):
                assert np.array_equal(got[k], ref[k]), (share, k)
            assert np.allclose(got["score"], ref["score"], rtol=2e-6, atol=0.0), share
            w = e.run(n_hyp, sc.thr, seed=77)
            b = e.best()
            assert (w.best_index, w.best_score, tuple(w.best_n_inl)) == (ref_w.best_index, ref_w.best_score, tuple(ref_w.best_n_inl)), share
            assert np.array_equal(b["scale_words"], ref_b["scale_words"]) and np.array_equal(b["orient_words"], ref_b["orient_words"]), share
            e.close()
    with pytest.raises(gcr_capi.GcrError):
        eng.set_sm_share(0)
    with pytest.raises(gcr_capi.GcrError):
        eng.set_sm_share(100000)
    eng.set_sm_share(1)


def test_ring_and_resident_scoring_kernels_agree(port):
    """Small feature sets are scored by k_score_resident (whole working set in shared memory, static cost partition over warps),
    large ones or big batches by k_score_fast (TMA ring).  Same arithmetic: masks, counts, flags and the winner are bit-identical,
    scores equal up to the grouping of the FP32 partial sums; both match the oracle.  Sizes chosen so that precise tiles, ragged
    last blocks (Ns, No not multiples of 32) and tiles split over several warps all occur."""
    import gcr_capi
    import gcr_scenes as gs
    engines = {}
    for name, env in (("ring", {"GCR_RESIDENT": "0"}), ("resident", {"GCR_RESIDENT_MAX_HYP": "100000000"})):
        os.environ.update(env)
        try:
            engines[name] = gcr_capi.Engine(0)
        finally:
            for k in env:
                del os.environ[k]
    try:
        for ns, no, n_hyp, seed in ((1999, 487, 20000, 5), (130, 33, 3000, 6), (4000, 1000, 150000, 7), (2, 2, 64, 8)):
            sc = gs.make_scene(ns, no, 0.3, seed)
            smp = gs.draw_samples(n_hyp, ns, no, seed + 100)
            out = {}
            for name, e in engines.items():
                e.set_features(sc.scale, sc.orient)
                res = e.run(n_hyp, sc.thr, samples=smp, flags=gcr_capi.RUN_EMIT_MASKS)
                out[name] = (res, e.fetch(masks=True))
            (ra, a), (rb, b) = out["ring"], out["resident"]
            for k in ("flags", "models", "n_inl", "masks"):
                assert np.array_equal(a[k], b[k]), (ns, no, k)
            assert np.allclose(a["score"], b["score"], rtol=2e-6, atol=0.0)  # FP32 partial sums grouped by chunk vs by piece
            assert (ra.best_index, ra.best_score, tuple(ra.best_n_inl)) == (rb.best_index, rb.best_score, tuple(rb.best_n_inl))
            if n_hyp <= 20000:
                ora = port.replay(sc.scale, sc.orient, smp, sc.thr, nthreads=0)
                _compare(b, ora, ns, no, label=f"resident {ns}+{no}")
    finally:
        for e in engines.values():
            e.close()


def test_back_to_back_feature_changes_with_chained_launches(port):
    """gcr_set_features + gcr_run back to back with page-locked inputs, alternating between two scenes of the same size on one
    engine: the kernels are chained with programmatic dependent launches (each starts under its predecessor and waits on the
    device for what it needs), so a kernel that read the feature planes before k_prep_features had finished would score the
    PREVIOUS scene's features.  Small batches (few CTAs of k_sample_solve, most SMs free for an early scoring kernel) are
    the exposed case; every run must reproduce its own scene's reference result."""
    import ctypes as C
    import torch
    import gcr_capi
    import gcr_scenes as gs
    ns, no, n_hyp = 1500, 400, 2048
    scenes = [gs.make_scene(ns, no, 0.3, 21), gs.make_scene(ns, no, 0.4, 22)]
    smps = [gs.draw_samples(n_hyp, ns, no, 31), gs.draw_samples(n_hyp, ns, no, 32)]
    want = []
    for sc, smp in zip(scenes, smps):
        ora = port.replay(sc.scale, sc.orient, smp, sc.thr, want_masks=False, nthreads=0)
        want.append(port.select_best(ora["score"], ora["flags"]))
    assert want[0] >= 0 and want[1] >= 0
    pinned = [tuple(torch.from_numpy(a).pin_memory() for a in (sc.scale, sc.orient, smp)) for sc, smp in zip(scenes, smps)]
    e = gcr_capi.Engine(0)
    try:
        out = gcr_capi.GcrResult()
        for it in range(60):
            k = it & 1
            ps, po, pp = pinned[k]
            thr_c = (C.c_double * 2)(*scenes[k].thr)
            assert e.lib.gcr_set_features(e.h, C.c_void_p(ps.data_ptr()), ns, C.c_void_p(po.data_ptr()), no, 0) == 0
            assert e.lib.gcr_run(e.h, n_hyp, 0, 0, C.c_void_p(pp.data_ptr()), thr_c, 0, None, C.byref(out)) == 0
            assert out.best_index == want[k], (it, out.best_index, want)
    finally:
        e.close()
