"""Winner-only runs (no per-hypothesis outputs asked for): the scoring kernels evaluate MSAC as the truncated quadratic
sum_f max(0, 1 - r^2/tsq) in FP32 without inlier decisions, models that isValidModel rejects are not scored, and the
selection re-scores in FP64 reference order every hypothesis whose upper bound reaches the best lower bound
(gcr_score.cuh score_chunk_w, gcr_kernels.cuh *_phase_w).  The selected hypothesis, its score, counts, model and packed
masks must be the reference's (GCRANSAC.h:440-462 on MSAC_scoring_function.hpp:22-130): checked here against the CPU
oracle's replay + selection rule on the same samples, through the C ABI."""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    import gcr_capi
    e = gcr_capi.Engine(0)
    yield e
    e.close()


def _check_winner(eng, port, scale, orient, smp, thr, label="", libm_ties=False, flags=0):
    """winner-only replay run vs oracle: index, score (1e-11: both FP64 reference order), counts, model, masks.
    libm_ties: where the oracle's scores of two hypotheses agree to 1e-13 relative their order is the last ulp of log / pow / atan2
    (glibc vs CUDA libm); either is accepted (thresholds of 1e-9, where only the minimal sample itself scores)."""
    import gcr_capi
    ora = port.replay(scale, orient, smp, thr, nthreads=0)
    want = port.select_best(ora["score"], ora["flags"])
    eng.set_features(scale, orient)
    res = eng.run(len(smp), thr, samples=smp, flags=flags)
    if libm_ties and want >= 0 and res.best_index >= 0 and res.best_index != want:
        got = res.best_index
        assert (ora["flags"][got] & 6) == 6 and abs(ora["score"][got] - ora["score"][want]) <= 1e-13 * abs(ora["score"][want]), label
        want = got
    assert res.best_index == want, (label, res.best_index, want, res.best_score, ora["score"][want] if want >= 0 else None)
    b = eng.best()
    if want < 0:
        assert b["index"] == -1 and b["score"] == 0.0
        return res
    assert abs(b["score"] - ora["score"][want]) <= 1e-11 * max(1.0, abs(ora["score"][want])), label
    assert np.array_equal(b["n_inl"], ora["n_inl"][want]), label
    assert np.allclose(b["model"], ora["models"][want], rtol=1e-6, atol=1e-12), label
    ws = (len(scale) + 31) // 32
    assert np.array_equal(b["scale_words"], ora["masks"][want][:ws]), label
    assert np.array_equal(b["orient_words"], ora["masks"][want][ws:]), label
    assert res.n_solved == int(((ora["flags"] & 2) > 0).sum())
    if flags == 0:
        # the slots of rejected models are reported (flags) but not scored; no inlier decisions, hence no guard-band pairs
        assert res.n_scored == int(((ora["flags"] & 4) > 0).sum())
        assert res.n_rechecked == 0
    else:
        assert res.n_scored == res.n_solved
    return res


def test_winner_only_matches_oracle_on_random_scenes(eng, port):
    import gcr_scenes as gs
    rng = np.random.default_rng(20261)
    for k in range(24):
        ns, no = int(rng.integers(2, 3000)), int(rng.integers(2, 900))
        ratio = float(rng.choice([0.05, 0.1, 0.3, 0.5, 0.8]))
        level = float(rng.choice([0.25, 0.5, 1.0, 2.0, 4.0, 8.0]))
        sc = gs.make_scene(ns, no, ratio, 9000 + k, thr_scale=0.05 * level, thr_orient=math.radians(level))
        smp = gs.draw_samples(int(rng.integers(1, 6000)), ns, no, 500 + k)
        _check_winner(eng, port, sc.scale, sc.orient, smp, sc.thr, label=f"scene {k}: {ns}+{no} ratio {ratio} level {level}")


def test_winner_only_config_c1_and_c2_slice(eng, port):
    import gcr_scenes as gs
    sc = gs.config_scene("C2")
    smp = gs.draw_samples(100_000, 2000, 500, gs.SAMPLE_SEED)
    r = _check_winner(eng, port, sc.scale, sc.orient, smp[:5000], sc.thr, "C1")
    assert r.n_exact_hyp < 0.06 * r.n_scored
    r = _check_winner(eng, port, sc.scale, sc.orient, smp[:40_000], sc.thr, "C2 slice")
    assert r.n_candidates <= 8 and r.n_refined < 2000  # the cascade keeps the FP64 re-scoring to a handful


def test_winner_only_equals_reporting_run_at_c2(eng):
    """100k replayed hypotheses: the winner-only run and the reporting run (per-hypothesis counts, guard bands, worklist) select the
    same hypothesis with the same FP64 score and masks; ring and resident kernels agree."""
    import os
    import gcr_capi
    import gcr_scenes as gs
    sc = gs.config_scene("C2")
    smp = gs.draw_samples(100_000, 2000, 500, gs.SAMPLE_SEED)
    eng.set_features(sc.scale, sc.orient)
    rep = eng.run(len(smp), sc.thr, samples=smp, flags=gcr_capi.RUN_KEEP_PER_HYP)
    brep = eng.best()
    win = eng.run(len(smp), sc.thr, samples=smp)
    bwin = eng.best()
    assert (win.best_index, win.best_score) == (rep.best_index, rep.best_score)
    assert np.array_equal(bwin["scale_words"], brep["scale_words"]) and np.array_equal(bwin["orient_words"], brep["orient_words"])
    assert np.array_equal(bwin["n_inl"], brep["n_inl"])
    assert win.score_kernel == 1
    os.environ["GCR_RESIDENT"] = "0"
    try:
        ring = gcr_capi.Engine(0)
    finally:
        del os.environ["GCR_RESIDENT"]
    ring.set_features(sc.scale, sc.orient)
    rr = ring.run(len(smp), sc.thr, samples=smp)
    assert rr.score_kernel == 0 and (rr.best_index, rr.best_score) == (win.best_index, win.best_score)
    ring.close()


def test_winner_only_ties_and_duplicates(eng, port):
    """Thousands of identical samples: every copy has the same score, the earliest index wins (S11) -- the candidate window must hold all
    of them, over several rounds of 1024 slots."""
    import gcr_scenes as gs
    sc = gs.make_scene(800, 300, 0.4, 77)
    smp = gs.draw_samples(3000, 800, 300, 78)
    ora = port.replay(sc.scale, sc.orient, smp, sc.thr, nthreads=0)
    best = port.select_best(ora["score"], ora["flags"])
    tied = np.tile(smp[best], (6000, 1))
    tied[:100] = smp[:100]
    tied[5000:5100] = smp[100:200]
    _check_winner(eng, port, sc.scale, sc.orient, tied.astype(np.int32), sc.thr, "ties")


def test_winner_only_degenerate_thresholds_and_sizes(eng, port):
    import gcr_scenes as gs
    for ns, no in ((2, 2), (3, 5), (33, 31), (513, 511), (1025, 64)):
        sc = gs.make_scene(ns, no, 0.5, 100 + ns)
        smp = gs.draw_samples(257, ns, no, ns)
        _check_winner(eng, port, sc.scale, sc.orient, smp, sc.thr, f"ragged {ns}/{no}")
    sc = gs.make_scene(400, 150, 0.5, 1)
    smp = gs.draw_samples(512, 400, 150, 2)
    for thr in ((1e-9, 1e-9), (0.0, 0.0), (5.0, 0.2), (0.2, math.radians(40.0)), (1e-4, math.radians(0.01)), (3.0, math.radians(25.0))):
        _check_winner(eng, port, sc.scale, sc.orient, smp, np.array(thr), f"thr {thr}", libm_ties=thr[0] < 1e-3)


def test_winner_only_irregular_features(eng, port):
    """Non-finite / absurd feature rows stay out of the FP32 planes and are evaluated in FP64 for every hypothesis."""
    import gcr_scenes as gs
    sc = gs.make_scene(600, 200, 0.4, 5)
    scale, orient = sc.scale.copy(), sc.orient.copy()
    scale[7, 2] = 0.0
    scale[100, 2] = 1e-40
    scale[300, 0] = 1e300
    orient[11, 2] = float("inf")
    orient[50, 0] = 1e300
    smp = gs.draw_samples(2000, 600, 200, 6)
    bad = np.isin(smp[:, 0], (7, 100, 300)) | np.isin(smp[:, 1], (7, 100, 300)) | np.isin(smp[:, 2], (11, 50)) | np.isin(smp[:, 3], (11, 50))
    _check_winner(eng, port, scale, orient, smp[~bad], sc.thr, "irregular")


def test_winner_only_reversed_fp32_order_is_resolved_in_fp64(eng, port):
    """Two hypotheses whose true scores differ by less than the FP32 path resolves (but by more than the last ulps of libm), at a tight
    (0.25 degree) and a loose (25 degree) threshold: the selection must follow the FP64 reference order (GCRANSAC.h:440-446,
    score.hpp:28-36).  The pair: the best minimal sample of a scene, and the same sample with one orientation feature replaced by a
    copy of it whose angle is moved by 1e-9 ... 1e-6 rad (appended rows), planted in both orders of appearance."""
    import gcr_scenes as gs
    checked = 0
    for level in (0.25, 3.0, 25.0):
        sc = gs.make_scene(1200, 500, 0.5, 31, thr_scale=0.05 * min(max(level, 1.0), 8.0), thr_orient=math.radians(level))
        base = gs.draw_samples(4000, 1200, 500, 32)
        ora = port.replay(sc.scale, sc.orient, base, sc.thr, nthreads=0)
        best = port.select_best(ora["score"], ora["flags"])
        assert best >= 0
        for k, eps in enumerate((1e-9, -1e-8, 1e-7, -1e-6)):
            orient = np.vstack([sc.orient, sc.orient[base[best, 3]] + np.array([0.0, 0.0, eps])])
            twin = base[best].copy()
            twin[3] = len(orient) - 1
            for order in ((twin, base[best]), (base[best], twin)):
                smp = np.concatenate([base[:50], np.array(order), base[50:300]]).astype(np.int32)
                o2 = port.replay(sc.scale, orient, smp[50:52], sc.thr)
                gap = abs(o2["score"][0] - o2["score"][1]) / o2["score"][0]
                if not (1e-12 < gap < 1e-5):
                    continue  # a libm-level tie, or a pair the FP32 values already order
                checked += 1
                import gcr_capi
                # the reporting modes select on proven bounds too (finalize_phase / collect_phase: exact counts, FP32 sums)
                rep = _check_winner(eng, port, sc.scale, orient, smp, sc.thr, f"twins level {level} eps {eps} (reporting)",
                                    flags=gcr_capi.RUN_KEEP_PER_HYP)
                assert rep.n_candidates >= 2 or rep.n_exact_hyp > 0 or gap > 1e-7
                res = _check_winner(eng, port, sc.scale, orient, smp, sc.thr, f"twins level {level} eps {eps}")
                assert res.n_refined >= 2 or res.n_exact_hyp > 0  # (25 degrees: every hypothesis is scored in FP64 from the start) both went through the refinement (exact counts, error ~1e-6 of the value) before the FP64 level
    assert checked >= 8


def test_winner_only_scale_only_problems(eng, port):
    import gcr_capi
    import gcr_scenes as gs
    sc = gs.make_scene(1500, 10, 0.4, 12)
    smp = gs.draw_samples_scale3(5000, 1500, 13)
    for problem, original in ((gcr_capi.PROBLEM_SCALE_ONLY, False), (gcr_capi.PROBLEM_SCALE_ONLY_ORIGINAL, True)):
        e = gcr_capi.Engine(0)
        e.set_problem(problem)
        e.set_features(sc.scale)
        ora = port.replay_scale(sc.scale, smp, sc.thr[0], original=original, nthreads=0)
        want = port.select_best(ora["score"], ora["flags"])
        res = e.run(len(smp), sc.thr, samples=smp)
        assert res.best_index == want
        b = e.best()
        assert abs(b["score"] - ora["score"][want]) <= 1e-11 * abs(ora["score"][want])
        assert int(b["n_inl"][0]) == int(ora["n_inl"][want][0] if ora["n_inl"].ndim == 2 else ora["n_inl"][want])
        assert np.array_equal(b["scale_words"], ora["masks"][want][:len(b["scale_words"])])
        e.close()


def test_selection_cascade_rare_paths(port):
    """The paths of the winner-only selection that ordinary data rarely takes, forced through the engine's test switches: the
    contender slots run out (GCR_CONT_CAP: the stragglers go straight to the FP64 level by their loose bound), and the first contender
    bar turns out too high (GCR_DEBUG_FLOOR_BUMP: no contender above it, or a refined lower bound below it -> the bar is lowered and
    the round repeats).  Same winner, score, counts and masks as the oracle every time."""
    import os
    import gcr_capi
    import gcr_scenes as gs
    sc = gs.make_scene(1500, 500, 0.3, 91)
    smp = gs.draw_samples(30_000, 1500, 500, 92)
    # near-duplicates of the best sample make a crowd of contenders
    ora = port.replay(sc.scale, sc.orient, smp[:4000], sc.thr, want_masks=False, nthreads=0)
    best = port.select_best(ora["score"], ora["flags"])
    smp[10_000:10_200] = smp[best]
    for env in ({"GCR_CONT_CAP": "3"}, {"GCR_CONT_CAP": "40"}, {"GCR_DEBUG_FLOOR_BUMP": "5.0"}, {"GCR_DEBUG_FLOOR_BUMP": "1e9"},
                {"GCR_DEBUG_FLOOR_BUMP": "0.5", "GCR_CONT_CAP": "17"}):
        os.environ.update(env)
        try:
            e = gcr_capi.Engine(0)
        finally:
            for k in env:
                del os.environ[k]
        for n in (len(smp), 12_000, 4000):
            _check_winner(e, port, sc.scale, sc.orient, smp[:n], sc.thr, f"{env} n={n}")
        e.close()


def test_full_size_c3_properties(port):
    """BASELINE config C3 at its full size (50 000 + 20 000 features, 1 000 000 Philox hypotheses) through the C ABI, by properties that do
    not need the oracle to replay a million hypotheses: idempotence, shard invariance (two half ranges and eight eighths combined by the
    selection rule == the whole run), the winner's score / counts / masks == the CPU oracle's MSAC of the winning model, the winning
    hypothesis' minimal sample == the oracle's Philox twin of that slot and its model == the oracle's solve of that sample, and no
    hypothesis of an oracle-replayed slice of 3 000 slots beats it."""
    import gcr_capi
    import gcr_parallel
    import gcr_scenes as gs
    sc = gs.config_scene("C3")
    H, seed = 1_000_000, 2025
    eng = gcr_capi.Engine(0)
    eng.set_features(sc.scale, sc.orient)
    r1 = eng.run(H, sc.thr, seed=seed)
    b1 = eng.best()
    r2 = eng.run(H, sc.thr, seed=seed)
    b2 = eng.best()
    assert (r1.best_index, r1.best_score) == (r2.best_index, r2.best_score) and np.array_equal(b1["scale_words"], b2["scale_words"])
    assert r1.n_scored < r1.n_solved < r1.n_sample_valid < H and r1.n_rechecked == 0
    for parts in (2, 8):
        recs = np.zeros(parts, dtype=gcr_capi.RECORD_DTYPE)
        for k in range(parts):
            first, count = gcr_parallel.shard_range(H, k, parts)
            r = eng.run(count, sc.thr, seed=seed, first_hyp=first)
            recs[k]["index"], recs[k]["score"] = r.best_index, r.best_score
        w = gcr_parallel.select_record(recs)
        assert int(recs[w]["index"]) == r1.best_index and float(recs[w]["score"]) == r1.best_score, parts
    # the winner against the CPU oracle
    val, n_inl, sm, om = port.score_model(sc.scale, sc.orient, b1["model"], sc.thr)
    assert np.array_equal(sm, b1["scale_mask"]) and np.array_equal(om, b1["orient_mask"])
    assert np.array_equal(np.asarray(n_inl, np.uint32), b1["n_inl"]) and abs(val - b1["score"]) <= 1e-10 * val
    smp = port.sample_philox(seed, r1.best_index, 1, len(sc.scale), len(sc.orient))
    one = port.replay(sc.scale, sc.orient, smp, sc.thr)
    assert (one["flags"][0] & 7) == 7 and np.allclose(one["models"][0], b1["model"], rtol=1e-6, atol=1e-12)
    assert abs(one["score"][0] - b1["score"]) <= 1e-10 * val
    # the planted plane is found, and nothing in an oracle-replayed slice beats the winner
    assert np.allclose(b1["model"][:2], sc.model[:2], atol=5e-5)
    sl = port.sample_philox(seed, 400_000, 3000, len(sc.scale), len(sc.orient))
    ora = port.replay(sc.scale, sc.orient, sl, sc.thr, want_masks=False, nthreads=0)
    ok = (ora["flags"] & 6) == 6
    assert ora["score"][ok].max() <= b1["score"]
    eng.close()
