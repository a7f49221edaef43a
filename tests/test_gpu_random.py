"""Randomised GPU sweeps beyond the fixed fixtures: scale-only problems, the device sampler at awkward sizes, and the full
estimator's self-consistency against the CPU oracle.  Seeds are fixed: failures are reproducible."""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SCORE_RTOL = 1e-5
MODEL_RTOL = 1e-6


@pytest.mark.parametrize("original", [False, True])
def test_randomised_scale_only_scenes_vs_oracle(port, original):
    import gcr_capi
    import gcr_scenes as gs
    e = gcr_capi.Engine(0)
    e.set_problem(gcr_capi.PROBLEM_SCALE_ONLY_ORIGINAL if original else gcr_capi.PROBLEM_SCALE_ONLY)
    rng = np.random.Generator(np.random.PCG64(424242 + int(original)))
    for case in range(24):
        n = int(rng.integers(3, 900))
        ratio = float(rng.choice([0.1, 0.4, 0.8]))
        thr = float(rng.choice([0.0125, 0.05, 0.2, 0.6]))
        sc = gs.make_scene(n, 4, ratio, 7000 + case, thr_scale=thr)
        scale = sc.scale.copy()
        if case % 4 == 0 and n > 10:
            scale[3] = scale[2]
            scale[5, 2] = 0.0
            scale[6, 2] = -1.5
            scale[8, 2] = 1e-11
            scale[9, 2] = 1e11
        H = int(rng.integers(1, 1200))
        smp = gs.draw_samples_scale3(H, n, 50 + case)
        ora = port.replay_scale(scale, smp, thr, original=original, nthreads=0)
        e.set_features(scale)
        res = e.run(H, [thr, 0.0], samples=smp, flags=gcr_capi.RUN_EMIT_MASKS)
        dev = e.fetch(masks=True)
        label = f"scale-only case {case} (n={n}, thr={thr}, H={H}, original={original})"
        assert np.array_equal(dev["flags"] & 7, ora["flags"] & 7), label
        solved = (ora["flags"] & 2) > 0
        assert np.allclose(dev["models"][solved], ora["models"][solved], rtol=MODEL_RTOL, atol=1e-300), label
        assert np.array_equal(dev["n_inl"], ora["n_inl"]), label
        assert np.array_equal(dev["masks"], ora["masks"]), label
        assert np.allclose(dev["score"], ora["score"], rtol=SCORE_RTOL, atol=1e-9), label
        want = port.select_best(ora["score"], ora["flags"])
        if want < 0 or np.sum(np.isclose(ora["score"], ora["score"][want], rtol=1e-12)) == 1:
            assert res.best_index == want, label
        res0 = e.run(H, [thr, 0.0], samples=smp)  # winner-only mode
        assert (res0.best_index, res0.best_score) == (res.best_index, res.best_score), label
    e.close()


def test_device_sampler_matches_oracle_at_awkward_sizes(port):
    """Philox4x32-10 draws (key = seed, counter = global slot): 2 distinct indices per class for class sizes from the
    minimum (2) up, window offsets beyond 2^32."""
    import gcr_capi
    import gcr_scenes as gs
    e = gcr_capi.Engine(0)
    rng = np.random.Generator(np.random.PCG64(99))
    for ns, no in [(2, 2), (2, 3), (3, 2), (5, 7), (33, 31), (1000, 2), (2, 1000), (65537, 4099)]:
        sc = gs.make_scene(ns, no, 0.5, 1)
        e.set_features(sc.scale, sc.orient)
        for first in (0, 12345, (1 << 32) - 17, (1 << 40) + 3):
            seed = int(rng.integers(0, 1 << 62))
            H = 777
            e.run(H, sc.thr, seed=seed, first_hyp=first, flags=gcr_capi.RUN_KEEP_PER_HYP)
            got = e.fetch(samples=True)["samples"]
            want = port.sample_philox(seed, first, H, ns, no)
            assert np.array_equal(got, want), (ns, no, first)
            assert (got[:, 0] != got[:, 1]).all() and (got[:, 2] != got[:, 3]).all()
            assert got[:, :2].max() < ns and got[:, 2:].max() < no and got.min() >= 0
    e.close()


def test_estimator_result_is_consistent_with_the_oracle(port):
    """Full call (device hypothesis stage, local optimisation, refit) on seeded scenes: whatever model comes back, its
    inlier masks are exactly the oracle's MSAC labelling of that model, its support is at least the raw winner's, and it
    explains the planted plane."""
    import gcr_scenes as gs
    import pygcransac
    rng = np.random.Generator(np.random.PCG64(7))
    for case in range(12):
        ns, no = int(rng.integers(60, 1500)), int(rng.integers(40, 500))
        ratio = float(rng.choice([0.25, 0.5, 0.8]))
        sc = gs.make_scene(ns, no, ratio, 9000 + case)
        H, sm, om, model, st = pygcransac.findRectifyingHomographySIFT(
            sc.scale, sc.orient, float(sc.thr[0]), float(sc.thr[1]), 0.0, 4000, 4000, 50, seed=case, return_stats=True)
        assert H is not None, case
        m4 = np.array([model.h7, model.h8, model.alpha, model.phi])
        val, n, smo, omo = port.score_model(sc.scale, sc.orient, m4, sc.thr)
        assert np.array_equal(sm, smo) and np.array_equal(om, omo), case
        assert st["final_score"] == pytest.approx(val, rel=1e-9), case
        assert st["final_score"] >= st["best_raw_score"] * (1 - 1e-12), case
        assert np.allclose(np.asarray(H).ravel()[[6, 7, 8]], [model.h7, model.h8, 1.0], rtol=1e-12), case  # model.h:211-226, identity normaliser
        err = np.hypot(model.h7 - sc.model[0], model.h8 - sc.model[1]) / np.hypot(sc.model[0], sc.model[1])
        assert err < 0.05, (case, err)


def test_engine_reuse_across_sizes_problems_and_modes(port):
    """One engine handle driven through growing / shrinking feature sets, batch sizes, problem modes and run modes in a
    shuffled order: buffer re-allocation, counter reset and stale state must never leak into a result."""
    import gcr_capi
    import gcr_scenes as gs
    e = gcr_capi.Engine(0)
    rng = np.random.Generator(np.random.PCG64(31415))
    for step in range(40):
        problem = int(rng.choice([0, 0, 0, 1, 2]))
        ns, no = int(rng.integers(3, 3000)), int(rng.integers(2, 900))
        H = int(rng.choice([1, 7, 129, 1000, 5000, 20000]))
        sc = gs.make_scene(ns, no, float(rng.choice([0.2, 0.6])), 100 + step)
        mode = int(rng.choice([0, gcr_capi.RUN_KEEP_PER_HYP, gcr_capi.RUN_EMIT_MASKS, gcr_capi.RUN_ALL_EXACT]))
        e.set_problem(problem)
        if problem == 0:
            e.set_features(sc.scale, sc.orient)
            smp = gs.draw_samples(H, ns, no, step)
            ora = port.replay(sc.scale, sc.orient, smp, sc.thr, nthreads=0)
            thr = sc.thr
        else:
            e.set_features(sc.scale)
            smp = gs.draw_samples_scale3(H, ns, step)
            ora = port.replay_scale(sc.scale, smp, sc.thr[0], original=problem == 2, nthreads=0)
            thr = [sc.thr[0], 0.0]
        res = e.run(H, thr, samples=smp, flags=mode)
        want = port.select_best(ora["score"], ora["flags"])
        label = f"step {step}: problem {problem}, {ns}+{no} features, H={H}, mode {mode}"
        assert res.n_solved == int(((ora["flags"] & 2) > 0).sum()), label
        if want < 0:
            assert res.best_index == -1, label
            continue
        ties = np.isclose(ora["score"], ora["score"][want], rtol=1e-12) & ((ora["flags"] & 4) > 0)
        if ties.sum() == 1:
            assert res.best_index == want, label
        assert res.best_score == pytest.approx(ora["score"][want], rel=1e-9), label
        b = e.best()
        if ties.sum() == 1:
            assert np.array_equal(np.concatenate([b["scale_words"], b["orient_words"]]), ora["masks"][want]), label
            assert tuple(b["n_inl"]) == tuple(ora["n_inl"][want]), label
        if mode & (gcr_capi.RUN_KEEP_PER_HYP | gcr_capi.RUN_EMIT_MASKS):
            dev = e.fetch(masks=bool(mode & gcr_capi.RUN_EMIT_MASKS))
            assert np.array_equal(dev["n_inl"], ora["n_inl"]), label
            if mode & gcr_capi.RUN_EMIT_MASKS:
                assert np.array_equal(dev["masks"], ora["masks"]), label
    e.close()


def test_engines_in_concurrent_threads(port):
    """Four engines driven from four host threads at once (one stream each): results equal the single-threaded ones."""
    import threading
    import gcr_capi
    import gcr_scenes as gs
    scenes = [gs.make_scene(900 + 50 * i, 300 + 10 * i, 0.3, 40 + i) for i in range(4)]
    want, got, errs = [], [None] * 4, []
    e0 = gcr_capi.Engine(0)
    for sc in scenes:
        e0.set_features(sc.scale, sc.orient)
        r = e0.run(30000, sc.thr, seed=5)
        want.append((r.best_index, r.best_score, tuple(r.best_n_inl)))
    e0.close()

    def work(i):
        try:
            e = gcr_capi.Engine(0)
            for _ in range(5):
                e.set_features(scenes[i].scale, scenes[i].orient)
                r = e.run(30000, scenes[i].thr, seed=5)
                got[i] = (r.best_index, r.best_score, tuple(r.best_n_inl))
                assert got[i] == want[i]
            e.close()
        except Exception as ex:  # surfaced in the main thread
            errs.append((i, repr(ex)))

    ts = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
    assert got == want
