"""gcr_lo_round (one pass of the local optimisation's loop as one launch sequence, include/gcr_b200.h) against the two synchronous calls it
replaces, through the C ABI: labels by gcr_score_models with the relabelling threshold, the sub-samples drawn HERE by the host driver's
sampler (SplitMix64, partial Fisher-Yates on the ascending list of labelled inliers: csrc/host/gcr_host.cpp drawSubset, the role of
GCRANSAC.h:53-80), fits and scores by gcr_fit_score_models, the strict-< update in trial order (GCRANSAC.h:1034).  Everything must agree bit
for bit: the update flag, the model, score, counts and mask of the new best, the generator state after the round, the number of trial
models.  Rounds are chained as the estimator chains them (the output of one is the input of the next) until no trial improves."""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLDEN = 0x9E3779B97F4A7C15
M64 = (1 << 64) - 1


class SplitMix:
    def __init__(self, s):
        self.s = s & M64

    def next(self):
        self.s = (self.s + GOLDEN) & M64
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
        return z ^ (z >> 31)

    def below(self, n):
        return ((self.next() >> 32) * n) >> 32


def draw_subset(src, k, rng):
    scratch = list(src)
    out = []
    for i in range(k):
        j = i + rng.below(len(scratch) - i)
        scratch[i], scratch[j] = scratch[j], scratch[i]
        out.append(scratch[i])
    return np.array(out, np.uint32)


def words_to_indices(words, n):
    bits = np.unpackbits(np.ascontiguousarray(words).view(np.uint8), bitorder="little")[:n]
    return np.flatnonzero(bits).astype(np.uint32)


def two_call_round(eng, model, score, rng_state, thr, n_trials, lim, min_inl, hybrid):
    """The round as the host driver ran it before gcr_lo_round -> dict of everything gcr_lo_round reports."""
    ws = (eng.ns + 31) // 32
    _, _, lab = eng.score_models([model], thr, truncated_sq=True, masks=True, strict=not hybrid)
    in_s = words_to_indices(lab[0, :ws], eng.ns)
    in_o = words_to_indices(lab[0, ws:], eng.no) if hybrid else np.zeros(0, np.uint32)
    out = dict(sampled=0, updated=0, n_models=0, rng_state=rng_state, model=np.array(model, float), score=score, n_labelled=(len(in_s), len(in_o)))
    if len(in_s) <= min_inl[0] or (hybrid and len(in_o) <= min_inl[1]):
        return out
    rng = SplitMix(rng_state)
    k_s, k_o = min(lim[0], len(in_s)), min(lim[1], len(in_o)) if hybrid else 0
    s_lists, o_lists = [], []
    for _ in range(n_trials):
        s_lists.append(draw_subset(in_s, k_s, rng) if k_s < len(in_s) else in_s)
        o_lists.append(draw_subset(in_o, k_o, rng) if k_o < len(in_o) else in_o)
    models, ok, sc, cnt, words = eng.fit_score_models(s_lists, o_lists if hybrid else None, thr)
    out.update(sampled=1, rng_state=rng.s, n_models=int(ok.sum()))
    for t in range(n_trials):
        if ok[t] and out["score"] < sc[t]:
            out.update(updated=1, score=float(sc[t]), model=models[t].copy(), n_inl=cnt[t].copy(), words=words[t].copy())
    return out


def run_both(eng, model0, thr, n_trials, lim, min_inl, hybrid, seed, max_rounds=6):
    import gcr_capi
    score0 = float(eng.score_models([model0], thr)[0][0])
    model, score, state = np.array(model0, float), score0, seed
    rounds = updates = 0
    flags = gcr_capi.SCORE_TRUNCATED_SQ | (0 if hybrid else gcr_capi.SCORE_STRICT)
    for _ in range(max_rounds):
        want = two_call_round(eng, model, score, state, thr, n_trials, lim, min_inl, hybrid)
        io, words = eng.lo_round(model, score, state, thr, n_trials, lim, min_inl, label_flags=flags)
        rounds += 1
        assert io.phi_mismatch == 0
        assert (io.sampled, io.updated) == (want["sampled"], want["updated"]), rounds
        assert tuple(io.n_labelled) == tuple(want["n_labelled"])
        if not io.sampled:
            assert io.rng_state == state and io.score == score
            break
        assert io.n_models_scored == want["n_models"] and io.rng_state == want["rng_state"], rounds
        if not io.updated:
            assert io.score == score and np.array_equal(np.array(io.model[:]), model)
            break
        updates += 1
        assert io.score == want["score"] and np.array_equal(np.array(io.model[:]), want["model"]), rounds
        assert np.array_equal(np.array(io.n_inl[:], np.uint32), want["n_inl"]) and np.array_equal(words, want["words"]), rounds
        model, score, state = np.array(io.model[:]), io.score, io.rng_state
    return rounds, updates


def test_lo_round_equals_the_two_call_round_hybrid():
    import gcr_capi
    import gcr_scenes as gs
    eng = gcr_capi.Engine(0)
    total_updates = 0
    cases = [("C1", None), ("C2", None), (None, gs.make_scene(700, 1300, 0.5, 502, thr_scale=0.1, thr_orient=math.radians(1.5))),
             (None, gs.make_scene(3000, 900, 0.3, 77)), (None, gs.make_scene(90, 40, 0.6, 5))]
    for k, (name, sc) in enumerate(cases):
        sc = gs.config_scene(name) if name else sc
        eng.set_features(sc.scale, sc.orient)
        # a start model a little off the planted plane: the first rounds improve it, a later one does not
        rng = np.random.default_rng(10 + k)
        start = np.array(sc.model) * (1.0 + 0.01 * rng.standard_normal(4))
        rounds, updates = run_both(eng, start, sc.thr, 50, (14, 14), (2, 2), True, seed=0xC0FFEE + k)
        total_updates += updates
        # other trial counts and sub-sample sizes (a single trial; the most the path takes)
        run_both(eng, start, sc.thr, 1, (14, 14), (2, 2), True, seed=7)
        run_both(eng, start, sc.thr, 200, (32, 5), (2, 2), True, seed=8, max_rounds=2)
    assert total_updates >= 3  # the comparison is not about rounds that never improve
    eng.close()


def test_lo_round_small_label_sets():
    """A class taken whole (no more labelled inliers than the sub-sample size: no draws for it), and too few labelled inliers (the
    reference's sampling fails, the optimisation ends: sampled = 0, nothing changes)."""
    import gcr_capi
    import gcr_scenes as gs
    eng = gcr_capi.Engine(0)
    for n_o in (40, 30, 50, 24, 60):  # a scene whose labelled orientation class is no larger than the sub-sample: it is taken whole
        sc = gs.make_scene(400, n_o, 0.15, 31)
        eng.set_features(sc.scale, sc.orient)
        n_lab = eng.score_models([sc.model], sc.thr, truncated_sq=True)[1][0]
        if 2 < n_lab[1] <= 14 < n_lab[0]:
            break
    assert 2 < n_lab[1] <= 14 < n_lab[0]
    rounds, _ = run_both(eng, np.array(sc.model) * [1.002, 0.998, 1.001, 1.0], sc.thr, 50, (14, 14), (2, 2), True, seed=99)
    assert rounds >= 1
    # a model far from everything: (almost) no labelled inliers
    io, words = eng.lo_round([5e-3, 7e-3, 3.0, 0.3], 0.0, 1234, sc.thr, 50, (14, 14), (2, 2))
    assert io.sampled == 0 and io.updated == 0 and words is None and io.rng_state == 1234
    want = two_call_round(eng, [5e-3, 7e-3, 3.0, 0.3], 0.0, 1234, sc.thr, 50, (14, 14), (2, 2), True)
    assert want["sampled"] == 0 and tuple(io.n_labelled) == tuple(want["n_labelled"])
    # argument errors
    with pytest.raises(gcr_capi.GcrError):
        eng.lo_round(sc.model, 0.0, 1, sc.thr, 50, (33, 14), (2, 2))
    with pytest.raises(gcr_capi.GcrError):
        eng.lo_round(sc.model, 0.0, 1, sc.thr, 2000, (14, 14), (2, 2))
    eng.close()


def test_lo_round_scale_only_problems():
    import gcr_capi
    import gcr_scenes as gs
    for problem in (gcr_capi.PROBLEM_SCALE_ONLY, gcr_capi.PROBLEM_SCALE_ONLY_ORIGINAL):
        eng = gcr_capi.Engine(0)
        eng.set_problem(problem)
        sc = gs.make_scene(1500, 10, 0.5, 41)
        eng.set_features(sc.scale)
        m = np.array(sc.model)
        if problem == gcr_capi.PROBLEM_SCALE_ONLY_ORIGINAL:
            m[2] = 1.0 / m[2]  # (its alpha is the reciprocal class scale)
        m[3] = 0.0
        start = m * [1.003, 0.997, 1.002, 1.0]
        rounds, updates = run_both(eng, start, [sc.thr[0], 0.0], 50, (21, 0), (3, 0), False, seed=5)
        assert rounds >= 1
        eng.close()
