"""GPU tests of the device non-minimal fits (gcr_fit_models): the least-squares step of the
local optimisation and of the final refit, against (a) the REFERENCE's own estimateModelNonminimal outputs stored in
tests/golden/host_refit.npz and scale_only.npz, (b) oracle/_ref (the reference compiled against the shims) on fresh
index subsets when that library travelled with the snapshot, (c) the host C++ fit of csrc/host/gcr_refit.hpp.

Bar: vanishing line (h7, h8, 1), alpha <= 1e-6 relative; phi identical up to the bin-mean rounding (1e-9).
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MODEL_RTOL = 1e-6
PHI_ATOL = 1e-9


@pytest.fixture(scope="module")
def eng():
    import gcr_capi
    e = gcr_capi.Engine(0)
    yield e
    e.close()


@pytest.fixture(scope="module")
def pg():
    import pygcransac
    return pygcransac


@pytest.fixture(scope="module")
def ref():
    from oracle_lib import Oracle
    try:
        return Oracle("reference")
    except Exception:  # oracle/_ref was not built (no /root/reference on this box and no prebuilt library)
        return None


def test_device_fit_matches_reference_golden(eng, golden_dir):
    g = np.load(os.path.join(golden_dir, "host_refit.npz"))
    for k in range(int(g["n_fits"])):
        s_rows, o_rows = g[f"fit{k}_scale"], g[f"fit{k}_orient"]
        eng.set_features(s_rows, o_rows)
        models, ok = eng.fit_models([np.arange(len(s_rows))], [np.arange(len(o_rows))])
        assert bool(ok[0]) == bool(int(g[f"fit{k}_ok"])), k
        if not ok[0]:
            continue
        want = g[f"fit{k}_model7"]
        assert np.allclose(models[0, :3], want[3:6], rtol=MODEL_RTOL, atol=1e-12), (k, models[0], want)
        assert abs(models[0, 3] - want[6]) < PHI_ATOL, (k, models[0, 3], want[6])


def test_device_scale_fit_matches_reference_golden(golden_dir):
    import gcr_capi
    g = np.load(os.path.join(golden_dir, "scale_only.npz"))
    for problem, name in ((gcr_capi.PROBLEM_SCALE_ONLY, "std"), (gcr_capi.PROBLEM_SCALE_ONLY_ORIGINAL, "orig")):
        e = gcr_capi.Engine(0)
        e.set_problem(problem)
        for k in range(int(g["n_fits"])):
            rows = g[f"fit{k}_rows"]
            e.set_features(rows)
            models, ok = e.fit_models([np.arange(len(rows))])
            assert bool(ok[0]) == bool(int(g[f"fit{k}_{name}_ok"])), (name, k)
            if ok[0]:
                want = g[f"fit{k}_{name}_model6"]
                assert np.allclose(models[0, :3], want[3:6], rtol=MODEL_RTOL, atol=1e-12), (name, k, models[0], want)
        rows = g["fit0_rows"]
        e.set_features(rows)
        assert not e.fit_models([np.arange(3)])[1][0]  # a minimal sample is not a non-minimal fit
        e.close()


def test_batched_device_fits_match_host_fit_and_reference(eng, pg, ref):
    """One call with 50 local-optimisation sized fits (14 + 14 rows), mid-sized ones and the all-inliers refit."""
    import gcr_scenes as gs
    sc = gs.config_scene("C1")
    eng.set_features(sc.scale, sc.orient)
    rng = np.random.default_rng(11)
    si, oi = np.flatnonzero(sc.scale_is_inlier), np.flatnonzero(sc.orient_is_inlier)
    s_lists = [rng.choice(si, 14, replace=False) for _ in range(50)]
    o_lists = [rng.choice(oi, 14, replace=False) for _ in range(50)]
    s_lists += [rng.choice(si, 90, replace=False), si, rng.permutation(len(sc.scale))[:300], si[:2], si[:1], si[:5]]
    o_lists += [rng.choice(oi, 60, replace=False), oi, rng.permutation(len(sc.orient))[:120], oi[:2], oi[:7], oi[:1]]
    models, ok = eng.fit_models(s_lists, o_lists)
    assert ok[:53].all() and not ok[53:].any()  # 2 + 2 rows is the minimal solver's case; < 2 rows of a class: no model
    for k in range(53):
        m = pg._fit_non_minimal(sc.scale[s_lists[k]], sc.orient[o_lists[k]])
        assert m is not None
        assert np.allclose(models[k, :3], [m.h7, m.h8, m.alpha], rtol=MODEL_RTOL, atol=1e-12), (k, models[k], m.h7, m.h8, m.alpha)
        assert abs(models[k, 3] - m.phi) < PHI_ATOL, (k, models[k, 3], m.phi)
        if ref is not None and k % 5 == 0:
            r7 = ref.fit_nonminimal(sc.scale[s_lists[k]], sc.orient[o_lists[k]])
            assert r7 is not None
            assert np.allclose(models[k, :3], r7[3:6], rtol=MODEL_RTOL, atol=1e-12), (k, models[k], r7)
            assert abs(models[k, 3] - r7[6]) < PHI_ATOL
    # a second call gives the same bits (fixed reduction order)
    again, _ = eng.fit_models(s_lists, o_lists)
    assert np.array_equal(models[:53], again[:53])


def test_large_refit_millions_of_pairs(eng, pg):
    """All inliers of a C3-sized scene: ~3.6k orientation inliers = 6.5 million pair rows, many row units."""
    import gcr_scenes as gs
    sc = gs.config_scene("C3")
    eng.set_features(sc.scale, sc.orient)
    si, oi = np.flatnonzero(sc.scale_is_inlier), np.flatnonzero(sc.orient_is_inlier)
    models, ok = eng.fit_models([si], [oi])
    assert ok[0]
    m = pg._fit_non_minimal(sc.scale[si], sc.orient[oi])
    assert np.allclose(models[0, :3], [m.h7, m.h8, m.alpha], rtol=MODEL_RTOL, atol=1e-12), (models[0], m.h7, m.h8, m.alpha)
    assert abs(models[0, 3] - m.phi) < PHI_ATOL
    truth = sc.model
    assert np.allclose(models[0, :2], truth[:2], rtol=0.05)  # and it is the scene's plane
    # whichever CTA stores the last of the 790 row units finishes the fit, the units summed in index order: the same bits every time;
    # and fits of very different sizes (row units of 1 024 ... 8 192 rows, lines in shared or global memory) side by side in one call
    for _ in range(3):
        again, _ok = eng.fit_models([si], [oi])
        assert np.array_equal(again, models)
    rng = np.random.default_rng(4)
    s_lists = [si, rng.choice(si, 14, replace=False), rng.choice(si, 600, replace=False), si[:3000], rng.choice(si, 40, replace=False)]
    o_lists = [oi, rng.choice(oi, 14, replace=False), rng.choice(oi, 150, replace=False), oi[:1100], rng.choice(oi, 900, replace=False)]
    mixed, ok = eng.fit_models(s_lists, o_lists)
    assert ok.all() and np.array_equal(mixed[0], models[0])
    for k in range(1, 5):
        alone, _ok = eng.fit_models([s_lists[k]], [o_lists[k]])
        assert np.array_equal(alone[0], mixed[k]), k
        m = pg._fit_non_minimal(sc.scale[s_lists[k]], sc.orient[o_lists[k]])
        assert np.allclose(mixed[k, :3], [m.h7, m.h8, m.alpha], rtol=MODEL_RTOL, atol=1e-12), k


def test_fit_argument_errors(eng):
    import gcr_capi
    import gcr_scenes as gs
    sc = gs.config_scene("C1")
    eng.set_features(sc.scale, sc.orient)
    with pytest.raises(gcr_capi.GcrError):
        eng.fit_models([[0, 1, len(sc.scale)]], [[0, 1, 2]])  # index out of range
    with pytest.raises(gcr_capi.GcrError):
        eng.fit_models([[0, 1, 2]])  # the hybrid problem needs orientation lists


def test_fit_score_models_equals_fit_then_score(eng):
    """gcr_fit_score_models (fits and their scores in one device pass, one synchronisation: the local optimisation's inner RANSAC and the
    final refit in csrc/host/gcr_host.cpp) == gcr_fit_models followed by gcr_score_models, bit for bit -- models (phi: the device's
    histogram mode vs the host's replay of the reference's unordered_map order), ok flags, scores, counts, masks."""
    import gcr_scenes as gs
    sc = gs.make_scene(900, 300, 0.5, 61)
    rng = np.random.default_rng(62)
    si, oi = np.flatnonzero(sc.scale_is_inlier), np.flatnonzero(sc.orient_is_inlier)
    s_lists = [rng.choice(si, int(rng.integers(2, 30)), replace=False) for _ in range(60)]
    o_lists = [rng.choice(oi, int(rng.integers(2, 16)), replace=False) for _ in range(60)]
    s_lists[5], o_lists[5] = s_lists[5][:2], o_lists[5][:2]      # exactly a minimal sample: declined (ok = 0)
    s_lists[7], o_lists[7] = si, oi                                # the refit case: all inliers
    o_lists[9] = np.concatenate([o_lists[9][:4], o_lists[9][:4]])  # duplicated rows: equally heavy histogram bins are likely
    eng.set_features(sc.scale, sc.orient)
    for truncated in (False, True):
        m1, ok1 = eng.fit_models(s_lists, o_lists)
        sc1, n1, w1 = eng.score_models(np.where(ok1[:, None], m1, 0.0), sc.thr, truncated_sq=truncated, masks=True)
        m2, ok2, sc2, n2, w2 = eng.fit_score_models(s_lists, o_lists, sc.thr, truncated_sq=truncated)
        assert np.array_equal(ok1, ok2) and not ok2[5] and ok2[7]
        assert np.array_equal(m1[ok1], m2[ok2])
        assert np.array_equal(sc1[ok1], sc2[ok2]) and np.array_equal(n1[ok1], n2[ok2]) and np.array_equal(w1[ok1], w2[ok2])
        assert (sc2[~ok2] == 0).all() and (n2[~ok2] == 0).all()
        assert sc2[7] > 0.5 * (len(si) + len(oi))
    # scores only
    m3, ok3, sc3, n3, w3 = eng.fit_score_models(s_lists, o_lists, sc.thr, masks=False)
    assert w3 is None and np.array_equal(ok3, ok2)
    # the device resolved equally heavy histogram bins as the host's std::unordered_map does (no fit needed the second scoring pass),
    # also for histograms that grow through every rehash of the map (index lists of up to 300 orientation features)
    all_o = np.arange(len(sc.orient))
    big_lists = [rng.choice(all_o, n, replace=False) for n in (13, 14, 15, 29, 30, 31, 59, 60, 61, 127, 128, 129, 257, 258, 300)]
    m4, ok4, sc4, _, _ = eng.fit_score_models([si] * len(big_lists), big_lists, sc.thr, masks=False)
    m5, ok5 = eng.fit_models([si] * len(big_lists), big_lists)
    assert np.array_equal(m4, m5) and ok4.all()
    assert eng.lib.gcr_debug_phi_fallbacks(eng.h) == 0
