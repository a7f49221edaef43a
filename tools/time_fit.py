"""Wall time of one gcr_fit_score_models call (C ABI, arguments packed beforehand) for the two shapes the estimator issues on a small
scene: the refit (1 fit over all inliers, masks) and a local-optimisation batch (50 fits of 14 + 14 features).  GPU box:
python tools/time_fit.py [C1 C3]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")]
import gcr_scenes as gs
import gcr_capi

def timed(lib, eng, s_lists, o_lists, thr, masks, reps=int(os.environ.get("REPS", "200"))):
    K = len(s_lists)
    so = np.zeros(K + 1, np.int64); so[1:] = np.cumsum([len(l) for l in s_lists])
    oo = np.zeros(K + 1, np.int64); oo[1:] = np.cumsum([len(l) for l in o_lists])
    si = np.ascontiguousarray(np.concatenate(s_lists), np.uint32); oi = np.ascontiguousarray(np.concatenate(o_lists), np.uint32)
    thr = np.ascontiguousarray(thr, np.float64)
    W = (eng.ns + 31) // 32 + (eng.no + 31) // 32
    models, ok, score, n_inl = np.zeros((K, 4)), np.zeros(K, np.int32), np.zeros(K), np.zeros((K, 2), np.uint32)
    words = np.zeros((K, W), np.uint32)
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        rc = lib.gcr_fit_score_models(eng.h, K, si.ctypes.data, so.ctypes.data, oi.ctypes.data, oo.ctypes.data, thr.ctypes.data, 0,
                                      models.ctypes.data, ok.ctypes.data, score.ctypes.data, n_inl.ctypes.data, words.ctypes.data if masks else None)
        ts.append((time.perf_counter() - t0) * 1e6)
        assert rc == 0
    ts = np.array(ts[min(10, len(ts) - 1):])
    return np.median(ts), ts.min()

lib = gcr_capi.load_library()
for name in (sys.argv[1:] or ["C1", "C3"]):
    sc = gs.config_scene(name)
    eng = gcr_capi.Engine(0)
    eng.set_features(sc.scale, sc.orient)
    si, oi = np.flatnonzero(sc.scale_is_inlier), np.flatnonzero(sc.orient_is_inlier)
    rng = np.random.default_rng(3)
    med, mn = timed(lib, eng, [si], [oi], sc.thr, True)
    print(f"{name}: refit shape (1 fit of {len(si)}+{len(oi)}, masks): median {med:.1f} us, min {mn:.1f} us")
    med, mn = timed(lib, eng, [rng.choice(si, 14, replace=False) for _ in range(50)], [rng.choice(oi, 14, replace=False) for _ in range(50)], sc.thr, False)
    print(f"{name}: LO batch shape (50 fits of 14+14, no masks): median {med:.1f} us, min {mn:.1f} us")
    med, mn = timed(lib, eng, [rng.choice(si, 14, replace=False)], [rng.choice(oi, 14, replace=False)], sc.thr, False)
    print(f"{name}: one small fit (14+14): median {med:.1f} us, min {mn:.1f} us")
    # one pass of the local optimisation's loop as gcr_lo_round issues it (labels, draws, 50 fits, scores, update)
    import ctypes as C
    start = np.array(sc.model) * (1.0 + 0.01 * np.random.default_rng(1).standard_normal(4))
    ts = []
    for _ in range(int(os.environ.get("REPS", "200"))):
        t0 = time.perf_counter()
        io, words = eng.lo_round(start, 0.0, 12345, sc.thr, 50, (14, 14), (2, 2))
        ts.append((time.perf_counter() - t0) * 1e6)
    ts = np.array(ts[min(10, len(ts) - 1):])
    print(f"{name}: gcr_lo_round (labels + 50 trials of 14+14 + update; Python wrapper included): median {np.median(ts):.1f} us, min {ts.min():.1f} us")
    eng.close()
