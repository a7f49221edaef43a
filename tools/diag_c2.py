"""Diagnostic (GPU box): C2 replay, fast path vs oracle, score-error anatomy."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "graph-cut-ransac_b200"))
import gcr_capi, gcr_scenes as gs
from oracle_lib import Oracle
sc = gs.config_scene("C2")
H = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
smp = gs.draw_samples(H, 2000, 500, gs.SAMPLE_SEED)
t = time.time(); ora = Oracle("port").replay(sc.scale, sc.orient, smp, sc.thr, nthreads=0); print("oracle s", time.time() - t, "threads", ora["threads"])
eng = gcr_capi.Engine(0); eng.set_features(sc.scale, sc.orient)
for rep in range(3):
    res = eng.run(H, sc.thr, samples=smp, flags=gcr_capi.RUN_KEEP_PER_HYP | gcr_capi.RUN_TIME_KERNELS)
    print(f"run {rep}: stage {res.device_ms:.3f} ms, score kernel {res.score_kernel_ms:.3f} ms, solved {res.n_solved}, rechecked {res.n_rechecked}, exact hyps {res.n_exact_hyp}, cands {res.n_candidates}")
dev = eng.fetch()
rel = np.abs(dev["score"] - ora["score"]) / np.maximum(ora["score"], 1e-9)
print("flags equal", np.array_equal(dev["flags"] & 7, ora["flags"]), "counts equal", np.array_equal(dev["n_inl"], ora["n_inl"]))
print("rel err percentiles (solved)", np.percentile(rel[(ora["flags"] & 2) > 0], [50, 90, 99, 99.9, 100]))
idx = np.argsort(-rel)[:15]
for i in idx:
    m = ora["models"][i]
    print(i, "rel %.2e" % rel[i], "score %.4f/%.4f" % (dev["score"][i], ora["score"][i]), "n", ora["n_inl"][i], "h7 %.2e h8 %.2e a %.3e" % (m[0], m[1], m[2]), "flags", dev["flags"][i])
bad = rel > 1e-5
print("n bad", bad.sum(), "of solved", ((ora["flags"] & 2) > 0).sum())
if bad.any():
    mm = ora["models"][bad]
    print("bad: |h|max pct", np.percentile(np.maximum(abs(mm[:, 0]), abs(mm[:, 1])), [0, 50, 100]), "alpha pct", np.percentile(mm[:, 2], [0, 50, 100]))
