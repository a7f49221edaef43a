"""Wall-time breakdown of the full estimator (pygcransac.findRectifyingHomographySIFT through the C ABI) on the
BASELINE scenes (GPU box): python tools/time_estimator.py [C1 C3 ...]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")]
import gcr_scenes as gs
import pygcransac
for name in (sys.argv[1:] or ["C1", "C3"]):
    c = gs.CONFIGS[name]; sc = gs.config_scene(name)
    iters = c["n_hyp"]
    reps = int(os.environ.get("REPS", "25"))
    rows = []
    for rep in range(reps):
        t0 = time.perf_counter()
        H, ms, mo, model, st = pygcransac.findRectifyingHomographySIFT(
            sc.scale, sc.orient, sc.thr[0], sc.thr[1], 0.0, iters, iters, 50, seed=7, return_stats=True)
        wall = (time.perf_counter() - t0) * 1e3
        if rep >= 2:
            rows.append([wall, st['setup_ms'], st['loop_ms'], st['hypothesis_stage_ms'], st['lo_ms'], st['refit_ms']])
    wall, setup, loop, dev, lo, refit = np.median(np.array(rows), axis=0)
    print(f"{name}: wall {wall:.3f} ms (median of {len(rows)}, min {min(r[0] for r in rows):.3f}) | setup {setup:.3f} loop {loop:.3f} (device {dev:.3f}) "
          f"lo {lo:.3f} refit {refit:.3f} | inliers {int(ms.sum())}+{int(mo.sum())} gc {st['graph_cut_number']} "
          f"lo_models {st['lo_models_scored']} refit_accepted {st['refit_accepted']}")
