"""Device time of the hypothesis stage for the scale-only problems (three_sift.hpp / three_sift_original.hpp through the same
kernels, GCR_PROBLEM_SCALE_ONLY[_ORIGINAL]) at C2's and C3's scale classes: python tools/time_scale_only.py   (GPU box)
Winner-only runs, Philox triples, L2 flushed between runs; CUDA events of GCR_RUN_TIME_KERNELS (whole stage, scoring kernel alone)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")]
import torch, gcr_capi, gcr_scenes as gs
PEAK = 148 * 128 * 2 * 1.965e9  # nominal FP32 flop/s
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for label, ns, n_hyp in (("C2-sized", 2000, 100_000), ("C3-sized", 50_000, 1_000_000)):
    sc = gs.make_scene(ns, 16, 0.3 if ns < 10000 else 0.1, 91)
    for problem, pname in ((gcr_capi.PROBLEM_SCALE_ONLY, "three_sift"), (gcr_capi.PROBLEM_SCALE_ONLY_ORIGINAL, "three_sift_original")):
        e = gcr_capi.Engine(0)
        e.set_problem(problem)
        e.set_features(sc.scale)
        st, kt, res = [], [], None
        for rep in range(12):
            flush.zero_(); torch.cuda.synchronize()
            res = e.run(n_hyp, [sc.thr[0], 0.0], seed=5, flags=gcr_capi.RUN_TIME_KERNELS)
            if rep >= 2:
                st.append(res.device_ms); kt.append(res.score_kernel_ms)
        plain = []
        for rep in range(12):  # without the timing events (chained launches): wall time of the synchronous call
            flush.zero_(); torch.cuda.synchronize()
            import time
            t0 = time.perf_counter(); e.run(n_hyp, [sc.thr[0], 0.0], seed=5); plain.append((time.perf_counter() - t0) * 1e3)
        k_ms, s_ms = float(np.median(kt)), float(np.median(st))
        evals = res.n_scored * ns
        print(f"{label} {pname}: {ns} scale features x {n_hyp} hypotheses ({res.n_scored} scored): stage {s_ms:.4f} ms, call {np.median(plain[2:]):.4f} ms, "
              f"scoring kernel {k_ms:.4f} ms = {evals / (k_ms * 1e-3):.3e} scored evals/s = {11 * evals / (k_ms * 1e-3) / PEAK:.3f} of the nominal FP32 peak "
              f"(11 flop per scale evaluation); all slots {n_hyp * ns / (s_ms * 1e-3):.3e} evals/s; winner {res.best_index} score {res.best_score:.3f} inliers {res.best_n_inl[0]}")
        e.close()
