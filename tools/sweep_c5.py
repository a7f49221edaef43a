"""C5 (BASELINE.json configs[4]): inlier-ratio x threshold sweep at 1M device-sampled hypotheses, plus roofline fraction
vs feature count.  Run on the GPU box; writes a table to gpurun_out/sweep_c5.txt (copied to profiles/ for the round).

Threshold "level" p in {1, 2, 4, 8}: thr_orientation = p degrees, thr_scale = 0.05 p (the reference's
thresholds are a log-scale ratio and radians, not pixels).  Every winner is checked against the CPU oracle's MSAC of the
same model (masks identical, score equal)."""
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")]
import gcr_capi  # noqa: E402
import gcr_scenes as gs  # noqa: E402
from oracle_lib import Oracle  # noqa: E402

FP32_PEAK = 148 * 128 * 2 * 1.965e9
H = int(os.environ.get("SWEEP_H", 1_000_000))


def run(eng, port, sc, n_hyp, seed):
    eng.set_features(sc.scale, sc.orient)
    for _ in range(2):
        res = eng.run(n_hyp, sc.thr, seed=seed)
    rs = [eng.run(n_hyp, sc.thr, seed=seed, flags=gcr_capi.RUN_TIME_KERNELS) for _ in range(3)]
    res = rs[0]
    stage = float(np.median([r.device_ms for r in rs]))
    kern = float(np.median([r.score_kernel_ms for r in rs]))
    b = eng.best()
    ok = "-"
    if res.best_index >= 0:
        val, n_inl, sm, om = port.score_model(sc.scale, sc.orient, b["model"], sc.thr)
        ok = "ok" if (np.array_equal(sm, b["scale_mask"]) and np.array_equal(om, b["orient_mask"]) and
                      abs(val - b["score"]) <= 1e-9 * max(1.0, abs(val))) else "MISMATCH"
    ns, no = len(sc.scale), len(sc.orient)
    n_fast = res.n_solved - res.n_exact_hyp
    frac = n_fast * (ns * 11.0 + no * 30.0) / (kern * 1e-3) / FP32_PEAK
    err = float(np.hypot(*(b["model"][:2] - sc.model[:2])) / np.hypot(*sc.model[:2])) if res.best_index >= 0 else float("nan")
    return dict(stage_ms=stage, kern_ms=kern, evals_s=n_hyp * (ns + no) / (stage * 1e-3), kernel_evals_s=n_fast * (ns + no) / (kern * 1e-3),
                frac=frac, solved=res.n_solved, gb=res.n_rechecked, exact=res.n_exact_hyp, cands=res.n_candidates, ok=ok, err=err,
                n_inl=tuple(int(x) for x in b["n_inl"]))


def main():
    port = Oracle("port")
    eng = gcr_capi.Engine(0)
    lines = [f"# C5 sweep, {H} Philox hypotheses per run, 2000 + 500 features (seed 777 + id), 1 x B200, device-timed stage",
             f"{'inliers':>8s} {'level':>5s} {'stage_ms':>9s} {'kernel_ms':>9s} {'evals/s (all slots)':>20s} {'kernel evals/s':>15s} {'fp32 frac':>9s} "
             f"{'solved':>8s} {'band pairs':>10s} {'exact':>6s} {'cands':>5s} {'winner inliers':>15s} {'line err':>9s} {'oracle':>8s}"]
    k = 0
    for ratio in (0.05, 0.10, 0.20, 0.30, 0.50):
        for p in (1, 2, 4, 8):
            sc = gs.make_scene(2000, 500, ratio, 777 + k, thr_scale=0.05 * p, thr_orient=math.radians(p))
            r = run(eng, port, sc, H, 5000 + k)
            lines.append(f"{int(ratio * 100):7d}% {p:5d} {r['stage_ms']:9.3f} {r['kern_ms']:9.3f} {r['evals_s']:20.3e} {r['kernel_evals_s']:15.3e} "
                         f"{r['frac']:9.3f} {r['solved']:8d} {r['gb']:10d} {r['exact']:6d} {r['cands']:5d} {str(r['n_inl']):>15s} {r['err']:9.2e} {r['ok']:>8s}")
            k += 1
    lines.append("")
    lines.append(f"# roofline fraction vs feature count: 10% inliers, level 3, {H // 4} hypotheses")
    lines.append(f"{'Ns':>7s} {'No':>7s} {'stage_ms':>9s} {'kernel_ms':>9s} {'kernel evals/s':>15s} {'fp32 frac':>9s} {'solved':>8s} {'oracle':>8s}")
    for ns, no in ((500, 125), (2000, 500), (8000, 2000), (20000, 8000), (50000, 20000)):
        sc = gs.make_scene(ns, no, 0.10, 2025, thr_scale=0.15, thr_orient=math.radians(3))
        r = run(eng, port, sc, H // 4, 9)
        lines.append(f"{ns:7d} {no:7d} {r['stage_ms']:9.3f} {r['kern_ms']:9.3f} {r['kernel_evals_s']:15.3e} {r['frac']:9.3f} {r['solved']:8d} {r['ok']:>8s}")
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    open(os.path.join(ROOT, "gpurun_out", "sweep_c5.txt"), "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
