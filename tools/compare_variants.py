"""Interleaved timing of engine builds (GPU box): python tools/compare_variants.py C3 lib_a.so lib_b.so ...
Each library is dlopened separately; rounds alternate between them so that clock / thermal drift cancels."""
import ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")]
import torch, gcr_capi, gcr_scenes as gs
name, libs = sys.argv[1], sys.argv[2:]
c = dict(gs.CONFIGS[name]); sc = gs.config_scene(name)
if os.environ.get("HYP"): c["n_hyp"] = int(os.environ["HYP"])
smp = torch.from_numpy(gs.draw_samples(c["n_hyp"], c["n_scale"], c["n_orient"], gs.SAMPLE_SEED)).cuda()
engines = []
for lib in libs:
    gcr_capi.LIB_PATH = os.path.join(ROOT, lib)
    e = gcr_capi.Engine(0); e.set_features(sc.scale, sc.orient); engines.append(e)
res = {lib: ([], []) for lib in libs}
win = {}
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda") if os.environ.get("FLUSH") else None  # > L2: cold caches per run, as bench.py
for rnd in range(12):
    for lib, e in zip(libs, engines):
        if flush is not None: flush.zero_(); torch.cuda.synchronize()
        r = e.run(c["n_hyp"], sc.thr, samples_ptr=smp.data_ptr(), flags=gcr_capi.RUN_TIME_KERNELS)
        win[lib] = (r.best_index, r.best_score, r.n_scored, r.n_refined, r.n_candidates, r.n_exact_hyp)
        if rnd >= 2:
            res[lib][0].append(r.score_kernel_ms); res[lib][1].append(r.device_ms)
for lib in libs:
    k, s = np.array(res[lib][0]), np.array(res[lib][1])
    print(f"{name} {lib:28s} kernel ms median {np.median(k):.4f} min {k.min():.4f} | stage ms median {np.median(s):.4f} min {s.min():.4f} | winner {win[lib]}")
