"""Interleaved timing of engine settings read at gcr_create (GPU box): python tools/ab_env.py C2 - GCR_RESIDENT=0 GCR_COST_O=27 ...
One engine per setting ("-" = defaults), rounds alternate between them so that clock / thermal drift cancels; FLUSH=1 flushes
L2 before every run as bench.py does.  The winners of all settings must agree.  Times are the engine's event brackets
(GCR_RUN_TIME_KERNELS: chained launches off, event clock ticks every ~2 us); tools/ab_step.py measures whole steps as bench.py does."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")]
import torch, gcr_capi, gcr_scenes as gs
name, settings = sys.argv[1], sys.argv[2:]
c = gs.CONFIGS[name]; sc = gs.config_scene(name)
smp = torch.from_numpy(gs.draw_samples(c["n_hyp"], c["n_scale"], c["n_orient"], gs.SAMPLE_SEED)).cuda()
engines = []
for st in settings:
    kv = [] if st == "-" else [x.split("=") for x in st.split(",")]
    for k, v in kv: os.environ[k] = v
    e = gcr_capi.Engine(0); e.set_features(sc.scale, sc.orient); engines.append(e)
    for k, _ in kv: del os.environ[k]
res = {st: ([], [], set()) for st in settings}
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda") if os.environ.get("FLUSH") else None
for rnd in range(14):
    for st, e in zip(settings, engines):
        if flush is not None: flush.zero_(); torch.cuda.synchronize()
        r = e.run(c["n_hyp"], sc.thr, samples_ptr=smp.data_ptr(), flags=gcr_capi.RUN_TIME_KERNELS)
        if rnd >= 2:
            res[st][0].append(r.score_kernel_ms); res[st][1].append(r.device_ms); res[st][2].add((r.best_index, r.best_score))
for st in settings:
    k, s = np.array(res[st][0]), np.array(res[st][1])
    print(f"{name} {st:28s} kernel ms median {np.median(k):.4f} min {k.min():.4f} | stage ms median {np.median(s):.4f} min {s.min():.4f} | winner {sorted(res[st][2])}")
