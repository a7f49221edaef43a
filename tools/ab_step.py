"""Interleaved A/B of engine settings read at gcr_create, measured the way bench.py measures (GPU box):
    python tools/ab_step.py C2 - GCR_PDL=0 GCR_RESIDENT=0,GCR_PDL=0 ...
per setting: device step (torch events around an asynchronous gcr_run on the engine stream, L2 flushed before every step) and the
end-to-end step (gcr_set_features + gcr_run + gcr_best with pinned host buffers, host clock).  Winners must agree."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")]
import torch, gcr_capi, gcr_scenes as gs
C = gcr_capi.C
name, settings = sys.argv[1], sys.argv[2:]
c = gs.CONFIGS[name]; sc = gs.config_scene(name)
ns, no, n_hyp = c["n_scale"], c["n_orient"], c["n_hyp"]
smp = gs.draw_samples(n_hyp, ns, no, gs.SAMPLE_SEED)
d_smp = torch.from_numpy(smp).cuda()
h_scale, h_orient, h_smp = (torch.from_numpy(a).pin_memory() for a in (sc.scale, sc.orient, smp))
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
d_rec = torch.zeros(8, dtype=torch.float64, device="cuda")
engines = []
DEFAULT_LIB = gcr_capi.LIB_PATH
for st in settings:
    kv = [] if st == "-" else [x.split("=") for x in st.split(",")]
    for k, v in kv: os.environ[k] = v
    gcr_capi.LIB_PATH = os.path.join(ROOT, os.environ["LIB"]) if "LIB" in os.environ else DEFAULT_LIB  # LIB=gpurun_variants/x.so: another build
    e = gcr_capi.Engine(0); e.set_stream(stream.cuda_stream); e.set_features(sc.scale, sc.orient)
    e2 = gcr_capi.Engine(0)
    engines.append((e, e2))
    for k, _ in kv: del os.environ[k]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
thr_c = (C.c_double * 2)(*sc.thr)
out = gcr_capi.GcrResult()
b_model, b_n = np.zeros(4), np.zeros(2, np.uint32)
b_sm, b_om = np.zeros((ns + 31) // 32, np.uint32), np.zeros((no + 31) // 32, np.uint32)
b_score, b_idx = C.c_double(), C.c_int64()
res = {st: ([], [], set()) for st in settings}
for rnd in range(8):
    for st, (e, e2) in zip(settings, engines):
        evs = []
        for _ in range(10):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            e.run(n_hyp, sc.thr, samples_ptr=d_smp.data_ptr(), record_ptr=d_rec.data_ptr(), sync=False)
            b.record(stream)
            evs.append((a, b))
        torch.cuda.synchronize()
        r = e.run(n_hyp, sc.thr, samples_ptr=d_smp.data_ptr(), sync=True)
        lib, hh = e2.lib, e2.h
        def e2e_step():
            rc = lib.gcr_set_features(hh, C.c_void_p(h_scale.data_ptr()), ns, C.c_void_p(h_orient.data_ptr()), no, 0)
            rc |= lib.gcr_run(hh, n_hyp, 0, 0, C.c_void_p(h_smp.data_ptr()), thr_c, 0, None, C.byref(out))
            rc |= lib.gcr_best(hh, C.c_void_p(b_model.ctypes.data), C.byref(b_score), C.c_void_p(b_n.ctypes.data), C.c_void_p(b_sm.ctypes.data),
                               C.c_void_p(b_om.ctypes.data), C.byref(b_idx))
            assert rc == 0, lib.gcr_last_error(hh)
        for _ in range(3): e2e_step()
        t0 = time.perf_counter()
        for _ in range(20): e2e_step()
        e2e = (time.perf_counter() - t0) / 20
        if rnd >= 1:
            res[st][0].append(np.mean([a.elapsed_time(b) for a, b in evs])); res[st][1].append(e2e * 1e3)
            res[st][2].add((r.best_index, r.best_score, b_idx.value, b_score.value))
for st in settings:
    d, h = np.array(res[st][0]), np.array(res[st][1])
    print(f"{name} {st:34s} device step ms median {np.median(d):.4f} min {d.min():.4f} | e2e ms median {np.median(h):.4f} min {h.min():.4f} | winner {sorted(res[st][2])}")
