"""Host-side timing of the three C-ABI calls of one end-to-end step (GPU box): python tools/diag_e2e.py [C2]"""
import ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")]
import numpy as np, torch, gcr_capi, gcr_scenes as gs
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
c = gs.CONFIGS[name]; sc = gs.config_scene(name)
ns, no, H = c["n_scale"], c["n_orient"], c["n_hyp"]
smp = gs.draw_samples(H, ns, no, gs.SAMPLE_SEED)
hs, ho, hp = torch.from_numpy(sc.scale).pin_memory(), torch.from_numpy(sc.orient).pin_memory(), torch.from_numpy(smp).pin_memory()
e = gcr_capi.Engine(0); lib, h = e.lib, e.h
thr = (C.c_double * 2)(*sc.thr); out = gcr_capi.GcrResult()
m, n = np.zeros(4), np.zeros(2, np.uint32); sm, om = np.zeros((ns + 31) // 32, np.uint32), np.zeros((no + 31) // 32, np.uint32)
sco, idx = C.c_double(), C.c_int64()
ps, po, pp = C.c_void_p(hs.data_ptr()), C.c_void_p(ho.data_ptr()), C.c_void_p(hp.data_ptr())
T = []
for it in range(60):
    t0 = time.perf_counter(); lib.gcr_set_features(h, ps, ns, po, no, 0)
    t1 = time.perf_counter(); lib.gcr_run(h, H, 0, 0, pp, thr, 0, None, C.byref(out))
    t2 = time.perf_counter(); lib.gcr_best(h, m.ctypes.data, C.byref(sco), n.ctypes.data, sm.ctypes.data, om.ctypes.data, C.byref(idx))
    t3 = time.perf_counter()
    if it >= 10: T.append((t1 - t0, t2 - t1, t3 - t2, out.device_ms * 1e-3, out.stage_phase_us[1] * 1e-6))
T = np.median(np.array(T), axis=0) * 1e6
print(f"{name}: set_features {T[0]:.1f} us | run (incl. sync) {T[1]:.1f} us | best {T[2]:.1f} us | total {T[:3].sum():.1f} us | "
      f"device stage (events) {T[3]:.1f} us, first kernel start -> end of tail {T[4]:.1f} us")
