"""Timeline of the scoring kernel (GPU box; needs gpurun_variants/lib_trace.so = the engine built with -DGCR_TRACE):
python tools/trace_score.py C2 [GCR_RESIDENT=0|GCR_TPW=4 ...]   FLUSH=1 flushes L2 before the traced run."""
import ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")]
import torch, gcr_capi, gcr_scenes as gs
name = sys.argv[1]
for kv in sys.argv[2:]:
    k, v = kv.split("="); os.environ[k] = v
gcr_capi.LIB_PATH = os.path.join(ROOT, "gpurun_variants", "lib_trace.so")
c = gs.CONFIGS[name]; sc = gs.config_scene(name)
smp = torch.from_numpy(gs.draw_samples(c["n_hyp"], c["n_scale"], c["n_orient"], gs.SAMPLE_SEED)).cuda()
e = gcr_capi.Engine(0); e.set_features(sc.scale, sc.orient)
lib = gcr_capi.load_library()
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda") if os.environ.get("FLUSH") else None
ROWS, COLS = 8192, 64
buf = np.zeros((ROWS, COLS), dtype=np.uint64)
for rep in range(4):
    if flush is not None: flush.zero_(); torch.cuda.synchronize()
    r = e.run(c["n_hyp"], sc.thr, samples_ptr=smp.data_ptr(), flags=gcr_capi.RUN_TIME_KERNELS)
    lib.gcr_debug_read_trace(buf.ctypes.data_as(C.c_void_p), C.c_size_t(buf.nbytes))
print(f"{name} {sys.argv[2:]} kernel {r.score_kernel_ms*1e3:.1f} us, stage {r.device_ms*1e3:.1f} us")
if os.environ.get("TRACE_DUMP"): np.save(os.environ["TRACE_DUMP"], buf[:4736])
rows = buf[buf[:, 1] > 0]
t0 = rows[:, 1].min()
T = (rows[:, 1:].astype(np.int64) - np.int64(t0)) / 1e3  # us
T[rows[:, 1:] == 0] = np.nan
entry, plan = T[:, 0], T[:, 1]
n_st = np.sum(~np.isnan(T), axis=1)
n_items = (n_st - 2) // 2
last = np.array([T[i, n_st[i] - 1] for i in range(len(T))])
print(f"rows {len(rows)}  entry us: min {entry.min():.2f} median {np.median(entry):.2f} max {entry.max():.2f}")
print(f"plan ready us: min {plan.min():.2f} median {np.median(plan):.2f} max {plan.max():.2f}")
print(f"items per row: min {n_items.min()} median {np.median(n_items):.1f} max {n_items.max()}  total {n_items.sum()}")
print(f"last stamp us: min {last.min():.2f} p10 {np.percentile(last,10):.2f} median {np.median(last):.2f} p90 {np.percentile(last,90):.2f} max {last.max():.2f}")
# item durations and claim gaps: stamps after plan come in pairs (begin, end)
dur, gap, first_begin = [], [], []
for i in range(len(T)):
    st = T[i, 2:n_st[i]]
    b, en = st[0::2], st[1::2]
    m = min(len(b), len(en))
    dur += list(en[:m] - b[:m])
    if m: first_begin.append(b[0])
    gap += list(b[1:m] - en[:m - 1])
dur, gap = np.array(dur), np.array(gap)
print(f"first item begins us: median {np.median(first_begin):.2f} max {np.max(first_begin):.2f}")
print(f"item duration us: min {dur.min():.2f} p10 {np.percentile(dur,10):.2f} median {np.median(dur):.2f} p90 {np.percentile(dur,90):.2f} max {dur.max():.2f}  sum/rows {dur.sum()/len(T):.2f}")
if len(gap): print(f"claim gap us: median {np.median(gap):.2f} p90 {np.percentile(gap,90):.2f} max {gap.max():.2f}")
# busy rows over time (how many rows are inside an item at time t)
grid = np.arange(0, np.nanmax(T) + 1, 2.0)
busy = []
for t in grid:
    cnt = 0
    for i in range(len(T)):
        st = T[i, 2:n_st[i]]
        b, en = st[0::2], st[1::2]
        m = min(len(b), len(en))
        cnt += int(np.any((b[:m] <= t) & (en[:m] > t)))
    busy.append(cnt)
print("rows busy at t (us):", " ".join(f"{int(t)}:{b}" for t, b in zip(grid, busy)))
