#!/usr/bin/env python
"""bench.py -- hybrid-MSAC hypothesis stage throughput (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C2]

A "step" is one pass of the hypothesis stage (replayed minimal samples -> degeneracy test -> minimal
solver -> multi-class MSAC over every feature -> best-model selection) over one batch of hypotheses of
the named configuration.  Metric: hypothesis x feature evaluations per second = H (Ns + No) / t, counted
over ALL hypothesis slots as the reference's loop counts iterations (GCRANSAC.h:293,339), for both arms.

`value`   device-resident inputs, CUDA-event timed on the engine stream, L2 flushed between steps.
`e2e`     the same step through the C ABI with HOST buffers: features + samples copied host->device,
          winner record + masks copied back, inside the timed region.
`--impl reference` times the reference's own CPU implementation of the stage (oracle/_ref when it was
built in the container, else the oracle port) on the host cores, on a bounded sample of the workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import gcr_scenes as gs  # noqa: E402

METRIC = "hypothesis-feature evals/sec (hybrid MSAC)"
UNIT = "evals/s"
# algorithmic FP32 work per evaluation (SURVEY.md 8d / DESIGN.md): scale 11 flop + 1 SFU, orientation 30 + 1
FLOP_SCALE, FLOP_ORIENT = 11.0, 30.0
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12  # nominal non-tensor FP32 peak at clocks.max.sm


def workload(name: str, rank: int = 0):
    c = gs.CONFIGS[name]
    scene = gs.config_scene(name)
    n_hyp = c["n_hyp"]
    samples = gs.draw_samples(n_hyp, c["n_scale"], c["n_orient"], gs.SAMPLE_SEED + 1000 * rank)
    return scene, samples, n_hyp


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons sampled DURING the timed regions (device-timed steps and the e2e loop): NVML
    polled every ~2 ms (pynvml), falling back to nvidia-smi when NVML is unavailable."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.sm, self.mx, self.reasons, self.stop_flag, self.active = index, [], [], set(), False, False
        self.interval = float(os.environ.get("GCR_BENCH_SAMPLE_MS", "2")) * 1e-3
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.mx.append(float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)))
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
        try:
            get = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
            mask = int(get(self.handle))
            self.reasons |= {k for k, b in self.BITS.items() if mask & b}
        except Exception:
            pass

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout.strip()
        r = [x.strip() for x in out.split(",")]
        if len(r) >= 7 and r[0].replace(".", "").isdigit():
            self.sm.append(float(r[0]))
            self.mx.append(float(r[1]))
            self.reasons |= {k for k, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7])
                             if v.lower().startswith("active")}

    def run(self):
        while not self.stop_flag:
            if self.active:
                try:
                    self._sample_nvml() if self.nvml else self._sample_smi()
                except Exception:
                    pass
            time.sleep(self.interval if self.nvml else 0.05)

    def summary(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml" if self.nvml else "nvidia-smi"}


def cpu_reference_leg(name: str, budget_hyp: int, nthreads: int = 0):
    """Times the reference's CPU hypothesis stage on a bounded sample of the workload."""
    from oracle_lib import Oracle, have_reference
    kind = "reference" if have_reference() else "port"
    o = Oracle(kind)
    if nthreads <= 0:  # every host core this process may use (torchrun exports OMP_NUM_THREADS=1: do not inherit that)
        nthreads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    scene, samples, n_hyp = workload(name)
    smp = samples[:budget_hyp]
    nfeat = len(scene.scale) + len(scene.orient)
    o.replay(scene.scale, scene.orient, smp[:256], scene.thr, want_masks=False, nthreads=nthreads)  # warm
    t0 = time.perf_counter()
    r = o.replay(scene.scale, scene.orient, smp, scene.thr, want_masks=False, nthreads=nthreads)
    dt = time.perf_counter() - t0
    return {"value": len(smp) * nfeat / dt, "unit": UNIT, "cores": int(r["threads"]), "kind": kind,
            "sample": f"first {len(smp)} of {n_hyp} replayed hypotheses of {name} x {nfeat} features, "
                      f"{'unmodified reference headers on Eigen/OpenCV shims' if kind == 'reference' else 'C oracle port'}, "
                      f"OpenMP over hypotheses, {dt:.2f} s"}, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the reference is single-process; other ranks exit without work
    name = args.config
    c = gs.CONFIGS[name]
    nfeat = c["n_scale"] + c["n_orient"]
    # bounded sample per step: ~2.5e8 evals of CPU work at most
    budget = max(512, min(c["n_hyp"], int(2.5e8 / nfeat)))
    for _ in range(args.warmup):
        cpu_reference_leg(name, min(budget, 2048))
    vals, times = [], []
    for _ in range(args.steps):
        leg, dt = cpu_reference_leg(name, budget)
        vals.append(leg["value"])
        times.append(dt)
    value = float(np.mean(vals))
    leg["value"] = value
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(times)), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "impl": "reference",
            "config": {"workload": f"{name}: {c['n_scale']} scale + {c['n_orient']} orientation features, "
                                   f"{int(100 * c['inlier_ratio'])}% inliers, replayed samples, {budget} hypotheses per step (bounded sample)"},
            "cpu_baseline": leg,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def run_ours(args):
    import torch
    import torch.distributed as dist
    import gcr_capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    name = args.config
    scene, samples, n_hyp = workload(name, rank)
    ns, no = len(scene.scale), len(scene.orient)
    nfeat = ns + no
    first_hyp = rank * n_hyp  # weak scaling: every rank owns its own block of hypothesis slots

    eng = gcr_capi.Engine(local)
    stream = torch.cuda.Stream()  # a real (non-null) stream shared by torch, NCCL and the engine
    torch.cuda.set_stream(stream)
    eng.set_stream(stream.cuda_stream)
    d_scale = torch.from_numpy(scene.scale).cuda()
    d_orient = torch.from_numpy(scene.orient).cuda()
    d_samples = torch.from_numpy(samples).cuda()
    d_record = torch.zeros(8, dtype=torch.float64, device="cuda")  # one 64-byte gcr_record: the NCCL send buffer
    d_gather = torch.zeros(8 * world, dtype=torch.float64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    eng.set_features_device(d_scale.data_ptr(), ns, d_orient.data_ptr(), no)

    # N > 1: the path's one exchange (64-byte winner records) is fused into the engine's last kernel: peer-memory stores
    # over NVLink + acquire loads, no collective launch; --exchange nccl times the all_gather form instead
    fused = world > 1 and args.exchange == "p2p"
    if fused:
        import gcr_parallel
        gcr_parallel.attach_peers(eng, rank, world)
    run_flags = gcr_capi.RUN_EXCHANGE if fused else 0

    def step():
        eng.run(n_hyp, scene.thr, first_hyp=first_hyp, samples_ptr=d_samples.data_ptr(), record_ptr=d_record.data_ptr(),
                flags=run_flags, sync=False)
        if world > 1 and not fused:
            dist.all_gather_into_tensor(d_gather, d_record)

    for _ in range(max(args.warmup, 3)):
        flush.zero_()
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    torch.cuda.synchronize()
    sampler.active = rank == 0 or os.environ.get("GCR_BENCH_SAMPLE_ALL") is not None  # one poller per job: NVML calls can stall that GPU's launches
    if world > 1:
        # The ranks leave the host barrier up to milliseconds apart, and every step ends in an exchange that waits for the
        # slowest rank: a few untimed steps right here absorb that start-up skew, so that the K timed steps measure the
        # lock-step rate of the job (they are enqueued back to back behind these, with no host synchronisation between).
        for _ in range(4):
            flush.zero_()
            step()
    evs = []
    for _ in range(args.steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        step()
        e1.record(stream)
        evs.append((e0, e1))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.active = False
    if world > 1:  # every rank must hold the same global winner (checked outside the timed region)
        mine = d_record.clone()
        allr = torch.empty(8 * world, dtype=torch.float64, device="cuda")
        dist.all_gather_into_tensor(allr, mine)
        recs = np.frombuffer(allr.cpu().numpy().tobytes(), dtype=gcr_capi.RECORD_DTYPE)
        if fused:
            assert all(r["index"] == recs[0]["index"] and r["score"] == recs[0]["score"] for r in recs), recs
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
    t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = world * n_hyp * nfeat / (ms_per_step * 1e-3)

    # ---- dominant kernel (FP32 MSAC scoring) timed alone, same launches, synchronous result read-back
    res = eng.run(n_hyp, scene.thr, first_hyp=first_hyp, samples_ptr=d_samples.data_ptr(), sync=True)
    k_ms = []
    for _ in range(20 if n_hyp * nfeat < 1e10 else 5):
        flush.zero_()
        r = eng.run(n_hyp, scene.thr, first_hyp=first_hyp, samples_ptr=d_samples.data_ptr(), sync=True, flags=gcr_capi.RUN_TIME_KERNELS)
        k_ms.append(r.score_kernel_ms)
    k_ms = float(np.mean(k_ms))  # mean, not median: CUDA event stamps tick every ~2 us here, which a 45 us kernel only averages out
    kernel_name = "k_score_resident" if r.score_kernel == 1 else "k_score_fast"
    n_fast = res.n_solved - res.n_exact_hyp
    flops = n_fast * (ns * FLOP_SCALE + no * FLOP_ORIENT)
    achieved = flops / (k_ms * 1e-3) / 1e12
    alg_bytes = (res.n_solved * (32 + 32 + 4) + (ns + no) * 12)  # hypothesis parameters in/out + one pass over the features
    roofline = {"bound": "fp32", "achieved": achieved, "peak": FP32_PEAK_TFLOPS, "unit": "TFLOP/s",
                "frac": achieved / FP32_PEAK_TFLOPS, "traffic": None,
                "note": "issue-bound FP32/SFU kernel (no dense contraction, HBM traffic negligible): peak = 148 SM x 128 lanes x 2 x "
                        "1.965 GHz nominal (not in MEASURED_PEAKS.json, which holds HBM and bf16-tensor peaks only)",
                "kernel": kernel_name, "kernel_ms": k_ms, "kernel_evals": int(n_fast) * nfeat,
                "kernel_evals_per_s": n_fast * nfeat / (k_ms * 1e-3),
                "hbm": {"algorithmic_bytes": int(alg_bytes), "achieved_gbs": alg_bytes / (k_ms * 1e-3) / 1e9}}
    try:  # DRAM bytes of this kernel from the committed `ncu --set full` capture (tools/summarize_profiles.py)
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[args.config][kernel_name]
        roofline["traffic"] = tr["dram_bytes_per_launch"]
        roofline["traffic_source"] = tr["source"]
    except Exception:
        pass
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        roofline["hbm"]["peak_gbs_measured"] = peaks.get("hbm_gbs")
    except Exception:
        pass

    # ---- end to end through the C ABI with HOST buffers (pinned), H2D + D2H inside the timed region
    h_scale = torch.from_numpy(scene.scale).pin_memory()
    h_orient = torch.from_numpy(scene.orient).pin_memory()
    h_samples = torch.from_numpy(samples).pin_memory()
    e2e_eng = gcr_capi.Engine(local)

    # the C ABI called directly with preallocated argument / result buffers (what a C++ caller does); unpacking the packed
    # masks into Python objects is not part of the path
    C = gcr_capi.C
    thr_c = (C.c_double * 2)(*scene.thr)
    out = gcr_capi.GcrResult()
    b_model, b_n_inl = np.zeros(4), np.zeros(2, np.uint32)
    b_sm, b_om = np.zeros((ns + 31) // 32, np.uint32), np.zeros((no + 31) // 32, np.uint32)
    b_score, b_idx = C.c_double(), C.c_int64()
    lib, hh = e2e_eng.lib, e2e_eng.h
    p_scale, p_orient, p_samples = C.c_void_p(h_scale.data_ptr()), C.c_void_p(h_orient.data_ptr()), C.c_void_p(h_samples.data_ptr())
    p_model, p_n_inl, p_sm, p_om = (C.c_void_p(a.ctypes.data) for a in (b_model, b_n_inl, b_sm, b_om))

    def e2e_step():
        rc = lib.gcr_set_features(hh, p_scale, ns, p_orient, no, 0)
        rc |= lib.gcr_run(hh, n_hyp, 0, first_hyp, p_samples, thr_c, 0, None, C.byref(out))
        rc |= lib.gcr_best(hh, p_model, C.byref(b_score), p_n_inl, p_sm, p_om, C.byref(b_idx))
        assert rc == 0, lib.gcr_last_error(hh)
        return {"index": b_idx.value, "score": b_score.value}

    for _ in range(3):
        best = e2e_step()
    if world > 1:
        dist.barrier()
    sampler.active = rank == 0 or os.environ.get("GCR_BENCH_SAMPLE_ALL") is not None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        best = e2e_step()
    e2e_s = (time.perf_counter() - t0) / args.steps
    sampler.active = False
    sampler.stop_flag = True
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    h2d = scene.scale.nbytes + scene.orient.nbytes + samples.nbytes
    d2h = 128 + 4 * ((ns + 31) // 32 + (no + 31) // 32)  # winner record + run counters + packed masks, one copy
    e2e = {"value": world * n_hyp * nfeat / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
           "ms_per_step": 1e3 * e2e_s, "api": "gcr_set_features + gcr_run + gcr_best (C ABI, pinned host buffers; the kernels read the pinned inputs in place "
                                           "over PCIe and write the result block into pinned staging, so the bytes cross the bus inside the timed region without separate copies)"}

    if rank == 0:
        c = gs.CONFIGS[name]
        cpu_leg, _ = cpu_reference_leg(name, max(512, min(n_hyp, int(2.5e8 / nfeat))))
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32 (+f64 solver, guard-band and winner re-scoring)", "data": "synthetic",
                "config": {"workload": f"{name}: {c['n_scale']} scale + {c['n_orient']} orientation features, "
                                       f"{int(100 * c['inlier_ratio'])}% inliers, {n_hyp} replayed hypotheses per GPU per step",
                           "l2": "flushed between timed steps (256 MiB write)",
                           "exchange": ("fused peer-memory stores in k_tail (NVLink P2P, 64 B per rank pair)" if fused else
                                        ("nccl all_gather of 64 B records" if world > 1 else "none (1 GPU)")), "hypotheses_scored_per_step": int(res.n_solved),
                           "guard_band_pairs": int(res.n_rechecked), "exact_hypotheses": int(res.n_exact_hyp),
                           "best_index": int(best["index"]), "best_score": float(best["score"])},
                "clocks": sampler.summary(), "e2e": e2e, "gpu_launches": 3 * args.steps, "roofline": roofline,
                "cpu_baseline": cpu_leg}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_batched(args):
    """C4: a batch of independent images per GPU (BASELINE.json configs[3]): 256 images x (4000 + 1000) features, 100k
    device-sampled hypotheses each, images/world per rank, no communication.  A step = the rank's whole image slice;
    images are spread over `--engines` engine handles, each on its own stream, so that the low-occupancy kernels of one
    image (sample+solve, tail) overlap the scoring kernel of another."""
    import torch
    import torch.distributed as dist
    import gcr_capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    c = gs.CONFIGS[args.config]
    n_img = c["n_images"] // 8 if args.images is None else args.images  # per GPU: the 256-image batch is cut for 8 GPUs
    n_hyp, nfeat = c["n_hyp"], c["n_scale"] + c["n_orient"]
    scenes = [gs.config_scene(args.config, image=rank * n_img + i) for i in range(n_img)]
    d_scale = [torch.from_numpy(sc.scale).cuda() for sc in scenes]
    d_orient = [torch.from_numpy(sc.orient).cuda() for sc in scenes]
    K = max(1, min(args.engines, n_img))
    main = torch.cuda.Stream()
    streams = [torch.cuda.Stream() for _ in range(K)]
    engines = [gcr_capi.Engine(local) for _ in range(K)]
    for e, st in zip(engines, streams):
        e.set_stream(st.cuda_stream)
    records = torch.zeros(n_img, 8, dtype=torch.float64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def step():
        start = torch.cuda.Event()
        start.record(main)
        for st in streams:
            st.wait_event(start)
        for i in range(n_img):
            e = engines[i % K]
            e.set_features_device(d_scale[i].data_ptr(), c["n_scale"], d_orient[i].data_ptr(), c["n_orient"])
            e.run(n_hyp, scenes[i].thr, seed=1000 + rank * n_img + i, record_ptr=records[i].data_ptr(), sync=False)
        for st in streams:
            done = torch.cuda.Event()
            done.record(st)
            main.wait_event(done)

    with torch.cuda.stream(main):
        for _ in range(max(args.warmup, 3)):
            flush.zero_()
            step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        sampler = ClockSampler(local)
        sampler.start()
        sampler.active = True
        evs = []
        for _ in range(args.steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(main)
            step()
            e1.record(main)
            evs.append((e0, e1))
        torch.cuda.synchronize()
        sampler.active = False
        sampler.stop_flag = True
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
    t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.barrier()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / args.steps
    value = world * n_img * n_hyp * nfeat / (ms_per_step * 1e-3)
    recs = np.frombuffer(records.cpu().numpy().tobytes(), dtype=gcr_capi.RECORD_DTYPE)
    err = [float(np.hypot(*(r["model"][:2] - sc.model[:2])) / np.hypot(*sc.model[:2])) for r, sc in zip(recs, scenes)]
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32 (+f64 solver, guard-band and winner re-scoring)", "data": "synthetic",
                "config": {"workload": f"{args.config}: {n_img} images per GPU x ({c['n_scale']} scale + {c['n_orient']} orientation features), "
                                       f"{n_hyp} device-sampled (Philox) hypotheses per image, {K} engines/streams in flight",
                           "l2": "flushed between timed steps (256 MiB write)", "images_with_model": int((recs["index"] >= 0).sum()),
                           "median_vanishing_line_error": float(np.median(err))},
                "clocks": sampler.summary(), "e2e": None, "gpu_launches": 4 * n_img * args.steps, "roofline": None, "cpu_baseline": None,
                "images_per_s": world * n_img / (ms_per_step * 1e-3)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=sorted(gs.CONFIGS))
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"], help="N>1: fused peer-memory exchange or NCCL all_gather")
    ap.add_argument("--engines", type=int, default=4, help="C4: engine handles / streams in flight per GPU")
    ap.add_argument("--images", type=int, default=None, help="C4: images per GPU (default 256 / 8)")
    args = ap.parse_args()
    if args.impl == "ours" and "n_images" in gs.CONFIGS[args.config]:
        run_batched(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
