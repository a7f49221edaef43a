#!/usr/bin/env python
"""bench.py -- hybrid-MSAC hypothesis stage throughput (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C3|C2|C1|C4|C5]

A "step" is one pass of the hypothesis stage (minimal samples -> degeneracy test -> minimal solver -> multi-class MSAC
over every feature -> best-model selection, GCRANSAC.h:286-485) over the hypotheses of the named configuration.
Metric: hypothesis x feature evaluations per second = H (Ns + No) / t, counted over ALL hypothesis slots as the
reference's loop counts iterations (GCRANSAC.h:293,339), for both arms; `value_scored` counts only the hypotheses
that are actually scored against the features.

Default workload = the north-star configuration C3 (BASELINE.json configs[2]): 50 000 + 20 000 features, 1 000 000
device-sampled (Philox) hypotheses IN TOTAL, sharded by global index over the N ranks (strong scaling); the winner
exchange is fused into the engine's last kernel (peer memory over NVLink).  It fits one GPU, so it is also the N = 1
line.  C2 (the replay-parity configuration) is reported in the same line as `secondary` (N = 1).

`value`   device-resident inputs, CUDA-event timed on the engine stream, L2 flushed between steps, max over ranks.
`e2e`     the same step through the C ABI with HOST buffers (gcr_set_features + gcr_run + gcr_best): features (+ replayed
          samples) cross PCIe inside the timed region, the winner record + packed masks come back.
`--impl reference` times the reference's own CPU implementation of the stage (oracle/_ref: the unmodified headers
compiled in the container; else the C port) on the host cores, on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import math
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
REPLAY_CONFIGS = ("C1", "C2")  # minimal samples replayed from an exported [H,4] index array; the others draw on the device (Philox)
PHILOX_SEED = 2025
CPU_BUDGET_EVALS = 2.5e8  # bounded CPU sample per reference step
DTYPE = "f32 (+f64 solver, candidate re-scoring and winner)"


def workload_string(name: str) -> str:
    """Identical in both arms (the driver compares the strings)."""
    c = gs.CONFIGS[name]
    src = "replayed samples" if name in REPLAY_CONFIGS else "Philox4x32-10 samples keyed by the global hypothesis index"
    img = f"{c['n_images']} images x " if "n_images" in c else ""
    sweep = ", 20-setting sweep (5-50% inliers x threshold levels 1/2/4/8) + feature-count ladder" if name == "C5" else ""
    return (f"{name}: {img}{c['n_scale']} scale + {c['n_orient']} orientation features, {int(100 * c['inlier_ratio'])}% inliers, "
            f"{c['n_hyp']} hypotheses in total{' per image' if img else ''} per step, {src}{sweep}")


def host_samples(name: str, count: int, first: int = 0, seed: int = PHILOX_SEED, image: int = 0) -> np.ndarray:
    """The minimal samples of global slots [first, first + count) as the GPU arm sees them."""
    c = gs.CONFIGS[name]
    if name in REPLAY_CONFIGS:
        return gs.draw_samples(c["n_hyp"], c["n_scale"], c["n_orient"], gs.SAMPLE_SEED)[first:first + count]
    from oracle_lib import Oracle  # Philox twin of the device sampler (CPU baseline legs only)
    return Oracle("port").sample_philox(seed, first, count, c["n_scale"], c["n_orient"])


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons sampled DURING the timed regions: NVML polled every ~2 ms (pynvml), falling back to
    nvidia-smi when NVML is unavailable."""
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


def host_threads() -> int:
    # every host core this process may use (torchrun exports OMP_NUM_THREADS=1: do not inherit that)
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def cpu_reference_leg(name: str, budget_hyp: int = 0, nthreads: int = 0, scene=None, seed: int = PHILOX_SEED):
    """Times the reference's CPU hypothesis stage (isValidSample -> estimateModel -> getScore -> isValidModel per sample,
    OpenMP over hypotheses) on a bounded sample of the workload: the first `budget_hyp` hypothesis slots."""
    from oracle_lib import Oracle, have_reference
    kind = "reference" if have_reference() else "port"
    o = Oracle(kind)
    nthreads = nthreads if nthreads > 0 else host_threads()
    c = gs.CONFIGS[name]
    scene = scene if scene is not None else gs.config_scene(name)
    nfeat = len(scene.scale) + len(scene.orient)
    if budget_hyp <= 0:
        budget_hyp = max(512, min(c["n_hyp"], int(CPU_BUDGET_EVALS / nfeat)))
    smp = host_samples(name, budget_hyp, seed=seed)
    o.replay(scene.scale, scene.orient, smp[:256], scene.thr, want_masks=False, nthreads=nthreads)  # warm
    t0 = time.perf_counter()
    r = o.replay(scene.scale, scene.orient, smp, scene.thr, want_masks=False, nthreads=nthreads)
    dt = time.perf_counter() - t0
    return {"value": len(smp) * nfeat / dt, "unit": UNIT, "cores": int(r["threads"]), "kind": kind,
            "sample": f"first {len(smp)} of {c['n_hyp']} hypothesis slots of {name} x {nfeat} features, "
                      f"{'unmodified reference headers on Eigen/OpenCV shims' if kind == 'reference' else 'C oracle port'}, "
                      f"OpenMP over hypotheses, {dt:.2f} s"}, dt


def reference_b1():
    """the UNMODIFIED reference estimator end to end (findRectifyingHomographySIFT_, gcransac_python.cpp:256-406:
    its own random_device sampler, graph-cut local optimisation and refit) on ONE core (the reference is single-threaded) at C1,
    min = max = 5000 iterations; best of 3 after a warm-up."""
    from oracle_lib import Oracle, have_reference
    if not have_reference():
        return None
    o = Oracle("reference")
    c = gs.CONFIGS["C1"]
    sc = gs.config_scene("C1")
    nfeat = c["n_scale"] + c["n_orient"]
    o.find_sift(sc.scale, sc.orient, sc.thr[0], sc.thr[1], 200, 200)
    best = None
    for _ in range(3):
        r = o.find_sift(sc.scale, sc.orient, sc.thr[0], sc.thr[1], c["n_hyp"], c["n_hyp"])
        if best is None or r["seconds"] < best["seconds"]:
            best = r
    return {"what": "unmodified findRectifyingHomographySIFT_ (own sampler, LO, refit), C1, min = max = 5000 iterations, 1 core",
            "seconds": best["seconds"], "iterations_per_s": c["n_hyp"] / best["seconds"],
            "evals_per_s": c["n_hyp"] * nfeat / best["seconds"], "inliers_found": int(best["n_inliers"])}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the reference is single-process; other ranks exit without work
    name = args.config
    for _ in range(args.warmup):
        cpu_reference_leg(name, 1024)
    vals, times = [], []
    for _ in range(args.steps):
        leg, dt = cpu_reference_leg(name)
        vals.append(leg["value"])
        times.append(dt)
    value = float(np.mean(vals))
    leg["value"] = value
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(times)), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "impl": "reference",
            "config": {"workload": workload_string(name)},
            "cpu_baseline": leg,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    try:
        line["b1_full_estimator_c1"] = reference_b1()
    except Exception as e:  # never lose the line over the extra leg
        line["b1_full_estimator_c1"] = {"error": repr(e)}
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------------------------------
def _load_json(path):
    try:
        return json.load(open(os.path.join(ROOT, path)))
    except Exception:
        return None


def roofline_block(name, kernel_name, k_ms, n_scored, n_exact, ns, no):
    n_fast = n_scored - n_exact
    flops = n_fast * (ns * FLOP_SCALE + no * FLOP_ORIENT)
    achieved = flops / (k_ms * 1e-3) / 1e12
    alg_bytes = n_scored * (32 + 32 + 4) + (ns + no) * 12  # hypothesis parameters in / sums out + one pass over the feature planes
    rf = {"bound": "fp32", "achieved": achieved, "peak": FP32_PEAK_TFLOPS, "unit": "TFLOP/s", "frac": achieved / FP32_PEAK_TFLOPS,
          "traffic": None,
          "note": "issue-bound FP32/SFU kernel (no dense contraction, HBM traffic negligible): peak = 148 SM x 128 lanes x 2 x 1.965 GHz "
                  "nominal (MEASURED_PEAKS.json holds HBM and bf16-tensor peaks only); peak_measured = FFMA microbenchmark of "
                  "tools/ubench.cu on a B200 of this pool (profiles/ubench_b200.json); flops = SURVEY 8d's algorithmic 11 (scale) / 30 "
                  "(orientation) per evaluation x evaluations performed",
          "kernel": kernel_name, "kernel_ms": k_ms, "kernel_evals": int(n_fast) * (ns + no),
          "kernel_evals_per_s": n_fast * (ns + no) / (k_ms * 1e-3),
          "sfu_frac": n_fast * (ns + no) / (k_ms * 1e-3) / (148 * 16 * 1.965e9),
          "hbm": {"algorithmic_bytes": int(alg_bytes), "achieved_gbs": alg_bytes / (k_ms * 1e-3) / 1e9}}
    ub = _load_json("profiles/ubench_b200.json")
    if ub:
        rf["peak_measured"] = ub.get("fp32_ffma_tflops_sustained")
        rf["peak_measured_burst"] = ub.get("fp32_ffma_tflops_burst")
        if rf["peak_measured"]:
            rf["frac_of_measured"] = achieved / rf["peak_measured"]
    tr = _load_json("profiles/traffic.json")
    try:
        t = tr[name][kernel_name]
        rf["traffic"] = t["dram_bytes_per_launch"]
        rf["traffic_source"] = t["source"]
    except Exception:
        pass
    peaks = _load_json("MEASURED_PEAKS.json")
    if peaks:
        rf["hbm"]["peak_gbs_measured"] = peaks.get("hbm_gbs")
    return rf


class StageBench:
    """One configuration of the hypothesis stage on this rank's GPU: device-timed steps, the scoring kernel alone, and the
    end-to-end step through the C ABI with host buffers."""

    def __init__(self, name, torch, gcr_capi, local, rank, world, stream, flush, fused, scene=None, seed=PHILOX_SEED, n_hyp=None):
        import gcr_parallel
        self.t, self.capi, self.name, self.rank, self.world, self.local, self.stream, self.flush = torch, gcr_capi, name, rank, world, local, stream, flush
        c = gs.CONFIGS[name]
        self.scene = scene if scene is not None else gs.config_scene(name)
        self.ns, self.no = len(self.scene.scale), len(self.scene.orient)
        self.nfeat = self.ns + self.no
        self.total = int(n_hyp if n_hyp is not None else c["n_hyp"])
        self.first, self.count = gcr_parallel.shard_range(self.total, rank, world)
        self.replay = name in REPLAY_CONFIGS
        self.seed = seed
        self.samples = host_samples(name, self.count, self.first) if self.replay else None
        self.eng = gcr_capi.Engine(local)
        self.eng.set_stream(stream.cuda_stream)
        self.d_scale = torch.from_numpy(self.scene.scale).cuda()
        self.d_orient = torch.from_numpy(self.scene.orient).cuda()
        self.d_samples = torch.from_numpy(self.samples).cuda() if self.replay else None
        self.d_record = torch.zeros(8, dtype=torch.float64, device="cuda")
        self.eng.set_features_device(self.d_scale.data_ptr(), self.ns, self.d_orient.data_ptr(), self.no)
        self.fused = fused
        if fused:
            gcr_parallel.attach_peers(self.eng, rank, world)
        self.flags = gcr_capi.RUN_EXCHANGE if fused else 0

    def run(self, sync, flags=None, record=True):
        return self.eng.run(self.count, self.scene.thr, seed=self.seed, first_hyp=self.first,
                            samples_ptr=self.d_samples.data_ptr() if self.replay else None,
                            record_ptr=self.d_record.data_ptr() if record else None,
                            flags=self.flags if flags is None else flags, sync=sync)

    def device_steps(self, steps, warmup, extra=None):
        """K steps, each bracketed by CUDA events on the engine stream, L2 flushed before each; returns the summed ms."""
        torch = self.t
        for _ in range(warmup):
            self.flush.zero_()
            self.run(False)
            if extra:
                extra()
        evs = []
        for _ in range(steps):
            self.flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(self.stream)
            self.run(False)
            if extra:
                extra()
            e1.record(self.stream)
            evs.append((e0, e1))
        return evs

    def kernel_alone(self, reps):
        """The dominant (scoring) kernel bracketed by events inside gcr_run (GCR_RUN_TIME_KERNELS), synchronous runs, no exchange."""
        k_ms, r = [], None
        for _ in range(reps):
            self.flush.zero_()
            r = self.run(True, flags=self.capi.RUN_TIME_KERNELS, record=False)
            k_ms.append(r.score_kernel_ms)
        # mean, not median: CUDA event stamps tick every ~2 us here, which a 45 us kernel only averages out
        return float(np.mean(k_ms)), r

    def make_e2e(self):
        """gcr_set_features + gcr_run + gcr_best with preallocated pinned host buffers (what a C++ caller does)."""
        torch, capi = self.t, self.capi
        C = capi.C
        self.h_scale = torch.from_numpy(self.scene.scale).pin_memory()
        self.h_orient = torch.from_numpy(self.scene.orient).pin_memory()
        # replayed minimal samples travel as uint16[H,4] where both classes allow it (GCR_RUN_SAMPLES_U16: half the PCIe bytes of int32)
        self.narrow = self.replay and max(self.ns, self.no) <= 65535
        self.h_samples = (torch.from_numpy(self.samples.astype(np.uint16) if self.narrow else self.samples).pin_memory()) if self.replay else None
        e = self.e2e_eng = capi.Engine(self.local)
        if self.fused:
            import gcr_parallel
            gcr_parallel.attach_peers(e, self.rank, self.world)
        thr_c = (C.c_double * 2)(*self.scene.thr)
        out = capi.GcrResult()
        b_model, b_n_inl = np.zeros(4), np.zeros(2, np.uint32)
        b_sm, b_om = np.zeros((self.ns + 31) // 32, np.uint32), np.zeros((self.no + 31) // 32, np.uint32)
        b_score, b_idx = C.c_double(), C.c_int64()
        lib, hh = e.lib, e.h
        p_scale, p_orient = C.c_void_p(self.h_scale.data_ptr()), C.c_void_p(self.h_orient.data_ptr())
        p_samples = C.c_void_p(self.h_samples.data_ptr()) if self.replay else None
        p_model, p_n_inl, p_sm, p_om = (C.c_void_p(a.ctypes.data) for a in (b_model, b_n_inl, b_sm, b_om))
        self._keep = (b_model, b_n_inl, b_sm, b_om)
        ns, no, count, seed, first = self.ns, self.no, self.count, self.seed, self.first
        flags = self.flags | (capi.RUN_SAMPLES_U16 if self.narrow else 0)

        def step():
            rc = lib.gcr_set_features(hh, p_scale, ns, p_orient, no, 0)
            rc |= lib.gcr_run(hh, count, seed, first, p_samples, thr_c, flags, None, C.byref(out))
            rc |= lib.gcr_best(hh, p_model, C.byref(b_score), p_n_inl, p_sm, p_om, C.byref(b_idx))
            assert rc == 0, lib.gcr_last_error(hh)
            return {"index": b_idx.value, "score": b_score.value}
        self.h2d = self.scene.scale.nbytes + self.scene.orient.nbytes + (self.h_samples.numel() * self.h_samples.element_size() if self.replay else 0)
        self.d2h = 128 + 4 * ((ns + 31) // 32 + (no + 31) // 32)  # winner record + run counters + packed masks
        return step


E2E_API = ("gcr_set_features + gcr_run + gcr_best (C ABI, pinned host buffers; the kernels read the pinned inputs in place over PCIe and write "
           "the result block into pinned staging, so the bytes cross the bus inside the timed region without separate copies)")


def winner_check(sb: StageBench, torch, dist, gcr_capi):
    """Outside the timed region: the winner every rank holds after a sharded, exchanging run must equal the UNSHARDED run's
    (index, score, model, counts, packed masks); at N = 1 the unsharded run is compared with two half-range runs combined by
    the selection rule (shard invariance) and with the FP64 scoring of the winning model."""
    import gcr_parallel
    W = (sb.ns + 31) // 32 + (sb.no + 31) // 32
    res = sb.run(True, record=False)
    mine = sb.eng.best()
    pack = torch.zeros(8 + W, dtype=torch.float64, device="cuda")
    ok = True
    if sb.world > 1:
        if sb.rank == 0:
            whole = gcr_capi.Engine(sb.local)
            whole.set_features(sb.scene.scale, sb.scene.orient)
            r = whole.run(sb.total, sb.scene.thr, seed=sb.seed, first_hyp=0, samples=host_samples(sb.name, sb.total) if sb.replay else None)
            b = whole.best()
            vals = [float(r.best_index), b["score"], *b["model"], float(b["n_inl"][0]), float(b["n_inl"][1]),
                    *np.concatenate([b["scale_words"], b["orient_words"]]).astype(np.float64)]
            pack.copy_(torch.tensor(vals, dtype=torch.float64))
            whole.close()
        dist.broadcast(pack, 0)
        ref = pack.cpu().numpy()
        words = np.concatenate([mine["scale_words"], mine["orient_words"]]).astype(np.float64)
        ok = (float(res.best_index) == ref[0] and mine["score"] == ref[1] and np.array_equal(mine["model"], ref[2:6])
              and float(mine["n_inl"][0]) == ref[6] and float(mine["n_inl"][1]) == ref[7] and np.array_equal(words, ref[8:]))
        flag = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = bool(flag.item())
    else:
        recs = np.zeros(2, dtype=gcr_capi.RECORD_DTYPE)
        half = gcr_capi.Engine(sb.local)
        half.set_features(sb.scene.scale, sb.scene.orient)
        for k in range(2):
            first, count = gcr_parallel.shard_range(sb.total, k, 2)
            r = half.run(count, sb.scene.thr, seed=sb.seed, first_hyp=first, samples=sb.samples[first:first + count] if sb.replay else None)
            recs[k]["index"], recs[k]["score"] = r.best_index, r.best_score
            recs[k]["model"] = np.array(r.best_model[:])
        w = gcr_parallel.select_record(recs)
        ok = w >= 0 and int(recs[w]["index"]) == res.best_index and float(recs[w]["score"]) == res.best_score and \
            np.array_equal(recs[w]["model"], mine["model"])
        score, n_inl, sm, om = half.score_model(mine["model"], sb.scene.thr)  # FP64 reference-order scoring of the winning model
        ok = ok and score == mine["score"] and np.array_equal(n_inl, mine["n_inl"]) and np.array_equal(sm, mine["scale_mask"]) and \
            np.array_equal(om, mine["orient_mask"])
        half.close()
    return ok, res, mine


def run_ours(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    name = args.config
    # the CPU baseline leg runs FIRST (N = 1, rank 0), before torch, NVML and the poller thread exist: with them alive it read 40 % low
    cpu_leg = cpu_leg_c2 = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu_leg, _ = cpu_reference_leg(name)

    import torch
    import torch.distributed as dist
    import gcr_capi

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.Stream()  # a real (non-null) stream shared by torch, NCCL and the engine
    torch.cuda.set_stream(stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    # N > 1: the path's one exchange (64-byte winner records) is fused into the engine's last kernel: peer-memory stores
    # over NVLink + acquire loads, no collective launch; --exchange nccl times the all_gather form instead
    fused = world > 1 and args.exchange == "p2p"
    sb = StageBench(name, torch, gcr_capi, local, rank, world, stream, flush, fused, n_hyp=args.hyp)
    d_gather = torch.zeros(8 * world, dtype=torch.float64, device="cuda")
    extra = (lambda: dist.all_gather_into_tensor(d_gather, sb.d_record)) if (world > 1 and not fused) else None

    warm = max(args.warmup, 3)
    sb.device_steps(0, warm, extra)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    torch.cuda.synchronize()
    sampler.active = rank == 0 or os.environ.get("GCR_BENCH_SAMPLE_ALL") is not None  # one poller per job: NVML calls can stall that GPU's launches
    # N > 1: the ranks leave the host barrier up to milliseconds apart, and every step ends in an exchange that waits for the
    # slowest rank: a few untimed steps right here absorb that start-up skew, so that the K timed steps measure the lock-step
    # rate of the job (they are enqueued back to back behind these, with no host synchronisation between).
    evs = sb.device_steps(args.steps, 4 if world > 1 else 0, extra)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.active = False
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
    t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / args.steps
    value = sb.total * sb.nfeat / (ms_per_step * 1e-3)

    # ---- global winner vs the unsharded run (outside the timed region)
    ok, res, best = winner_check(sb, torch, dist, gcr_capi)
    scored = torch.tensor([float(res.n_scored), float(res.n_solved)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(scored)
    n_scored_total, n_solved_total = int(scored[0].item()), int(scored[1].item())

    # ---- dominant kernel (FP32 MSAC scoring) timed alone on this rank's shard
    k_ms, kr = sb.kernel_alone(20 if sb.count * sb.nfeat < 1e10 else 5)
    kernel_name = "k_score_resident" if kr.score_kernel == 1 else "k_score_fast"
    roofline = roofline_block(name, kernel_name, k_ms, kr.n_scored, kr.n_exact_hyp, sb.ns, sb.no)

    # ---- end to end through the C ABI with HOST buffers (pinned), H2D + D2H inside the timed region
    e2e_step = sb.make_e2e()
    for _ in range(3):
        e2e_best = e2e_step()
    if world > 1:
        dist.barrier()
    sampler.active = rank == 0 or os.environ.get("GCR_BENCH_SAMPLE_ALL") is not None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_best = e2e_step()
    e2e_s = (time.perf_counter() - t0) / args.steps
    sampler.active = False
    sampler.stop_flag = True
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    ok = ok and e2e_best["index"] == res.best_index and e2e_best["score"] == res.best_score
    e2e = {"value": sb.total * sb.nfeat / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(sb.h2d), "d2h_bytes_per_step": int(sb.d2h),
           "ms_per_step": 1e3 * e2e_s, "exchange_in_e2e": bool(fused), "api": E2E_API}

    # ---- secondary: the replay-parity configuration C2 (N = 1 line only)
    secondary = None
    if world == 1 and name != "C2" and not args.no_secondary:
        if rank == 0 and not args.no_cpu:
            pass  # (C2's CPU rate is in the reference arm's `--config C2` line; one bounded CPU leg per run keeps the default run short)
        s2 = StageBench("C2", torch, gcr_capi, local, 0, 1, stream, flush, False)
        ev2 = s2.device_steps(max(args.steps, 20), 3)
        torch.cuda.synchronize()
        ms2 = float(np.mean([a.elapsed_time(b) for a, b in ev2]))
        ok2, r2, _ = winner_check(s2, torch, dist, gcr_capi)
        k2, kr2 = s2.kernel_alone(20)
        rf2 = roofline_block("C2", "k_score_resident" if kr2.score_kernel == 1 else "k_score_fast", k2, kr2.n_scored, kr2.n_exact_hyp, s2.ns, s2.no)
        st2 = s2.make_e2e()
        for _ in range(3):
            st2()
        t0 = time.perf_counter()
        for _ in range(max(args.steps, 20)):
            st2()
        e2 = (time.perf_counter() - t0) / max(args.steps, 20)
        secondary = {"workload": workload_string("C2"), "value": s2.total * s2.nfeat / (ms2 * 1e-3), "ms_per_step": ms2,
                     "value_scored": r2.n_scored * s2.nfeat / (ms2 * 1e-3), "hypotheses_scored_per_step": int(r2.n_scored),
                     "winner_check": "ok" if ok2 else "MISMATCH",
                     "roofline": {k: rf2[k] for k in ("kernel", "kernel_ms", "achieved", "peak", "frac", "kernel_evals_per_s", "sfu_frac") if k in rf2},
                     "e2e": {"value": s2.total * s2.nfeat / e2, "ms_per_step": 1e3 * e2, "h2d_bytes_per_step": int(s2.h2d),
                             "d2h_bytes_per_step": int(s2.d2h)}}
        ok = ok and ok2

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": warm, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
                "value_scored": n_scored_total * sb.nfeat / (ms_per_step * 1e-3),
                "hypotheses_per_s": sb.total / (ms_per_step * 1e-3),
                "config": {"workload": workload_string(name),
                           "sharding": f"{sb.total} hypothesis slots split by contiguous global index range over {world} rank(s) (gcr_parallel.shard_range)",
                           "l2": "flushed between timed steps (256 MiB write)",
                           "exchange": ("fused peer-memory stores in k_tail (NVLink P2P, 64 B per rank pair)" if fused else
                                        ("nccl all_gather of 64 B records" if world > 1 else "none (1 GPU)")),
                           "winner_check": "ok" if ok else "MISMATCH",
                           "winner_check_what": ("every rank's exchanged winner == rank 0's unsharded run (index, score, model, counts, masks)" if world > 1 else
                                                 "unsharded run == two half-range runs combined by the selection rule == FP64 scoring of the winning model; e2e winner identical"),
                           "hypotheses_solved_per_step": n_solved_total, "hypotheses_scored_per_step": n_scored_total,
                           "guard_band_pairs": int(res.n_rechecked), "exact_hypotheses": int(res.n_exact_hyp), "candidates_rescored": int(res.n_candidates),
                           "best_index": int(res.best_index), "best_score": float(res.best_score)},
                "clocks": sampler.summary(), "e2e": e2e, "gpu_launches": 3 * args.steps, "roofline": roofline,
                "cpu_baseline": cpu_leg}
        if secondary:
            line["secondary"] = secondary
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if not ok:
        sys.exit(3)


# ----------------------------------------------------------------------------------------------------------------------
def run_batched(args):
    """C4 (BASELINE.json configs[3]): 256 independent images x (4000 + 1000) features, 100k device-sampled hypotheses each,
    256 / world images per rank, no communication.  A step = the rank's whole image slice; images are spread over `--engines`
    engine handles, each on its own stream, so that the low-occupancy kernels of one image (sample+solve, tail) overlap the
    scoring kernel of another.  e2e: pinned host feature arrays in (gcr_set_features per image), one device-to-host copy of
    all winner records out."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    c = gs.CONFIGS[args.config]
    cpu_leg = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu_leg, _ = cpu_reference_leg(args.config, scene=gs.config_scene(args.config, image=0), seed=1000)
        cpu_leg["sample"] += " (image 0 of the batch)"

    import torch
    import torch.distributed as dist
    import gcr_capi

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n_img = c["n_images"] // world if args.images is None else args.images
    n_hyp, nfeat, ns, no = c["n_hyp"], c["n_scale"] + c["n_orient"], c["n_scale"], c["n_orient"]
    scenes = [gs.config_scene(args.config, image=rank * n_img + i) for i in range(n_img)]
    d_scale = [torch.from_numpy(sc.scale).cuda() for sc in scenes]
    d_orient = [torch.from_numpy(sc.orient).cuda() for sc in scenes]
    h_scale = [torch.from_numpy(sc.scale).pin_memory() for sc in scenes]
    h_orient = [torch.from_numpy(sc.orient).pin_memory() for sc in scenes]
    K = max(1, min(args.engines, n_img))
    main = torch.cuda.Stream()
    streams = [torch.cuda.Stream() for _ in range(K)]
    engines = [gcr_capi.Engine(local) for _ in range(K)]
    for e, st in zip(engines, streams):
        e.set_stream(st.cuda_stream)
        e.set_sm_share(max(1, args.sm_share))
    records = torch.zeros(n_img, 8, dtype=torch.float64, device="cuda")
    h_records = torch.zeros(n_img, 8, dtype=torch.float64).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    seeds = [1000 + rank * n_img + i for i in range(n_img)]

    def step(host):
        start = torch.cuda.Event()
        start.record(main)
        for st in streams:
            st.wait_event(start)
        for i in range(n_img):
            e = engines[i % K]
            if host:
                e._check(e.lib.gcr_set_features(e.h, h_scale[i].data_ptr(), ns, h_orient[i].data_ptr(), no, 0))
            else:
                e.set_features_device(d_scale[i].data_ptr(), ns, d_orient[i].data_ptr(), no)
            e.ns, e.no = ns, no
            e.run(n_hyp, scenes[i].thr, seed=seeds[i], record_ptr=records[i].data_ptr(), sync=False)
        for st in streams:
            done = torch.cuda.Event()
            done.record(st)
            main.wait_event(done)

    with torch.cuda.stream(main):
        for _ in range(max(args.warmup, 3)):
            flush.zero_()
            step(False)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        sampler = ClockSampler(local)
        sampler.start()
        sampler.active = rank == 0
        evs = []
        for _ in range(args.steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(main)
            step(False)
            e1.record(main)
            evs.append((e0, e1))
        torch.cuda.synchronize()
        sampler.active = False
        total_ms = sum(a.elapsed_time(b) for a, b in evs)
        t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.barrier()
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_per_step = float(t.item()) / args.steps
        total_img = world * n_img
        value = total_img * n_hyp * nfeat / (ms_per_step * 1e-3)
        recs = np.frombuffer(records.cpu().numpy().tobytes(), dtype=gcr_capi.RECORD_DTYPE).copy()

        # e2e: host features in, all records out
        for _ in range(2):
            step(True)
            h_records.copy_(records, non_blocking=True)
            torch.cuda.synchronize()
        sampler.active = rank == 0
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step(True)
            h_records.copy_(records, non_blocking=True)
            main.synchronize()
        e2e_s = (time.perf_counter() - t0) / args.steps
        sampler.active = False
        sampler.stop_flag = True
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
        recs_e2e = np.frombuffer(h_records.numpy().tobytes(), dtype=gcr_capi.RECORD_DTYPE)
        same = bool(np.array_equal(recs["index"], recs_e2e["index"]) and np.array_equal(recs["score"], recs_e2e["score"]))

    # the scoring kernel alone (image 0 of this rank, one engine, nothing else in flight) + winner check of that image
    one = gcr_capi.Engine(local)
    one.set_features(scenes[0].scale, scenes[0].orient)
    ks, kr = [], None
    for _ in range(10):
        kr = one.run(n_hyp, scenes[0].thr, seed=seeds[0], flags=gcr_capi.RUN_TIME_KERNELS)
        ks.append(kr.score_kernel_ms)
    b0 = one.best()
    score0, n0, sm0, om0 = one.score_model(b0["model"], scenes[0].thr)
    same = same and int(recs[0]["index"]) == kr.best_index and float(recs[0]["score"]) == kr.best_score and score0 == b0["score"] and \
        np.array_equal(sm0, b0["scale_mask"]) and np.array_equal(om0, b0["orient_mask"])
    roofline = roofline_block(args.config, "k_score_resident" if kr.score_kernel == 1 else "k_score_fast", float(np.mean(ks)), kr.n_scored,
                              kr.n_exact_hyp, ns, no)
    roofline["note"] += "; C4: the kernel timed alone on ONE image of the batch (in the step, kernels of several images overlap)"
    err = [float(np.hypot(*(r["model"][:2] - sc.model[:2])) / np.hypot(*sc.model[:2])) for r, sc in zip(recs, scenes)]
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": DTYPE, "data": "synthetic",
                "config": {"workload": workload_string(args.config),
                           "sharding": f"{total_img} images, {n_img} per rank, {K} engines/streams in flight per GPU, each sized for 1/{max(1, args.sm_share)} of the SMs "
                                       "(gcr_set_sm_share), no communication",
                           "l2": "flushed between timed steps (256 MiB write)", "images_with_model": int((recs["index"] >= 0).sum()),
                           "median_vanishing_line_error": float(np.median(err)),
                           "winner_check": "ok" if same else "MISMATCH",
                           "winner_check_what": "device-input step == host-input (e2e) step for every image (index, score); image 0: batch record == "
                                                "single-engine run == FP64 scoring of the winning model (score, masks)"},
                "clocks": sampler.summary(),
                "e2e": {"value": total_img * n_hyp * nfeat / e2e_s, "unit": UNIT, "ms_per_step": 1e3 * e2e_s,
                        "h2d_bytes_per_step": int(n_img * 24 * nfeat), "d2h_bytes_per_step": int(n_img * 64),
                        "api": "per image gcr_set_features (pinned host rows, read in place) + asynchronous gcr_run with a device record; one copy of all records back"},
                "gpu_launches": 4 * n_img * args.steps, "roofline": roofline, "cpu_baseline": cpu_leg,
                "images_per_s": total_img / (ms_per_step * 1e-3)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if not same:
        sys.exit(3)


# ----------------------------------------------------------------------------------------------------------------------
def run_sweep(args):
    """C5 (BASELINE.json configs[4]): inlier-ratio x threshold sweep (5 ratios x levels 1/2/4/8: thr_orientation = p degrees,
    thr_scale = 0.05 p) at 1M device-sampled hypotheses per setting on 2000 + 500 features, plus the roofline fraction against
    the feature count.  A step = one pass over the 20 settings.  Every winner is checked against the CPU oracle's MSAC of
    the same model (masks identical, score equal).  One GPU (N > 1: every rank runs the same sweep, replicas only)."""
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    c = gs.CONFIGS["C5"]
    H = int(args.hyp or c["n_hyp"])
    settings = []
    k = 0
    for ratio in (0.05, 0.10, 0.20, 0.30, 0.50):
        for p in (1, 2, 4, 8):
            settings.append((ratio, p, gs.make_scene(c["n_scale"], c["n_orient"], ratio, c["seed"] + k, thr_scale=0.05 * p,
                                                     thr_orient=math.radians(p)), 5000 + k))
            k += 1
    cpu_leg = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu_leg, _ = cpu_reference_leg("C5", scene=settings[13][2], seed=settings[13][3])
        cpu_leg["sample"] += " (setting 30% inliers, level 2)"
    from oracle_lib import Oracle
    port = Oracle("port")

    import torch
    import gcr_capi
    torch.cuda.set_device(local)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    eng = gcr_capi.Engine(local)
    eng.set_stream(stream.cuda_stream)
    nfeat = c["n_scale"] + c["n_orient"]
    dev = [(torch.from_numpy(sc.scale).cuda(), torch.from_numpy(sc.orient).cuda()) for _, _, sc, _ in settings]

    def sweep_pass(timed):
        evs = []
        for (ratio, p, sc, seed), (ds, do) in zip(settings, dev):
            eng.set_features_device(ds.data_ptr(), len(sc.scale), do.data_ptr(), len(sc.orient))
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            eng.run(H, sc.thr, seed=seed, sync=False)
            e1.record(stream)
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return [a.elapsed_time(b) for a, b in evs]

    for _ in range(max(args.warmup, 3)):
        sweep_pass(False)
    per = np.array([sweep_pass(True) for _ in range(max(3, args.steps // 4))])  # [passes, 20] ms: per-setting stage times, one engine, one at a time
    sequential_ms = float(per.sum(axis=1).mean())

    # The timed step: the 20 settings are independent runs, spread over K engines / streams, each engine sized for 1/share of the SMs
    # (gcr_set_sm_share) so that one setting's latency-bound kernels run under another's scoring; the L2 is flushed before every sweep.
    K, share = max(1, min(args.sweep_engines, len(settings))), max(1, args.sweep_sm_share)
    main_stream = torch.cuda.Stream()
    streams = [torch.cuda.Stream() for _ in range(K)]
    engines = [gcr_capi.Engine(local) for _ in range(K)]
    for e, st in zip(engines, streams):
        e.set_stream(st.cuda_stream)
        e.set_sm_share(share)
    records = torch.zeros(len(settings), 8, dtype=torch.float64, device="cuda")
    h_records = torch.zeros(len(settings), 8, dtype=torch.float64).pin_memory()
    hs = [(torch.from_numpy(sc.scale).pin_memory(), torch.from_numpy(sc.orient).pin_memory()) for _, _, sc, _ in settings]

    def sweep_step(host):
        start = torch.cuda.Event()
        start.record(main_stream)
        for st in streams:
            st.wait_event(start)
        for i, ((ratio, p, sc, seed), (ds, do), (ha, hb)) in enumerate(zip(settings, dev, hs)):
            e = engines[i % K]
            if host:
                e._check(e.lib.gcr_set_features(e.h, ha.data_ptr(), len(sc.scale), hb.data_ptr(), len(sc.orient), 0))
                e.ns, e.no = len(sc.scale), len(sc.orient)
            else:
                e.set_features_device(ds.data_ptr(), len(sc.scale), do.data_ptr(), len(sc.orient))
            e.run(H, sc.thr, seed=seed, record_ptr=records[i].data_ptr(), sync=False)
        for st in streams:
            done = torch.cuda.Event()
            done.record(st)
            main_stream.wait_event(done)

    sampler = ClockSampler(local)
    sampler.start()
    with torch.cuda.stream(main_stream):
        for _ in range(max(args.warmup, 3)):
            flush.zero_()
            sweep_step(False)
        torch.cuda.synchronize()
        sampler.active = rank == 0
        evs = []
        for _ in range(args.steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(main_stream)
            sweep_step(False)
            e1.record(main_stream)
            evs.append((e0, e1))
        torch.cuda.synchronize()
        sampler.active = False
        ms_per_step = float(np.mean([a.elapsed_time(b) for a, b in evs]))
        recs_dev = np.frombuffer(records.cpu().numpy().tobytes(), dtype=gcr_capi.RECORD_DTYPE).copy()
        # e2e: pinned host features per setting in, all records out
        for _ in range(2):
            sweep_step(True)
            h_records.copy_(records, non_blocking=True)
            torch.cuda.synchronize()
        sampler.active = rank == 0
        t0 = time.perf_counter()
        for _ in range(args.steps):
            sweep_step(True)
            h_records.copy_(records, non_blocking=True)
            main_stream.synchronize()
        e2e_s = (time.perf_counter() - t0) / args.steps
        sampler.active = False
        sampler.stop_flag = True
        recs_e2e = np.frombuffer(h_records.numpy().tobytes(), dtype=gcr_capi.RECORD_DTYPE).copy()
    value = len(settings) * H * nfeat / (ms_per_step * 1e-3)

    def one_setting(sc, seed, n_hyp):
        eng.set_features(sc.scale, sc.orient)
        rs = [eng.run(n_hyp, sc.thr, seed=seed, flags=gcr_capi.RUN_TIME_KERNELS) for _ in range(5)]
        kern = float(np.mean([r.score_kernel_ms for r in rs[1:]]))
        res = eng.run(n_hyp, sc.thr, seed=seed)
        b = eng.best()
        good = True
        if res.best_index >= 0:
            val, n_inl, sm, om = port.score_model(sc.scale, sc.orient, b["model"], sc.thr)
            good = bool(np.array_equal(sm, b["scale_mask"]) and np.array_equal(om, b["orient_mask"]) and abs(val - b["score"]) <= 1e-9 * max(1.0, abs(val)))
        rf = roofline_block("C5", "k_score_resident" if res.score_kernel == 1 else "k_score_fast", kern, res.n_scored, res.n_exact_hyp,
                            len(sc.scale), len(sc.orient))
        return rf, good, res

    rows, all_ok, n_scored = [], True, 0
    for i, (ratio, p, sc, seed) in enumerate(settings):
        rf, good, res = one_setting(sc, seed, H)
        # ... and the timed sweep's own record of this setting (concurrent engines of a share of the SMs, device and host inputs) == this
        # full-size, one-at-a-time run == the oracle's scoring of the model
        good = good and int(recs_dev[i]["index"]) == res.best_index and float(recs_dev[i]["score"]) == res.best_score and \
            int(recs_e2e[i]["index"]) == res.best_index and float(recs_e2e[i]["score"]) == res.best_score
        all_ok &= good
        n_scored += res.n_scored
        rows.append({"inliers": ratio, "level": p, "stage_ms": float(per[:, i].mean()), "kernel_ms": rf["kernel_ms"], "frac": rf["frac"],
                     "kernel_evals_per_s": rf["kernel_evals_per_s"], "scored": int(res.n_scored), "candidates": int(res.n_candidates),
                     "oracle": "ok" if good else "MISMATCH"})
    ladder = []
    for ns_, no_ in ((500, 125), (2000, 500), (8000, 2000), (20000, 8000), (50000, 20000)):
        sc = gs.make_scene(ns_, no_, 0.10, 2025, thr_scale=0.15, thr_orient=math.radians(3))
        rf, good, res = one_setting(sc, 9, H // 4)
        all_ok &= good
        ladder.append({"n_scale": ns_, "n_orient": no_, "kernel": rf["kernel"], "kernel_ms": rf["kernel_ms"], "frac": rf["frac"],
                       "kernel_evals_per_s": rf["kernel_evals_per_s"], "oracle": "ok" if good else "MISMATCH"})
    fr = np.array([r["frac"] for r in rows])
    mid = rows[int(np.argsort(fr)[len(fr) // 2])]
    sc_mid = [s for s in settings if s[0] == mid["inliers"] and s[1] == mid["level"]][0]
    roofline, _, _ = one_setting(sc_mid[2], sc_mid[3], H)
    roofline.update({"frac_min": float(fr.min()), "frac_median": float(np.median(fr)), "frac_max": float(fr.max()),
                     "frac_vs_feature_count": ladder, "per_setting": rows})
    roofline["note"] += "; C5: the block's kernel figures are the median setting's, frac_min/median/max over the 20 settings"

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
                "value_scored": n_scored * nfeat / (ms_per_step * 1e-3),
                "config": {"workload": workload_string("C5"), "l2": "flushed before every timed sweep (256 MiB write)",
                           "sharding": f"the 20 settings are independent runs: {K} engines / streams in flight, each sized for 1/{share} of the SMs "
                                       "(gcr_set_sm_share)",
                           "sequential_ms_per_sweep": sequential_ms,
                           "sequential_what": "the same 20 runs one at a time on one full-size engine, L2 flushed before every run (the per_setting stage_ms)",
                           "winner_check": "ok" if all_ok else "MISMATCH",
                           "winner_check_what": "every winner of the 20 settings and the 5 ladder sizes == the CPU oracle's MSAC of the same model (masks identical, "
                                                "score equal); the timed sweep's records (device and host inputs) == the one-at-a-time runs"},
                "clocks": sampler.summary(),
                "e2e": {"value": len(settings) * H * nfeat / e2e_s, "unit": UNIT, "ms_per_step": 1e3 * e2e_s,
                        "h2d_bytes_per_step": int(len(settings) * 24 * nfeat), "d2h_bytes_per_step": int(len(settings) * 64),
                        "api": "per setting gcr_set_features (pinned host rows, read in place) + asynchronous gcr_run with a device record; one copy of all records back"},
                "gpu_launches": 3 * len(settings) * args.steps, "roofline": roofline, "cpu_baseline": cpu_leg}
        print(json.dumps(line))
    if not all_ok:
        sys.exit(3)


def front_keypoints(n: int, seed: int) -> np.ndarray:
    """Synthetic cv2.KeyPoint rows (x, y, size, angle in degrees) as float32: a 4096 x 4096 image (about one keypoint in eight shares its
    integer coordinate with an earlier one), one keypoint in five without an orientation (angle == -1)."""
    rng = np.random.default_rng(seed)
    kp = np.empty((n, 4), np.float32)
    kp[:, 0] = rng.uniform(0.0, 4096.0, n)
    kp[:, 1] = rng.uniform(0.0, 4096.0, n)
    kp[:, 2] = rng.uniform(1.5, 40.0, n)
    kp[:, 3] = np.where(rng.uniform(size=n) < 0.2, -1.0, rng.uniform(0.0, 360.0, n))
    return kp


def run_front(args):
    """--config FRONT (not a BASELINE.json configuration): the feature front end gcr_set_keypoints on 4 Mi keypoints per
    GPU -- de-duplication by integer coordinate, stable compaction, orientation rows, derived feature planes.  HBM-bound integer / copy
    work: the roofline is the measured copy bandwidth of MEASURED_PEAKS.json."""
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    n = 1 << 22
    kp = front_keypoints(n, 77 + rank)
    cpu = None
    if rank == 0 and (not args.no_cpu or args.impl == "reference"):  # before torch: the numpy restatement of examples/utils.py:5-49 on a bounded sample, one core
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import frontend_oracle as fo
        m = 1 << 19
        t0 = time.perf_counter()
        want_s = fo.scale_features_from_rows(kp[:m])
        want_o, want_sz = fo.orientation_features_from_rows(kp[:m])
        dt = time.perf_counter() - t0
        cpu = {"value": m / dt, "unit": "keypoints/s", "cores": 1, "kind": "port",
               "sample": f"first {m} of {n} keypoint rows, numpy restatement of examples/utils.py:5-49 (oracle/frontend_oracle.py), {dt:.2f} s"}
    if args.impl == "reference":  # the reference's front end is host Python: its numpy restatement, one core, rank 0 only
        if rank == 0:
            print(json.dumps({"impl": "reference", "metric": "keypoints/s (feature front end, examples/utils.py:5-49)", "value": cpu["value"], "unit": "keypoints/s",
                              "n_gpus": world, "steps": 1, "warmup": 0, "ms_per_step": 1e3 * (1 << 19) / cpu["value"], "higher_is_better": True, "scaling": "weak",
                              "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": {"workload": f"FRONT: {n} float32 keypoint rows per GPU"},
                              "cpu_baseline": cpu, "e2e": {"value": cpu["value"], "unit": "keypoints/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    import torch
    import torch.distributed as dist
    import gcr_capi
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    C = gcr_capi.C
    eng = gcr_capi.Engine(local)
    stream = torch.cuda.Stream()
    eng.set_stream(stream.cuda_stream)
    d_kp = torch.from_numpy(kp).cuda()
    h_kp = torch.from_numpy(kp).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    check = None
    if cpu is not None:  # parity of this very input against the restatement (bit for bit), on the sample the CPU leg ran
        ns_, no_, sz = eng.set_keypoints(device_ptr=d_kp.data_ptr(), n=1 << 19, is_f32=True)
        got_s, got_o = eng.get_features()
        check = "ok" if (np.array_equal(got_s, want_s) and np.array_equal(got_o, want_o) and np.array_equal(sz, want_sz)) else "MISMATCH"
    n_s, n_o = C.c_int64(0), C.c_int64(0)

    def step(host, rows=n):
        rc = eng.lib.gcr_set_keypoints(eng.h, C.c_void_p(h_kp.data_ptr() if host else d_kp.data_ptr()), rows, 1, 0 if host else 1, C.byref(n_s), C.byref(n_o),
                                       h_sizes.ctypes.data if host else None)
        assert rc == 0, eng.lib.gcr_last_error(eng.h)
    h_sizes = np.zeros(n)
    # one image's worth of keypoints (65 536 rows): the call is six dependent launches and a synchronisation, not bandwidth
    small = []
    with torch.cuda.stream(stream):
        for k in range(13):
            flush.zero_()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            step(False, 1 << 16)
            if k >= 3:
                small.append((time.perf_counter() - t0) * 1e6)
    small_us = float(np.median(small))
    with torch.cuda.stream(stream):
        for _ in range(max(args.warmup, 3)):
            flush.zero_()
            step(False)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        sampler = ClockSampler(local)
        sampler.start()
        sampler.active = rank == 0
        evs = []
        for _ in range(args.steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            step(False)
            e1.record(stream)
            evs.append((e0, e1))
        torch.cuda.synchronize()
        sampler.active = False
        t = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.barrier()
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_per_step = float(t.item()) / args.steps
        for _ in range(2):
            step(True)
        sampler.active = rank == 0
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step(True)
        e2e_s = (time.perf_counter() - t0) / args.steps
        sampler.active = False
        sampler.stop_flag = True
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    if rank == 0:
        peaks = _load_json(os.path.join(ROOT, "MEASURED_PEAKS.json")) or {}
        peak = float(peaks.get("hbm_gbs", 0.0)) or 7700.0
        ns_out, no_out = int(n_s.value), int(n_o.value)
        # algorithmic bytes of one call: every keypoint row read once, every output row written once (scale rows, orientation rows, sizes)
        alg = 16 * n + 24 * ns_out + 24 * no_out + 8 * no_out
        achieved = alg / (ms_per_step * 1e-3) / 1e9
        line = {"metric": "keypoints/s (feature front end, examples/utils.py:5-49)", "value": world * n / (ms_per_step * 1e-3), "unit": "keypoints/s",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32 rows in, f64 rows out (integer keys)", "data": "synthetic",
                "config": {"workload": f"FRONT: {n} float32 keypoint rows per GPU on a 4096 x 4096 image -> {ns_out} scale rows (unique integer "
                                       f"coordinates, first wins, order of first appearance) + {no_out} orientation rows (angle != -1)",
                           "l2": "flushed between timed steps (256 MiB write)", "note": "SURVEY 8f row 4; not a BASELINE.json configuration",
                           "one_image_call_us": small_us, "one_image_what": "wall time of gcr_set_keypoints on 65 536 device-resident rows (L2 flushed): launch / "
                           "synchronisation bound, six dependent kernels",
                           "front_check": check, "front_check_what": "first 524288 rows: scale rows, orientation rows and sizes == the numpy restatement "
                           "of the reference's functions, bit for bit"},
                "clocks": sampler.summary(),
                "e2e": {"value": world * n / e2e_s, "unit": "keypoints/s", "ms_per_step": e2e_s * 1e3, "h2d_bytes_per_step": 16 * n, "d2h_bytes_per_step": 8 * no_out + 16,
                        "api": "gcr_set_keypoints on a page-locked host array (copied to the device inside the call), sizes copied back"},
                "gpu_launches": 6 * args.steps,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                             "note": "the whole call (memset of the hash table, insert, count, scan, compact, derived feature planes) against the measured "
                                     "copy bandwidth of MEASURED_PEAKS.json" + ("" if peaks else " (absent: nominal 7.7 TB/s)") +
                                     "; algorithmic bytes = 16 B per keypoint in + 24 B per scale row + 32 B per orientation row out"},
                "cpu_baseline": cpu}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C3", choices=sorted(gs.CONFIGS) + ["FRONT"])
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"], help="N>1: fused peer-memory exchange or NCCL all_gather")
    ap.add_argument("--engines", type=int, default=8, help="C4: engine handles / streams in flight per GPU")
    ap.add_argument("--sm-share", type=int, default=8, help="C4: every engine sizes its launches for 1/share of the SMs (gcr_set_sm_share)")
    ap.add_argument("--images", type=int, default=None, help="C4: images per GPU (default 256 / world)")
    ap.add_argument("--sweep-engines", type=int, default=4, help="C5: engine handles / streams in flight")
    ap.add_argument("--sweep-sm-share", type=int, default=4, help="C5: every engine sizes its launches for 1/share of the SMs")
    ap.add_argument("--hyp", type=int, default=None, help="override the configuration's total hypothesis count (diagnostics)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the bounded CPU baseline leg (diagnostics)")
    ap.add_argument("--no-secondary", action="store_true", help="skip the C2 secondary block (diagnostics)")
    args = ap.parse_args()
    if args.config == "FRONT":
        run_front(args)
    elif args.impl == "reference":
        run_reference(args)
    elif "n_images" in gs.CONFIGS[args.config]:
        run_batched(args)
    elif args.config == "C5":
        run_sweep(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
