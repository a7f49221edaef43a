"""Prints the device-timer breakdown of one hypothesis-stage run (GPU box): stage, score kernel, tail phases."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")]
import numpy as np, torch, gcr_capi, gcr_scenes as gs
for name in sys.argv[1:] or ["C2"]:
    c = dict(gs.CONFIGS[name]); sc = gs.config_scene(name)
    if os.environ.get("HYP"): c["n_hyp"] = int(os.environ["HYP"])  # e.g. one rank's shard of a sharded run
    smp = gs.draw_samples(c["n_hyp"], c["n_scale"], c["n_orient"], gs.SAMPLE_SEED)
    e = gcr_capi.Engine(0); e.set_features(sc.scale, sc.orient)
    d = torch.from_numpy(smp).cuda()
    for _ in range(3): r = e.run(c["n_hyp"], sc.thr, samples_ptr=d.data_ptr())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda") if os.environ.get("FLUSH") else None  # cold L2 per run, as bench.py
    rs = []
    for _ in range(10):
        if flush is not None: flush.zero_(); torch.cuda.synchronize()
        rs.append(e.run(c["n_hyp"], sc.thr, samples_ptr=d.data_ptr(), flags=gcr_capi.RUN_TIME_KERNELS if os.environ.get("TIME_KERNELS") else 0))  # the event bracket switches the chained (PDL) launches off
    ph = np.median([list(r.tail_phase_us) + list(r.stage_phase_us) for r in rs], axis=0)
    print(name, "stage_ms %.4f score_ms %.4f" % (np.median([r.device_ms for r in rs]), np.median([r.score_kernel_ms for r in rs])),
          "tail phases us (fix+exact | finalize | collect | rescoring | select):", np.round(ph[:5], 2), "sum %.2f" % ph[:5].sum(),
          "| ss->score %.2f score->tail %.2f ss->end %.2f |" % (ph[6] if len(ph) > 6 else 0, ph[5], ph[7] if len(ph) > 7 else 0),
          "refined", rs[0].n_refined, "cands", rs[0].n_candidates, "gb", rs[0].n_rechecked, "exact", rs[0].n_exact_hyp)
    e.close()
