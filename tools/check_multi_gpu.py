"""torchrun check of the multi-GPU path (run on a box with >= 2 GPUs):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tools/check_multi_gpu.py
Every rank shards the same global run; the fused peer-memory exchange, the NCCL all_gather form and an unsharded run on
rank 0 must all name the same winner (index, score, model, masks)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "graph-cut-ransac_b200"), os.path.join(ROOT, "tests")]
import gcr_capi  # noqa: E402
import gcr_parallel  # noqa: E402
import gcr_scenes as gs  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sc = gs.make_scene(3000, 900, 0.25, 2026)
    n_total, seed = 200_000, 99
    eng = gcr_capi.Engine(local)
    eng.set_features(sc.scale, sc.orient)
    gcr_parallel.attach_peers(eng, rank, world)
    ok = True
    for rep in range(5):  # repeated exchanges exercise the sequence numbers / slot parity
        fused = gcr_parallel.run_sharded_fused(eng, n_total, sc.thr, seed + rep, rank, world)
        rec = torch.zeros(8, dtype=torch.float64, device="cuda")
        nccl = gcr_parallel.run_sharded(eng, n_total, sc.thr, seed + rep, rank, world, rec)
        same = (fused["index"] == nccl["index"] and fused["score"] == nccl["score"] and np.array_equal(fused["model"], nccl["model"])
                and np.array_equal(fused["scale_mask"], nccl["scale_mask"]) and np.array_equal(fused["orient_mask"], nccl["orient_mask"])
                and fused["rank"] == nccl["rank"])
        if rank == 0:
            whole = gcr_capi.Engine(local)
            whole.set_features(sc.scale, sc.orient)
            r = whole.run(n_total, sc.thr, seed=seed + rep)
            b = whole.best()
            same = same and r.best_index == fused["index"] and r.best_score == fused["score"] and np.array_equal(
                b["scale_mask"], fused["scale_mask"]) and np.array_equal(b["orient_mask"], fused["orient_mask"])
            whole.close()
        flag = torch.tensor([1 if same else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = ok and bool(flag.item())
        if rank == 0:
            print(f"rep {rep}: winner #{fused['index']} on rank {fused['rank']} score {fused['score']:.4f} -> {'ok' if flag.item() else 'MISMATCH'}")
    # timing of the two exchange forms at this size (device events, max over ranks)
    for name, flags in (("fused", gcr_capi.RUN_EXCHANGE), ("nccl", 0)):
        first, count = gcr_parallel.shard_range(n_total, rank, world)
        rec = torch.zeros(8, dtype=torch.float64, device="cuda")
        allr = torch.zeros(8 * world, dtype=torch.float64, device="cuda")
        st = torch.cuda.Stream()
        torch.cuda.set_stream(st)
        eng.set_stream(st.cuda_stream)
        def step():
            eng.run(count, sc.thr, seed=seed, first_hyp=first, flags=flags, record_ptr=rec.data_ptr(), sync=False)
            if not flags:
                dist.all_gather_into_tensor(allr, rec)
        for _ in range(5):
            step()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(50):
            step()
        e1.record(st)
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 50], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"{name}: {t.item() * 1e3:.1f} us per sharded run ({count} slots per rank, world {world})")
        eng.set_stream(0)
        torch.cuda.set_stream(torch.cuda.default_stream())
        torch.cuda.synchronize()
    eng.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("MULTI_GPU_CHECK", "PASS" if ok else "FAIL")
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
