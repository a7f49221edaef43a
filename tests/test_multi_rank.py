"""world_size-2 tests of the multi-GPU plumbing on CPU (gloo): slot sharding and the winner-record exchange.
The GPU form (NCCL, the record buffer written by k_select) is exercised by `bench.py --gpus N` and the -m gpu test."""
import os
import sys
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_tile_the_slot_space():
    from gcr_parallel import shard_range
    for n in (0, 1, 7, 100_000, 1_000_003):
        for world in (1, 2, 3, 8):
            nxt = 0
            for r in range(world):
                first, count = shard_range(n, r, world)
                assert first == nxt and count >= 0
                nxt = first + count
            assert nxt == n
            counts = [shard_range(n, r, world)[1] for r in range(world)]
            assert max(counts) - min(counts) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_selection_rule_across_ranks():
    from gcr_capi import RECORD_DTYPE
    from gcr_parallel import select_record
    recs = np.zeros(4, RECORD_DTYPE)
    recs["index"] = -1
    assert select_record(recs) == -1
    recs[1] = (10.0, 500, [1, 2, 3, 4], [5, 6], 0)
    recs[3] = (10.0, 499, [1, 2, 3, 4], [5, 6], 0)  # same score, earlier slot -> wins (strict '<', GCRANSAC.h:440)
    assert select_record(recs) == 3
    recs[2] = (10.5, 900, [0, 0, 1, 0], [2, 2], 0)
    assert select_record(recs) == 2
    recs[0] = (0.0, 3, [0, 0, 1, 0], [0, 0], 0)  # score 0 never beats the initial best
    assert select_record(recs) == 2


def _worker(rank, world, init_file, out_dir):
    sys.path.insert(0, os.path.join(ROOT, "graph-cut-ransac_b200"))
    import torch
    import torch.distributed as dist
    from gcr_capi import RECORD_DTYPE
    from gcr_parallel import gather_records, select_record, shard_range
    dist.init_process_group("gloo", init_method=f"file://{init_file}", rank=rank, world_size=world)
    # each rank "scores" its own block of slots: score(slot) is a fixed function of the GLOBAL slot index
    n = 1001
    first, count = shard_range(n, rank, world)
    slots = np.arange(first, first + count)
    score = ((slots * 2654435761) % 1000).astype(np.float64)  # several slots share the maximum 999
    k = int(np.argmax(score))  # first maximum inside the block, as k_select picks
    rec = np.zeros(1, RECORD_DTYPE)
    rec[0] = (score[k], slots[k], [1e-4, 2e-4, 0.5, 0.1 * rank], [rank, rank + 1], 0)
    t = torch.from_numpy(rec.view(np.float64).copy())
    recs = gather_records(t, world)
    w = select_record(recs)
    np.save(os.path.join(out_dir, f"r{rank}.npy"), np.array([w, recs[w]["index"], recs[w]["score"]]))
    dist.barrier()
    dist.destroy_process_group()


def test_record_exchange_two_ranks_gloo():
    import torch.multiprocessing as mp
    with tempfile.TemporaryDirectory() as d:
        init = os.path.join(d, "init")
        mp.spawn(_worker, args=(2, init, d), nprocs=2, join=True)
        a, b = np.load(os.path.join(d, "r0.npy")), np.load(os.path.join(d, "r1.npy"))
    assert np.array_equal(a, b)  # every rank reaches the same verdict
    slots = np.arange(1001)
    score = (slots * 2654435761) % 1000
    assert a[1] == int(np.argmax(score)) and a[2] == score.max()  # = the single-process answer


@pytest.mark.gpu
def test_sharded_run_equals_single_run():
    """Two shards on one GPU (two engines, gloo-free): the combined winner equals the unsharded winner."""
    import gcr_capi
    import gcr_scenes as gs
    import torch
    from gcr_parallel import select_record, shard_range
    sc = gs.make_scene(1200, 400, 0.3, 21)
    n = 30_000
    whole = gcr_capi.Engine(0)
    whole.set_features(sc.scale, sc.orient)
    res = whole.run(n, sc.thr, seed=77)
    recs = np.zeros(3, gcr_capi.RECORD_DTYPE)
    for r in range(3):
        first, count = shard_range(n, r, 3)
        e = gcr_capi.Engine(0)
        e.set_features(sc.scale, sc.orient)
        buf = torch.zeros(8, dtype=torch.float64, device="cuda")
        e.run(count, sc.thr, seed=77, first_hyp=first, record_ptr=buf.data_ptr())
        recs[r] = np.frombuffer(buf.cpu().numpy().tobytes(), dtype=gcr_capi.RECORD_DTYPE)[0]
        e.close()
    w = select_record(recs)
    assert recs[w]["index"] == res.best_index and recs[w]["score"] == res.best_score
    assert np.array_equal(recs[w]["model"], np.array(res.best_model[:]))
    whole.close()


@pytest.mark.gpu
def test_fused_peer_exchange_two_gpus():
    """Two processes, two GPUs: fused NVLink exchange == NCCL all_gather form == unsharded run (tools/check_multi_gpu.py)."""
    import subprocess
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run by `gpurun --gpus 2`)")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                          "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tools", "check_multi_gpu.py")],
                         capture_output=True, text=True, timeout=600)
    assert "MULTI_GPU_CHECK PASS" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
