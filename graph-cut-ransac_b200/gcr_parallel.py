"""Multi-GPU plumbing of the hypothesis stage: one process per GPU (torchrun), hypothesis slots sharded by
contiguous global index range, features replicated, ONE exchange of 64-byte winner records.

The reference is single-process; what is mirrored here is its best-model rule
(GCRANSAC.h:440-462: strict '<' on the score, so the earliest iteration wins ties; a model needs score > 0),
applied across ranks exactly as `k_select` applies it inside one GPU.  Philox counters are the GLOBAL slot
indices, so the drawn samples -- and therefore the winner -- do not depend on the number of GPUs.
"""
from __future__ import annotations

import numpy as np

from gcr_capi import RECORD_DTYPE


def shard_range(n_hyp: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block of global hypothesis slots owned by `rank`: (first, count).  Blocks differ by at most one."""
    if not (0 <= rank < world) or n_hyp < 0:
        raise ValueError("bad shard request")
    base, extra = divmod(n_hyp, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def select_record(records: np.ndarray) -> int:
    """Index (rank) of the winning record among `records` (RECORD_DTYPE[world]), -1 when no rank has a model.
    Highest score; on equal scores the lowest global hypothesis index (S11)."""
    best, best_rank = None, -1
    for r, rec in enumerate(records):
        if rec["index"] < 0 or not rec["score"] > 0.0:
            continue
        if best is None or rec["score"] > best["score"] or (rec["score"] == best["score"] and rec["index"] < best["index"]):
            best, best_rank = rec, r
    return best_rank


def gather_records(record_tensor, world: int, group=None):
    """all_gather of one 64-byte record per rank.  `record_tensor`: 8 float64 on the device the backend
    communicates from (CUDA for NCCL -- the buffer k_select wrote, no staging; CPU for gloo).
    Returns RECORD_DTYPE[world] on the host."""
    import torch
    import torch.distributed as dist
    out = torch.empty(8 * world, dtype=torch.float64, device=record_tensor.device)
    dist.all_gather_into_tensor(out, record_tensor, group=group)
    return np.frombuffer(out.cpu().numpy().tobytes(), dtype=RECORD_DTYPE)


def attach_peers(engine, rank: int, world: int, group=None) -> None:
    """One-time setup of the fused winner exchange (include/gcr_b200.h, gcr_peer_*): every rank exports the CUDA IPC
    handle of its exchange buffer, the handles are all-gathered (torch.distributed is only the courier), every rank
    maps its peers' buffers.  After this, runs with RUN_EXCHANGE return the GLOBAL winner on every rank with no
    collective launched: the records travel as peer-memory stores inside the engine's last kernel."""
    import torch.distributed as dist
    handles = [None] * world
    dist.all_gather_object(handles, engine.peer_export(), group=group)
    engine.peer_attach(b"".join(handles), world, rank)
    dist.barrier(group=group)  # nobody runs before every buffer has been cleared and mapped


def run_sharded_fused(engine, n_hyp_total: int, thr, seed: int, rank: int, world: int, samples=None):
    """As run_sharded, with the exchange fused into the engine's last kernel (needs attach_peers).  Returns the same
    dict on every rank, or None."""
    import gcr_capi
    first, count = shard_range(n_hyp_total, rank, world)
    if count < 1:
        raise ValueError("every rank needs at least one hypothesis slot for a fused exchange")
    smp = None if samples is None else np.ascontiguousarray(samples[first:first + count])
    res = engine.run(count, thr, seed=seed, first_hyp=first, samples=smp, flags=gcr_capi.RUN_EXCHANGE, sync=True)
    if res.best_index < 0:
        return None
    b = engine.best()  # masks of a remote winner are recomputed locally from the replicated features
    return dict(model=b["model"], score=b["score"], index=b["index"], n_inl=b["n_inl"], scale_mask=b["scale_mask"],
                orient_mask=b["orient_mask"], rank=int(res.winner_rank))


def run_sharded(engine, n_hyp_total: int, thr, seed: int, rank: int, world: int, record_tensor, samples=None, group=None):
    """Hypothesis stage over `n_hyp_total` global slots split across `world` ranks.  Every rank returns the same
    dict(model, score, index, n_inl, scale_mask, orient_mask, rank) or None.  The winner's masks are recomputed
    locally with gcr_score_model (features are replicated), so no second collective is needed."""
    first, count = shard_range(n_hyp_total, rank, world)
    smp = None if samples is None else np.ascontiguousarray(samples[first:first + count])
    if count > 0:
        engine.run(count, thr, seed=seed, first_hyp=first, samples=smp, record_ptr=record_tensor.data_ptr(), sync=True)
    else:
        record_tensor.zero_()
        record_tensor.view(dtype=__import__("torch").int64)[1] = -1
    recs = gather_records(record_tensor, world, group) if world > 1 else np.frombuffer(
        record_tensor.cpu().numpy().tobytes(), dtype=RECORD_DTYPE)
    w = select_record(recs)
    if w < 0:
        return None
    rec = recs[w]
    score, n_inl, smask, omask = engine.score_model(rec["model"], thr)
    return dict(model=np.array(rec["model"]), score=float(rec["score"]), index=int(rec["index"]), n_inl=n_inl,
                scale_mask=smask, orient_mask=omask, rank=w, rescored=score)
