"""Multi-GPU host logic: residue-balanced sharding of a database and the deterministic top-k merge.

The scoring path shards by independent targets (the reference's `omp parallel for` over batches,
src/alignment_cmdline.c:504-507 of the reference tree): no data-path collective.  The only exchange is
the gather of the per-rank top-k lists, merged with the rule of SURVEY.md A.7 (score descending, then
file-order entry ascending), which makes the result independent of the number of GPUs."""
from __future__ import annotations

import numpy as np


def deal_shards(offsets, world):
    """Deal targets, longest first, round-robin over `world` ranks.  Returns a list of int64 arrays of
    file-order entry indices; every rank gets the same length mix and (within one target per round) the
    same residue count.  Inside a shard the entries stay in ascending file order, so the engine's tie-break
    (score desc, shard-local entry asc) is the global rule and the merged top-k is independent of `world`.
    Same rule as make_shards() in host/alignment_cmdline.c."""
    offsets = np.asarray(offsets, dtype=np.int64)
    lens = offsets[1:] - offsets[:-1]
    order = np.argsort(-lens, kind="stable")
    return [np.sort(order[r::world]) for r in range(world)]


def take_shard(codes, offsets, entries):
    offsets = np.asarray(offsets, dtype=np.int64)
    lens = (offsets[1:] - offsets[:-1])[entries]
    off = np.zeros(len(entries) + 1, dtype=np.int64)
    np.cumsum(lens, out=off[1:])
    out = np.empty(int(off[-1]), dtype=np.int8)
    for k, e in enumerate(entries):
        out[off[k]:off[k + 1]] = codes[offsets[e]:offsets[e + 1]]
    return out, off


def merge_topk(score_entry_lists, k):
    """score_entry_lists: iterable of int64 arrays [m_i, 2] = (score, global entry); rows with entry < 0 are
    padding.  Returns the k best as an int64 array [<=k, 2]."""
    allh = np.concatenate([np.asarray(x, dtype=np.int64).reshape(-1, 2) for x in score_entry_lists])
    allh = allh[allh[:, 1] >= 0]
    order = np.lexsort((allh[:, 1], -allh[:, 0]))
    return allh[order[:k]]


def gather_topk(local_hits, k, dist=None, device=None):
    """All-gathers fixed-size [k, 2] hit lists over torch.distributed (NCCL on GPUs, gloo in CPU tests)
    and merges them.  With dist=None it is the single-rank merge."""
    import torch
    pad = np.full((k, 2), -1, dtype=np.int64)
    lh = np.asarray(local_hits, dtype=np.int64).reshape(-1, 2)[:k]
    pad[: len(lh)] = lh
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return merge_topk([pad], k)
    t = torch.from_numpy(pad)
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return merge_topk([o.cpu().numpy() for o in out], k)


def merge_topk_batch(lists, k):
    """lists: int64 [world, n_queries, k, 2] (score, global entry; entry < 0 = padding).  One lexsort for all the
    queries of a batch: key (query asc, score desc, entry asc).  Returns int64 [n_queries, k, 2], padded with -1."""
    a = np.asarray(lists, dtype=np.int64)
    w, nq, kk, _ = a.shape
    rows = a.transpose(1, 0, 2, 3).reshape(nq, w * kk, 2)
    qid = np.repeat(np.arange(nq, dtype=np.int64), w * kk)
    flat = rows.reshape(-1, 2)
    valid = flat[:, 1] >= 0
    neg = np.where(valid, -flat[:, 0], np.iinfo(np.int64).max)      # padding sorts last within its query
    order = np.lexsort((flat[:, 1], neg, qid))
    out = flat[order].reshape(nq, w * kk, 2)[:, :k, :].copy()
    out[out[:, :, 1] < 0] = -1
    return out


def gather_topk_batch(local, k, dist=None, device=None):
    """ONE exchange for a whole batch of queries: local int64 [n_queries, <=k, 2] hit lists (padded to k with -1)
    -> torch.distributed all_gather (NCCL on GPUs, gloo in the CPU tests) -> merge_topk_batch.  Returns
    (merged [n_queries, k, 2], exchange_ms, merge_ms): exchange_ms is CUDA-event time around H2D + all_gather + D2H
    when `device` is a CUDA device (else host wall clock), merge_ms the host wall clock of the lexsort."""
    import time
    import torch
    loc = np.asarray(local, dtype=np.int64)
    nq = loc.shape[0]
    pad = np.full((nq, k, 2), -1, dtype=np.int64)
    pad[:, : loc.shape[1], :] = loc[:, :k, :]
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        t0 = time.perf_counter()
        out = merge_topk_batch(pad[None], k)
        return out, 0.0, 1e3 * (time.perf_counter() - t0)
    world = dist.get_world_size()
    cuda = device is not None and str(device).startswith("cuda")
    t = torch.from_numpy(pad)
    if cuda:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t = t.pin_memory().to(device, non_blocking=True)
    else:
        w0 = time.perf_counter()
    buf = torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)   # ranks concatenated along dim 0
    dist.all_gather_into_tensor(buf, t)
    host = buf.cpu().reshape((world,) + tuple(t.shape))
    if cuda:
        e1.record(); e1.synchronize()
        ex_ms = e0.elapsed_time(e1)
    else:
        ex_ms = 1e3 * (time.perf_counter() - w0)
    t0 = time.perf_counter()
    out = merge_topk_batch(host.numpy(), k)
    return out, ex_ms, 1e3 * (time.perf_counter() - t0)
