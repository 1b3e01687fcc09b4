"""Multi-GPU plumbing: one process per GPU, torch.distributed for the rendezvous.

* HNSW search shards by QUERY: the graph is replicated, rank r takes queries
  [r·nq/W, (r+1)·nq/W); there is no data-path collective. `gather_search_results` only exists to
  hand the concatenated answer to a caller that wants it on every rank.
* Exact k-NN shards the BASE rows: every rank scans its rows for all queries, the per-shard
  [nq × k] lists are exchanged with ONE all-gather (NCCL over NVLink on GPUs, gloo in the CPU
  tests) and merged under (distance, row index) by muninn_cuda_topk_merge_device.
"""
from __future__ import annotations

import numpy as np


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced split: sizes differ by at most one, earlier ranks get the extra."""
    base, extra = divmod(total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_search_results(ids, dist, cnt, nq_total: int, group=None):
    """All-gather per-rank search results (torch tensors [nq_r × k], [nq_r × k], [nq_r]) into the
    full batch order. Ranks may hold different nq_r (shard_range); padding is stripped."""
    import torch
    import torch.distributed as dist_

    world = dist_.get_world_size(group)
    k = ids.shape[1]
    pad = max(shard_range(nq_total, r, world)[1] - shard_range(nq_total, r, world)[0] for r in range(world))

    def padded(t, fill):
        out = torch.full((pad,) + tuple(t.shape[1:]), fill, dtype=t.dtype, device=t.device)
        out[: t.shape[0]] = t
        return out

    parts = []
    for t, fill in ((ids, -1), (dist, 0), (cnt, 0)):
        buf = [torch.empty_like(padded(t, fill)) for _ in range(world)]
        dist_.all_gather(buf, padded(t, fill), group=group)
        parts.append(buf)
    out = []
    for buf in parts:
        pieces = []
        for r in range(world):
            lo, hi = shard_range(nq_total, r, world)
            pieces.append(buf[r][: hi - lo])
        out.append(torch.cat(pieces, 0))
    assert out[0].shape == (nq_total, k)
    return tuple(out)


def allgather_topk_parts(idx, dist, cnt, group=None):
    """One all-gather of the per-shard exact-k-NN lists → ([W × nq × k], [W × nq × k], [W × nq]),
    the layout muninn_cuda_topk_merge_device consumes."""
    import torch
    import torch.distributed as dist_

    world = dist_.get_world_size(group)
    outs = []
    for t in (idx, dist, cnt):
        buf = torch.empty((world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
        dist_.all_gather(list(buf.unbind(0)), t.contiguous(), group=group)  # one collective per array; gloo-compatible
        outs.append(buf)
    return tuple(outs)


def merge_topk_parts_reference(idx_parts: np.ndarray, dist_parts: np.ndarray, cnt_parts: np.ndarray, k: int):
    """Host statement of the merge rule, for tests of the gathered layout only: ascending
    (distance, index) over all parts' valid entries, first k."""
    parts, nq, _ = idx_parts.shape
    o_idx = np.full((nq, k), -1, np.int64)
    o_dist = np.zeros((nq, k), np.float32)
    o_cnt = np.zeros(nq, np.int32)
    for q in range(nq):
        ii = np.concatenate([idx_parts[p, q, : cnt_parts[p, q]] for p in range(parts)])
        dd = np.concatenate([dist_parts[p, q, : cnt_parts[p, q]] for p in range(parts)])
        order = np.lexsort((ii, dd))[:k]
        o_cnt[q] = len(order)
        o_idx[q, : len(order)] = ii[order]
        o_dist[q, : len(order)] = dd[order]
    return o_idx, o_dist, o_cnt
