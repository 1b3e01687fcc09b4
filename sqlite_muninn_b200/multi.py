"""Multi-GPU plumbing: one process per GPU, torch.distributed for the rendezvous.

* HNSW search shards by QUERY: the graph is replicated, rank r takes queries
  [r·nq/W, (r+1)·nq/W); there is no data-path collective. `gather_search_results` only exists to
  hand the concatenated answer to a caller that wants it on every rank.
* Exact k-NN shards the BASE rows: every rank scans its rows for all queries, the per-shard
  [nq × k] lists are packed into 64-bit keys (ordered distance << 32 | global row index, unused
  tail = all ones) and exchanged with ONE all-gather of nq·k·8 bytes per rank, then merged under
  (distance, row index). On GPUs all of it happens inside the C ABI
  (muninn_cuda_exact_knn_sharded_device: scan + ncclAllGather + merge on the handle's stream);
  `ShardedExactKnn` only does the rendezvous. The host statements of the packing and the merge
  below exist for the gloo CPU tests of the gathered layout.
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


KEY_NONE = np.uint64(0xFFFFFFFFFFFFFFFF)


def pack_keys_host(idx: np.ndarray, dist: np.ndarray, cnt: np.ndarray) -> np.ndarray:
    """Host statement of topk_pack_keys (csrc/exact_knn.cuh): ordered-distance << 32 | index, tail = all ones.
    The float → uint32 map is order preserving with -0 folded into +0 (csrc/muninn_dev.cuh okey)."""
    d = (np.asarray(dist, np.float32) + np.float32(0.0)).view(np.uint32)
    ok = np.where(d & np.uint32(0x80000000), ~d, d | np.uint32(0x80000000)).astype(np.uint64)
    keys = (ok << np.uint64(32)) | (np.asarray(idx, np.int64).astype(np.uint64) & np.uint64(0xFFFFFFFF))
    k = keys.shape[1]
    keys[np.arange(k)[None, :] >= np.asarray(cnt)[:, None]] = KEY_NONE
    return keys


def unpack_keys_host(keys: np.ndarray):
    """Inverse of pack_keys_host → (idx int64 with -1 tail, dist float32 with 0 tail, cnt int32)."""
    valid = keys != KEY_NONE
    ok = (keys >> np.uint64(32)).astype(np.uint32)
    bits = np.where(ok & np.uint32(0x80000000), ok & np.uint32(0x7FFFFFFF), ~ok).astype(np.uint32)
    dist = np.where(valid, bits.view(np.float32), np.float32(0.0)).astype(np.float32)
    idx = np.where(valid, (keys & np.uint64(0xFFFFFFFF)).astype(np.int64), np.int64(-1))
    return idx, dist, valid.sum(axis=1).astype(np.int32)


def allgather_topk_keys(keys, group=None):
    """THE collective of the sharded exact k-NN: one all-gather of the [nq × k] key array → [W × nq × k].
    `keys` is a torch tensor of int64 (the uint64 keys reinterpreted; gloo and NCCL both move int64)."""
    import torch
    import torch.distributed as dist_

    world = dist_.get_world_size(group)
    buf = torch.empty((world,) + tuple(keys.shape), dtype=keys.dtype, device=keys.device)
    dist_.all_gather(list(buf.unbind(0)), keys.contiguous(), group=group)
    return buf


def merge_topk_keys_reference(keys_parts: np.ndarray, k: int):
    """Host statement of topk_merge_keys: the k smallest keys over all parts, per query."""
    parts, nq, _ = keys_parts.shape
    allk = np.sort(np.transpose(keys_parts, (1, 0, 2)).reshape(nq, -1), axis=1)[:, :k]
    return unpack_keys_host(allk)


class ShardedExactKnn:
    """Rendezvous for muninn_cuda_exact_knn_sharded_device under torch.distributed: rank 0 draws an ncclUniqueId through
    the C ABI, torch.distributed ships the 128 bytes, every rank joins its own communicator (so the library's
    collective never interleaves with torch's on one communicator). `use_torch_comm=True` hands the library torch's
    ncclComm_t instead (ProcessGroupNCCL._comm_ptr)."""

    def __init__(self, idx, device: int, use_torch_comm: bool = False, group=None):
        import torch
        import torch.distributed as dist_
        from . import cuda as C_
        self.idx, self.own = idx, not use_torch_comm
        self.world, self.rank = dist_.get_world_size(group), dist_.get_rank(group)
        if use_torch_comm:
            pg = group or dist_.distributed_c10d._get_default_group()
            self.comm = pg._get_backend(torch.device("cuda", device))._comm_ptr()
        else:
            box = [C_.comm_unique_id() if self.rank == 0 else None]
            dist_.broadcast_object_list(box, src=0, group=group)
            self.comm = C_.comm_create(device, self.world, self.rank, box[0])

    def __call__(self, q_ptr: int, nq: int, k: int, index_base: int, idx_ptr: int, dist_ptr: int, cnt_ptr: int):
        self.idx.exact_knn_sharded_device(self.comm, q_ptr, nq, k, index_base, idx_ptr, dist_ptr, cnt_ptr)

    def close(self):
        if self.own and self.comm:
            from . import cuda as C_
            C_.comm_destroy(self.comm)
        self.comm = None


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
