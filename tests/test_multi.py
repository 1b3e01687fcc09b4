"""N>1 host logic on CPU: world_size-2 gloo processes (query sharding for search, the gathered
layout of the sharded exact k-NN), plus the 2-shard merge on one GPU (`-m gpu`)."""
import os
import socket

import numpy as np
import pytest

from sqlite_muninn_b200.multi import (merge_topk_keys_reference, merge_topk_parts_reference, pack_keys_host, shard_range,
                                      unpack_keys_host)
from tests.conftest import uniform


def test_shard_range_covers_everything():
    for total in (0, 1, 7, 10, 10000, 10001):
        for world in (1, 2, 3, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, tmp):
    import torch
    import torch.distributed as dist
    from oracle.bindings import Oracle
    from sqlite_muninn_b200.builder import build_graph
    from sqlite_muninn_b200.multi import allgather_topk_keys, allgather_topk_parts, gather_search_results
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = Oracle()
    base, q = uniform(1, 1200, 16), uniform(2, 101, 16)  # 101: uneven split
    g = build_graph(base, 0, 8, 40)
    # search: graph replicated, queries sharded, results concatenated in batch order
    lo, hi = shard_range(len(q), rank, world)
    ids, d, c, _, _ = orc.graph(g).search(q[lo:hi], 5, 40)
    a_ids, a_d, a_c = gather_search_results(torch.from_numpy(ids), torch.from_numpy(d), torch.from_numpy(c), len(q))
    f_ids, f_d, f_c, _, _ = orc.graph(g).search(q, 5, 40)
    ok1 = np.array_equal(a_ids.numpy(), f_ids) and np.array_equal(a_d.numpy(), f_d) and np.array_equal(a_c.numpy(), f_c)
    # exact k-NN: base sharded, one all-gather, merge under (distance, index)
    r0, r1 = shard_range(len(base), rank, world)
    idx, dd, cc, _ = orc.exact_knn(base[r0:r1], q, 7, 0)
    idx = np.where(idx >= 0, idx + r0, -1)
    p_idx, p_d, p_c = allgather_topk_parts(torch.from_numpy(idx), torch.from_numpy(dd), torch.from_numpy(cc))
    m_idx, m_d, m_c = merge_topk_parts_reference(p_idx.numpy(), p_d.numpy(), p_c.numpy(), 7)
    t_idx, t_d, t_c, _ = orc.exact_knn(base, q, 7, 0)
    ok2 = np.array_equal(m_idx, t_idx) and np.array_equal(m_d, t_d) and np.array_equal(m_c, t_c)
    # the product's exchange: ONE all-gather of 64-bit keys (nq·k·8 bytes per rank), merged by key order
    keys = torch.from_numpy(pack_keys_host(idx, dd, cc).view(np.int64))
    k_idx, k_d, k_c = merge_topk_keys_reference(allgather_topk_keys(keys).numpy().view(np.uint64), 7)
    ok2 = ok2 and np.array_equal(k_idx, t_idx) and np.array_equal(k_d, t_d) and np.array_equal(k_c, t_c)
    open(os.path.join(tmp, f"ok{rank}"), "w").write(f"{int(ok1)}{int(ok2)}")
    dist.destroy_process_group()


def test_two_rank_gloo(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert [open(tmp_path / f"ok{r}").read() for r in range(2)] == ["11", "11"]


def test_key_packing_round_trip():
    rng = np.random.default_rng(0)
    d = np.sort(rng.normal(size=(6, 9)).astype(np.float32), axis=1)
    d[0, :2] = [-0.0, 0.0]
    idx = rng.integers(0, 1 << 32, size=(6, 9)).astype(np.int64)
    cnt = np.array([9, 3, 0, 9, 5, 1], np.int32)
    keys = pack_keys_host(idx, d, cnt)
    i2, d2, c2 = unpack_keys_host(keys)
    m = np.arange(9)[None, :] < cnt[:, None]
    assert np.array_equal(c2, cnt) and np.array_equal(i2[m], idx[m]) and np.array_equal(d2[m], d[m] + np.float32(0))
    assert (i2[~m] == -1).all() and (d2[~m] == 0).all()
    hi = (keys >> np.uint64(32)).astype(np.int64)  # the distance half is order preserving, -0 == +0
    assert hi[0, 0] == hi[0, 1] and (np.diff(hi[3]) >= 0).all()


@pytest.mark.gpu
def test_sharded_exact_knn_merge_on_gpu(oracle):
    """Two row shards on one GPU (two handles), lists concatenated as an all-gather would, merged by
    muninn_cuda_topk_merge_device: bit-exact vs the oracle over the whole base."""
    import torch
    from sqlite_muninn_b200.cuda import HnswIndexCuda
    from sqlite_muninn_b200.graph import FlatGraph
    n, dim, k, nq = 5000, 32, 20, 150
    base = np.round(uniform(3, n, dim) * 4).astype(np.float32)  # ties across shards
    q = np.round(uniform(4, nq, dim) * 4).astype(np.float32)
    dq = torch.from_numpy(q).cuda()
    parts = []
    handles = []
    for r in range(2):
        lo, hi = shard_range(n, r, 2)
        g = FlatGraph(dim, 0, 8, 10, np.arange(hi - lo, dtype=np.int64), np.zeros(hi - lo, np.int32),
                      np.zeros(hi - lo, np.uint8), np.ascontiguousarray(base[lo:hi]), np.zeros(hi - lo + 1, np.int64),
                      np.empty(0, np.int32), 0, 0)
        h = HnswIndexCuda.from_graph(g)
        handles.append(h)
        i = torch.empty((nq, k), dtype=torch.int64, device="cuda")
        d = torch.empty((nq, k), dtype=torch.float32, device="cuda")
        c = torch.empty(nq, dtype=torch.int32, device="cuda")
        h.exact_knn_device(dq.data_ptr(), nq, k, lo, i.data_ptr(), d.data_ptr(), c.data_ptr())
        parts.append((i, d, c))
    for h in handles:
        h.last_kernel_ms()  # synchronises on the handle's own stream
    pi, pd, pc = (torch.stack([p[j] for p in parts]).contiguous() for j in range(3))
    oi = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    od = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    oc = torch.empty(nq, dtype=torch.int32, device="cuda")
    handles[0].topk_merge_device(2, nq, k, pi.data_ptr(), pd.data_ptr(), pc.data_ptr(), oi.data_ptr(), od.data_ptr(),
                                 oc.data_ptr())
    handles[0].last_kernel_ms()
    t_idx, t_d, t_c, _ = oracle.exact_knn(base, q, k, 0)
    assert np.array_equal(oi.cpu().numpy(), t_idx) and np.array_equal(od.cpu().numpy(), t_d)
    assert np.array_equal(oc.cpu().numpy(), t_c)
    # the key form of the same exchange (what muninn_cuda_exact_knn_sharded_device moves with its one all-gather)
    keys = torch.empty((2, nq, k), dtype=torch.int64, device="cuda")
    for r, (i, d, c) in enumerate(parts):
        handles[0].topk_pack_keys_device(nq, k, i.data_ptr(), d.data_ptr(), c.data_ptr(), keys[r].data_ptr())
    handles[0].topk_merge_keys_device(2, nq, k, keys.data_ptr(), oi.data_ptr(), od.data_ptr(), oc.data_ptr())
    torch.cuda.synchronize()
    handles[0].last_kernel_ms()
    for r, (i, d, c) in enumerate(parts):  # device packing == host statement of it
        assert np.array_equal(keys[r].cpu().numpy().view(np.uint64), pack_keys_host(i.cpu().numpy(), d.cpu().numpy(), c.cpu().numpy()))
    for h in handles:
        h.close()  # destroy synchronises the handle's stream
    assert np.array_equal(oi.cpu().numpy(), t_idx) and np.array_equal(od.cpu().numpy(), t_d)
    assert np.array_equal(oc.cpu().numpy(), t_c)


def _nccl_worker(rank, world, port, tmp, use_torch_comm):
    import torch
    import torch.distributed as dist
    from oracle.bindings import Oracle
    from sqlite_muninn_b200.cuda import HnswIndexCuda
    from sqlite_muninn_b200.graph import FlatGraph
    from sqlite_muninn_b200.multi import ShardedExactKnn
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    n, dim, k, nq = 30011, 48, 100, 257
    base = np.round(uniform(5, n, dim) * 4).astype(np.float32)  # ties across the shard boundary
    q = np.round(uniform(6, nq, dim) * 4).astype(np.float32)
    lo, hi = shard_range(n, rank, world)
    g = FlatGraph(dim, 0, 8, 10, np.arange(hi - lo, dtype=np.int64), np.zeros(hi - lo, np.int32), np.zeros(hi - lo, np.uint8),
                  np.ascontiguousarray(base[lo:hi]), np.zeros(hi - lo + 1, np.int64), np.empty(0, np.int32), 0, 0)
    ok = True
    with HnswIndexCuda.from_graph(g, device=rank) as h:
        st = torch.cuda.Stream()
        h.set_stream(st.cuda_stream)
        sk = ShardedExactKnn(h, rank, use_torch_comm=use_torch_comm)
        dq = torch.from_numpy(q).cuda()
        oi = torch.empty((nq, k), dtype=torch.int64, device="cuda")
        od = torch.empty((nq, k), dtype=torch.float32, device="cuda")
        oc = torch.empty(nq, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        for _ in range(2):  # twice: buffers and communicator are reused
            sk(dq.data_ptr(), nq, k, lo, oi.data_ptr(), od.data_ptr(), oc.data_ptr())
            st.synchronize()
            t_idx, t_d, t_c, _ = Oracle().exact_knn(base, q, k, 0)
            ok = ok and np.array_equal(oi.cpu().numpy(), t_idx) and np.array_equal(od.cpu().numpy(), t_d) \
                and np.array_equal(oc.cpu().numpy(), t_c)
        sk.close()
    open(os.path.join(tmp, f"nccl{rank}"), "w").write(str(int(ok)))
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("use_torch_comm", [False, True])
def test_sharded_exact_knn_two_nccl_ranks(tmp_path, use_torch_comm):
    """The product's multi-GPU exact k-NN as shipped: two processes, two GPUs, base rows sharded, ONE ncclAllGather of
    64-bit keys inside muninn_cuda_exact_knn_sharded_device, every rank ends with the oracle's answer over the whole
    base. Needs two visible GPUs (the single-GPU box skips it; `gpurun --gpus 2` runs it)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_nccl_worker, args=(2, _free_port(), str(tmp_path), use_torch_comm), nprocs=2, join=True)
    assert [open(tmp_path / f"nccl{r}").read() for r in range(2)] == ["1", "1"]
