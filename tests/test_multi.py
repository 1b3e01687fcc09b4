"""N>1 host logic on CPU: world_size-2 gloo processes (query sharding for search, the gathered
layout of the sharded exact k-NN), plus the 2-shard merge on one GPU (`-m gpu`)."""
import os
import socket

import numpy as np
import pytest

from sqlite_muninn_b200.multi import merge_topk_parts_reference, shard_range
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
    from sqlite_muninn_b200.multi import allgather_topk_parts, gather_search_results
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
    open(os.path.join(tmp, f"ok{rank}"), "w").write(f"{int(ok1)}{int(ok2)}")
    dist.destroy_process_group()


def test_two_rank_gloo(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert [open(tmp_path / f"ok{r}").read() for r in range(2)] == ["11", "11"]


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
    for h in handles:
        h.close()  # destroy synchronises the handle's stream
    t_idx, t_d, t_c, _ = oracle.exact_knn(base, q, k, 0)
    assert np.array_equal(oi.cpu().numpy(), t_idx) and np.array_equal(od.cpu().numpy(), t_d)
    assert np.array_equal(oc.cpu().numpy(), t_c)
