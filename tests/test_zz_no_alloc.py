"""`-m gpu`: muninn_cuda_search_batch_device_static — the search entry point that allocates nothing (its scratch is a
file-scope device array inside libmuninn_cuda; include/muninn_cuda.h). It has no counterpart in the reference,
so its oracle is the existing entry point: ids, distances and counts must be
bit-identical to muninn_cuda_search_batch_device on the same inputs, on calls in a row that share the array.

Collected last on purpose: nothing in this file may hide another test's result. Runs alone in a
fresh process (the static calls come BEFORE any call of the allocating entry point) and after the other tests."""
import numpy as np
import pytest

from tests.conftest import uniform

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cuda():
    from sqlite_muninn_b200 import cuda as c
    assert c.device_count() > 0, "no CUDA device: the product has no CPU path"
    return c


@pytest.fixture(scope="module")
def graph():
    from sqlite_muninn_b200.builder import build_graph
    base = uniform(7001, 20_000, 128)
    return build_graph(base, 0, 16, 48, batch=1024, warm=2048)


def _buffers(torch, nq, k):
    dev = torch.device("cuda:0")
    return (torch.full((nq, k), -7, dtype=torch.int64, device=dev), torch.full((nq, k), -7.0, dtype=torch.float32, device=dev),
            torch.full((nq,), -7, dtype=torch.int32, device=dev))


def _run(torch, idx, fn, d_q, k, ef):
    nq = d_q.shape[0]
    ids, dist, cnt = _buffers(torch, nq, k)
    fn(d_q.data_ptr(), nq, k, ef, ids.data_ptr(), dist.data_ptr(), cnt.data_ptr(), None)
    torch.cuda.synchronize()
    return ids.cpu().numpy(), dist.cpu().numpy(), cnt.cpu().numpy()


@pytest.mark.parametrize("vis_factor", [2.0, 0.5])  # 0.5: first-level visited tables overflow → the retry lists in the array are used
def test_static_scratch_search_equals_allocating_entry_point(cuda, graph, vis_factor):
    import torch
    lim = cuda.static_scratch_limits()
    assert lim == {"max_nq": 65536, "max_nodes": 2097152 - 256, "bytes": 256 + 2 * 65536 * 4 + (4 << 20) * 4}
    k = 10
    # three calls in a row on the shared array: a full batch (one warp per query), a small one (query teams of several
    # warps, fewer teams, other table sizes), and the first again — stale counters, lists or tables would show here
    shapes = [(5000, 64, 11), (100, 32, 12), (5000, 64, 11)]
    qs = [torch.from_numpy(uniform(seed, nq, 128)).to("cuda:0") for nq, _, seed in shapes]
    torch.cuda.synchronize()
    with cuda.HnswIndexCuda.from_graph(graph) as idx:
        idx.set_option("vis_factor", vis_factor)
        got = []
        for i, ((nq, ef, _), d_q) in enumerate(zip(shapes, qs)):
            before = torch.cuda.mem_get_info()[0]
            got.append(_run(torch, idx, idx.search_device_static, d_q, k, ef))
            if i == 2:  # same shape as call 0, kernels already loaded: the call itself must not take device memory
                ids, dist, cnt = _buffers(torch, nq, k)
                torch.cuda.synchronize()
                free0 = torch.cuda.mem_get_info()[0]
                idx.search_device_static(d_q.data_ptr(), nq, k, ef, ids.data_ptr(), dist.data_ptr(), cnt.data_ptr(), None)
                torch.cuda.synchronize()
                assert torch.cuda.mem_get_info()[0] == free0, (before, free0)
                assert np.array_equal(ids.cpu().numpy(), got[0][0])
        want = [_run(torch, idx, idx.search_device, d_q, k, ef) for (nq, ef, _), d_q in zip(shapes, qs)]
        for i, (g, w) in enumerate(zip(got, want)):
            assert np.array_equal(g[0], w[0]), f"ids differ, call {i}"
            assert g[1].tobytes() == w[1].tobytes(), f"distances differ, call {i}"
            assert np.array_equal(g[2], w[2]), f"counts differ, call {i}"
            assert (g[2] == k).all() and (np.diff(g[1], axis=1) >= 0).all()
        assert np.array_equal(got[0][0], got[2][0]) and got[0][1].tobytes() == got[2][1].tobytes()


def test_static_scratch_limits_are_enforced(cuda, graph):
    import torch
    with cuda.HnswIndexCuda.from_graph(graph) as idx:
        nq = 65537
        d_q = torch.zeros((nq, 128), dtype=torch.float32, device="cuda:0")
        ids, dist, cnt = _buffers(torch, nq, 10)
        with pytest.raises(cuda.MuninnCudaError, match="static scratch"):
            idx.search_device_static(d_q.data_ptr(), nq, 10, 64, ids.data_ptr(), dist.data_ptr(), cnt.data_ptr(), None)
        # the handle is still usable, and an empty batch is a no-op as in the allocating entry point
        idx.search_device_static(d_q.data_ptr(), 0, 10, 64, ids.data_ptr(), dist.data_ptr(), cnt.data_ptr(), None)
        g = _run(torch, idx, idx.search_device_static, d_q[:64], 10, 64)
        w = _run(torch, idx, idx.search_device, d_q[:64], 10, 64)
        assert np.array_equal(g[0], w[0]) and g[1].tobytes() == w[1].tobytes() and np.array_equal(g[2], w[2])
