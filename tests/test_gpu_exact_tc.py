"""`-m gpu`: exact k-NN through the tensor-core filter (tcgen05 3xTF32 / 1xTF32, exact_tc.cuh) is
bit-identical to the CPU oracle and to the CUDA-core path — ids, distances and counts — including
deleted rows, ragged sizes, dims that are not multiples of 4 or 32, tie-heavy data that overflows the
survivor lists, and zero vectors; and the filter's error stays well inside the margin its soundness
argument assumes (test/test_hnsw_algo.c:10-34 of the reference)."""
import numpy as np
import pytest

from tests.conftest import uniform

pytestmark = pytest.mark.gpu
L2, COS, IP = 0, 1, 2


@pytest.fixture(scope="module")
def cuda():
    from sqlite_muninn_b200 import cuda as c
    assert c.device_count() > 0, "no CUDA device: the product has no CPU path"
    return c


def flat(base, metric, deleted=None):
    from sqlite_muninn_b200.graph import FlatGraph
    n, dim = base.shape
    deleted = np.zeros(n, np.uint8) if deleted is None else deleted
    return FlatGraph(dim, metric, 8, 10, np.arange(n, dtype=np.int64), np.zeros(n, np.int32), deleted, base,
                     np.zeros(n + 1, np.int64), np.empty(0, np.int32), 0 if n else -1, 0 if n else -1)


def latent(seed, n, dim, rank=8):
    """low intrinsic dimension: distances spread out, so the filter has something to prune"""
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((rank, dim)).astype(np.float32) / np.float32(np.sqrt(rank))
    return (rng.standard_normal((n, rank)).astype(np.float32) @ A
            + np.float32(0.05) * rng.standard_normal((n, dim)).astype(np.float32)).astype(np.float32)


CASES = [  # metric, dim, n, nq, k
    (L2, 128, 20000, 300, 10), (L2, 128, 9000, 130, 100), (COS, 768, 6000, 129, 10), (IP, 96, 12000, 257, 10),
    (COS, 24, 5000, 64, 7), (L2, 3, 3000, 33, 5), (IP, 10, 2500, 1, 3), (L2, 131, 4100, 200, 20),
    (L2, 100, 1500, 128, 300), (COS, 33, 1024, 5, 1), (L2, 16, 700, 40, 10), (L2, 64, 30000, 1100, 10)]


@pytest.mark.parametrize("metric,dim,n,nq,k", CASES)
def test_tensor_modes_equal_oracle(cuda, oracle, metric, dim, n, nq, k):
    base = latent(n + dim, n, dim) if dim >= 16 else uniform(n + dim, n, dim)
    q = latent(n + dim, nq + n, dim)[n:] if dim >= 16 else uniform(n + dim + 1, nq, dim)
    if metric == IP:
        base /= np.linalg.norm(base, axis=1, keepdims=True)
    deleted = np.zeros(n, np.uint8)
    deleted[::13] = 1
    i_o, d_o, c_o, _ = oracle.exact_knn(base, q, k, metric, deleted=deleted, threads=8)
    with cuda.HnswIndexCuda.from_graph(flat(base, metric, deleted)) as idx:
        for mode, tile in ((cuda.EXACT_TC_3XTF32, 256), (cuda.EXACT_TC_3XTF32, 128), (cuda.EXACT_TC_1XTF32, 256),
                           (cuda.EXACT_TC_1XTF32, 128), (cuda.EXACT_TC_AUTO, 0), (cuda.EXACT_CUDA_CORE, 0)):
            idx.set_exact_mode(mode, tile)
            i_g, d_g, c_g = idx.exact_knn(q, k, index_base=7)
            st = idx.exact_stats()
            assert np.array_equal(c_g, c_o), (mode, tile)
            assert d_g.tobytes() == d_o.tobytes(), (mode, tile)
            assert np.array_equal(i_g, np.where(i_o >= 0, i_o + 7, -1)), (mode, tile)
            # AUTO = one TF32 product off the raw slab at every batch size (query-stationary filter from 1024 queries)
            want_mode = mode if mode != cuda.EXACT_TC_AUTO else cuda.EXACT_TC_1XTF32
            assert st["mode"] == want_mode
            if mode != cuda.EXACT_CUDA_CORE:
                # counters are live: every query re-ranks at least its own top-k (deleted rows pass the filter too)
                assert st["chunks"] >= 1 and st["candidates"] >= nq * min(k, int((deleted == 0).sum()))
                assert st["filter_ms"] > 0.0


QS_CASES = [  # metric, dim, n, nq, k, super-block MB: the query-stationary filter (nq >= 1024, one TF32 product)
    (L2, 128, 40000, 2500, 10, 1), (L2, 128, 9000, 1024, 100, 32), (IP, 96, 30000, 1300, 10, 1), (COS, 131, 12000, 1100, 10, 2),
    (L2, 256, 8000, 1200, 5, 1), (L2, 300, 6000, 1100, 10, 1), (L2, 16, 70000, 4000, 10, 1), (COS, 768, 3000, 1030, 10, 32)]


@pytest.mark.parametrize("metric,dim,n,nq,k,sb_mb", QS_CASES)
def test_query_stationary_filter_equals_oracle(cuda, oracle, metric, dim, n, nq, k, sb_mb):
    """Large batches: query tile resident in shared memory, row tiles swept in super-blocks (forced small here so that a
    launch crosses many of them), CTAs switching query tiles mid-run; dims with 1..8 k-blocks (beyond 8 the streaming
    kernel takes over). Bit-exact against the oracle, and identical to the streaming filter's survivors' result."""
    base = latent(n + dim, n, dim)
    q = latent(n + dim, nq + n, dim)[n:]
    if metric == IP:
        base /= np.linalg.norm(base, axis=1, keepdims=True)
    deleted = np.zeros(n, np.uint8)
    deleted[::29] = 1
    i_o, d_o, c_o, _ = oracle.exact_knn(base, q, k, metric, deleted=deleted, threads=8)
    with cuda.HnswIndexCuda.from_graph(flat(base, metric, deleted)) as idx:
        idx.set_option("exact_superblock_mb", sb_mb)
        for qs, pair in ((1, 0), (1, 1), (0, 0)):  # query-stationary, the same on CTA pairs (cta_group::2, M = 256), streaming
            idx.set_option("exact_qs", qs)
            idx.set_option("exact_pair", pair)
            idx.set_exact_mode(cuda.EXACT_TC_AUTO, 0)
            i_g, d_g, c_g = idx.exact_knn(q, k)
            st = idx.exact_stats()
            assert st["mode"] == cuda.EXACT_TC_1XTF32 and st["overflow_queries"] == 0
            assert np.array_equal(c_g, c_o) and d_g.tobytes() == d_o.tobytes() and np.array_equal(i_g, i_o), (qs, pair)


def test_filter_prunes(cuda, oracle):
    """On spread-out data the 3xTF32 filter leaves ≈ (growth factor)·k survivors per chunk, not n. A batch of
    256 queries runs the small-batch policy: 2048-row first chunk (all survive), ×32 growth."""
    n, dim, nq, k = 60000, 128, 256, 10
    base, q = latent(1, n, dim), latent(2, nq, dim)
    with cuda.HnswIndexCuda.from_graph(flat(base, L2)) as idx:
        idx.set_exact_mode(cuda.EXACT_TC_3XTF32, 256)
        i_g, d_g, c_g = idx.exact_knn(q, k)
        st = idx.exact_stats()
    i_o, d_o, c_o, _ = oracle.exact_knn(base, q, k, L2, threads=8)
    assert np.array_equal(i_g, i_o) and d_g.tobytes() == d_o.tobytes()
    assert st["overflow_queries"] == 0
    per_query = st["candidates"] / nq
    assert per_query < 2048 + 60 * k * st["chunks"], per_query  # the first chunk passes whole
    assert per_query < 0.05 * n


def test_tie_heavy_data_overflows_and_stays_exact(cuda, oracle):
    """Duplicated rows: every copy of a top-k row ties with τ and must survive; the survivor lists
    overflow, the flagged queries are recomputed on the CUDA cores, ties still go to the lower index."""
    rng = np.random.default_rng(5)
    proto = np.round(rng.uniform(-2, 2, (6, 8))).astype(np.float32)
    base = proto[rng.integers(0, 6, 30000)]
    q = np.concatenate([proto[:3] + np.float32(0.25), latent(9, 20, 8)]).astype(np.float32)
    i_o, d_o, c_o, _ = oracle.exact_knn(base, q, 10, L2, threads=8)
    with cuda.HnswIndexCuda.from_graph(flat(base, L2)) as idx:
        for mode in (cuda.EXACT_TC_3XTF32, cuda.EXACT_TC_1XTF32):
            idx.set_exact_mode(mode, 256)
            i_g, d_g, c_g = idx.exact_knn(q, 10)
            st = idx.exact_stats()
            assert st["overflow_queries"] > 0
            assert np.array_equal(c_g, c_o) and d_g.tobytes() == d_o.tobytes() and np.array_equal(i_g, i_o), mode


def test_zero_vectors_cosine(cuda, oracle):
    n, dim = 3000, 20
    base = uniform(3, n, dim)
    base[::7] = 0.0
    q = uniform(4, 50, dim)
    q[3] = 0.0
    i_o, d_o, c_o, _ = oracle.exact_knn(base, q, 10, COS, threads=4)
    with cuda.HnswIndexCuda.from_graph(flat(base, COS)) as idx:
        idx.set_exact_mode(cuda.EXACT_TC_3XTF32, 256)
        i_g, d_g, c_g = idx.exact_knn(q, 10)
    assert np.array_equal(c_g, c_o) and d_g.tobytes() == d_o.tobytes() and np.array_equal(i_g, i_o)


def test_fewer_rows_than_k_and_empty(cuda, oracle):
    base, q = uniform(1, 5, 12), uniform(2, 9, 12)
    i_o, d_o, c_o, _ = oracle.exact_knn(base, q, 8, L2)
    with cuda.HnswIndexCuda.from_graph(flat(base, L2)) as idx:
        i_g, d_g, c_g = idx.exact_knn(q, 8)
    assert np.array_equal(c_g, c_o) and np.array_equal(i_g, i_o) and d_g.tobytes() == d_o.tobytes()
    with cuda.HnswIndexCuda.from_graph(flat(np.empty((0, 12), np.float32), L2)) as idx:
        i_g, d_g, c_g = idx.exact_knn(q, 8)
    assert (c_g == 0).all() and (i_g == -1).all()


@pytest.mark.parametrize("metric,dim", [(L2, 128), (L2, 768), (COS, 768), (IP, 96), (COS, 100), (L2, 5)])
@pytest.mark.parametrize("mode", [1, 2])
def test_filter_error_is_inside_the_margin(cuda, metric, dim, mode):
    """|d̃ − d| of the tensor-core contraction against float64, as a fraction of the margin ε the
    filter adds to τ (the soundness argument needs < 1). ε carries a factor 2 over the worst case by
    construction, so 1/2 is the ceiling; 3xTF32 is dominated by the (pessimistic) accumulation term
    and must stay below 1/10, 1xTF32 by the input truncation itself and may approach 1/2."""
    n, nq = 2048, 200
    base = uniform(dim, n, dim) * np.float32(3.0)
    q = uniform(dim + 1, nq, dim) * np.float32(3.0)
    with cuda.HnswIndexCuda.from_graph(flat(base, metric)) as idx:
        idx.set_exact_mode(mode, 256)
        s_tc = idx.exact_filter_probe(q, n).astype(np.float64)  # the tensor-core contraction Σ q·x
        c_eps = idx.exact_stats()["c_eps"]
    b64, q64 = base.astype(np.float64), q.astype(np.float64)
    dot = q64 @ b64.T
    qn, xn = (q64 * q64).sum(1)[:, None], (b64 * b64).sum(1)[None, :]
    if metric == L2:  # d̃ = qn + xn − 2s
        err, eps = 2 * np.abs(s_tc - dot), c_eps * (qn + xn)
    elif metric == IP:  # d̃ = −s
        err, eps = np.abs(s_tc - dot), c_eps * (qn + xn)
    else:  # d̃ = 1 − s/(|q||x|)
        err, eps = np.abs(s_tc - dot) / np.sqrt(qn * xn), c_eps * np.ones_like(dot)
    ratio = float((err / eps).max())
    print(f"metric={metric} dim={dim} mode={mode}: max |err|/eps = {ratio:.4f} (c_eps={c_eps:.3e})")
    assert ratio < (0.1 if mode == 1 else 0.5)


def _margin_ratio(cuda, base, q, metric, mode):
    n = base.shape[0]
    with cuda.HnswIndexCuda.from_graph(flat(base, metric)) as idx:
        idx.set_exact_mode(mode, 256)
        s_tc = idx.exact_filter_probe(q, n).astype(np.float64)
        c_eps = idx.exact_stats()["c_eps"]
    b64, q64 = base.astype(np.float64), q.astype(np.float64)
    dot = q64 @ b64.T
    qn, xn = (q64 * q64).sum(1)[:, None], (b64 * b64).sum(1)[None, :]
    if metric == COS:
        return float((np.abs(s_tc - dot) / np.sqrt(qn * xn) / c_eps).max())
    return float(((2 if metric == L2 else 1) * np.abs(s_tc - dot) / (c_eps * (qn + xn))).max())


@pytest.mark.parametrize("metric", [L2, IP, COS])
@pytest.mark.parametrize("mode", [1, 2])
@pytest.mark.parametrize("kind", ["low_bits_set", "mixed_magnitudes", "cancellation"])
def test_filter_margin_adversarial(cuda, metric, mode, kind):
    """Inputs built against the margin (DESIGN.md §5.2 derives it): (a) every one of the 13 mantissa bits the TF32
    input path drops is SET, all signs equal and |q_i| = |x_i|, so that every truncation error has the same sign and
    the AM-GM step of the bound is tight — the worst case of the single product, where |err|/eps approaches but must
    stay below 1/2; (b) magnitudes from 1e-3 to 1e3 inside one vector; (c) q·x cancelling to ~0 with large terms."""
    dim, n, nq = 128, 1024, 64
    rng = np.random.default_rng(42)
    if kind == "low_bits_set":
        m = rng.uniform(1.0, 1.002, (n + nq, dim)).astype(np.float32)  # mantissa just above 1.0: the dropped bits weigh most
        v = (m.view(np.uint32) | np.uint32(0x1FFF)).view(np.float32)
        base, q = v[:n].copy(), v[n:].copy()
        q[:] = base[:nq]  # |q_i| = |x_i| for the matching pairs
    elif kind == "mixed_magnitudes":
        e = rng.integers(-10, 11, (n + nq, dim))
        v = (rng.uniform(1, 2, (n + nq, dim)) * (2.0 ** e) * rng.choice([-1, 1], (n + nq, dim))).astype(np.float32)
        v = (v.view(np.uint32) | np.uint32(0x1FFF)).view(np.float32)
        base, q = v[:n].copy(), v[n:].copy()
    else:
        a = rng.uniform(50, 100, (n + nq, dim)).astype(np.float32)
        a[:, 1::2] = -a[:, 0::2]  # pairs (+a, -a) against a same-sign partner → Σ q·x ≈ 0 out of terms of 1e4
        base = np.abs(a[:n]) * np.float32(1.0)
        q = a[n:]
        base = (base.view(np.uint32) | np.uint32(0x1FFF)).view(np.float32)
        q = (q.view(np.uint32) | np.uint32(0x1FFF)).view(np.float32)
    ratio = _margin_ratio(cuda, base, q, metric, mode)
    print(f"{kind} metric={metric} mode={mode}: max |err|/eps = {ratio:.4f}")
    assert ratio < (0.2 if mode == 1 else 0.55)


@pytest.mark.parametrize("metric", [L2, IP, COS])
def test_single_product_stays_exact_on_adversarial_rows(cuda, oracle, metric):
    """End to end on the worst-case rows above (all dropped bits set, near-duplicates a few ulps apart so that the
    filter's margin decides which rows reach the exact re-rank): ids and distances still bit-exact."""
    dim, n, nq, k = 128, 20000, 200, 10
    rng = np.random.default_rng(7)
    centers = rng.uniform(1.0, 1.002, (50, dim)).astype(np.float32)
    base = centers[rng.integers(0, 50, n)] + rng.uniform(0, 3e-4, (n, dim)).astype(np.float32)
    base = (base.view(np.uint32) | np.uint32(0x1FFF)).view(np.float32)
    q = (centers[rng.integers(0, 50, nq)].view(np.uint32) | np.uint32(0x1FFF)).view(np.float32)
    i_o, d_o, c_o, _ = oracle.exact_knn(base, q, k, metric, threads=8)
    with cuda.HnswIndexCuda.from_graph(flat(base, metric)) as idx:
        for mode in (cuda.EXACT_TC_1XTF32, cuda.EXACT_TC_3XTF32):
            idx.set_exact_mode(mode, 256)
            i_g, d_g, c_g = idx.exact_knn(q, k)
            assert np.array_equal(c_g, c_o) and d_g.tobytes() == d_o.tobytes() and np.array_equal(i_g, i_o), mode


def test_rows_only_handle_from_host_and_device_memory(cuda, oracle):
    """muninn_cuda_index_upload_rows: a shard of base rows without a graph (BASELINE config 5), adopted from host OR
    device memory; exact k-NN equals the oracle, search and patch refuse the handle."""
    import torch
    n, dim, nq, k = 30000, 128, 1100, 20
    base, q = latent(71, n, dim), latent(72, nq, dim)
    i_o, d_o, c_o, _ = oracle.exact_knn(base, q, k, L2, threads=8)
    for where in ("host", "device"):
        with cuda.HnswIndexCuda(dim, L2, 8) as idx:
            if where == "host":
                idx.upload_rows(base.ctypes.data, n)
            else:
                d_rows = torch.from_numpy(base).cuda()
                torch.cuda.synchronize()
                idx.upload_rows(d_rows.data_ptr(), n)
                del d_rows  # the handle owns its copy
            i_g, d_g, c_g = idx.exact_knn(q, k, index_base=5)
            assert np.array_equal(c_g, c_o) and d_g.tobytes() == d_o.tobytes() and np.array_equal(i_g, i_o + 5), where
            with pytest.raises(cuda.MuninnCudaError, match="rows only"):
                idx.search(q[:4], 5, 10)
