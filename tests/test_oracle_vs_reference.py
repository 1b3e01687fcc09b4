"""oracle/hnsw_oracle.c pinned against the unmodified reference compiled into oracle/_ref
(skipped where /root/reference was absent at build time; the golden replay covers that case)."""
import numpy as np
import pytest

from tests.conftest import assert_results_match, lcg_vectors, uniform

L2, COS, IP = 0, 1, 2


# ── distances: known answers of test/test_vec_math.c:8-61 + random bit-exactness ──
@pytest.mark.parametrize("metric,a,b,want", [
    (L2, [1, 2, 3], [1, 2, 3], 0.0), (L2, [1, 0, 0], [0, 1, 0], 2.0), (L2, [3], [7], 16.0),
    (COS, [1, 2, 3], [1, 2, 3], 0.0), (COS, [1, 0, 0], [0, 1, 0], 1.0), (COS, [1, 0, 0], [-1, 0, 0], 2.0),
    (COS, [0, 0, 0], [1, 2, 3], 1.0), (IP, [1, 2, 3], [4, 5, 6], -32.0), (IP, [1, 0, 0], [0, 1, 0], 0.0)])
def test_distance_known_answers(oracle, metric, a, b, want):
    eps = 1e-7 if metric == L2 else 1e-6
    assert abs(float(oracle.distance(metric, np.float32(a), np.float32(b))) - want) <= eps


@pytest.mark.parametrize("dim", [1, 2, 3, 4, 5, 7, 8, 31, 96, 128, 131, 768])
@pytest.mark.parametrize("metric", [L2, COS, IP])
def test_distance_bit_exact(oracle, reference, metric, dim):
    a, b = uniform(dim, 64, dim), uniform(dim + 1000, 64, dim)
    for i in range(64):
        assert oracle.distance(metric, a[i], b[i]).tobytes() == reference.distance(metric, a[i], b[i]).tobytes()


def test_metric_names(oracle, reference):
    for name in ("l2", "cosine", "inner_product", "hamming", ""):
        assert oracle.parse_metric(name) == reference.parse_metric(name)


# ── search: ids, distances and distance-call counts identical, fresh and reloaded state ──
@pytest.mark.parametrize("metric,M,dim,n", [(L2, 8, 16, 1200), (COS, 16, 24, 800), (IP, 6, 10, 600), (L2, 16, 128, 700)])
def test_search_identical(oracle, reference, metric, M, dim, n):
    base = uniform(n, n, dim)
    if metric == IP:
        base /= np.linalg.norm(base, axis=1, keepdims=True)
    q = uniform(n + 1, 100, dim)
    ri = reference.build(base, metric=metric, M=M, efc=60)
    for reloaded in (False, True):
        if reloaded:
            ri.sort_lists()
        g = ri.export()
        og = oracle.graph(g)
        for k, ef in ((1, 1), (10, 10), (10, 20), (10, 64), (5, 200)):
            r_ids, r_d, r_c, r_nd = ri.search(q, k, ef)
            ids, d, c, st, _ = og.search(q, k, ef)
            assert np.array_equal(c, r_c) and np.array_equal(ids, r_ids) and np.array_equal(d, r_d)
            assert np.array_equal(st["n_dist"], r_nd)
            ids2, d2, c2, st2, _ = og.search(q, k, ef, canonical=True)
            assert_results_match(ids, d, c, ids2, d2, c2, "canonical")
            assert np.array_equal(st["n_dist"], st2["n_dist"])


def test_reference_recall_case(oracle, reference):
    """test/test_hnsw_algo.c:96-145: 50×8 LCG data, M=8, efc=50, ef=64, k=5, recall >= 4/5."""
    base = lcg_vectors(50, 8)
    ri = reference.build(base, metric=L2, M=8, efc=50)
    g = ri.export()
    q = base[7:8] + np.float32(0.01)
    ids, d, c, _, _ = oracle.graph(g).search(q, 5, 64)
    truth, _, _, _ = oracle.exact_knn(base, q, 5, L2)
    assert len(set(ids[0]) & set(truth[0])) >= 4


def test_delete_and_empty(oracle, reference):
    """test/test_hnsw_algo.c:147-180: a deleted id is never returned; an empty index gives 0."""
    base = uniform(5, 300, 12)
    ri = reference.build(base, metric=L2, M=8, efc=60)
    for rid in range(5, 300, 11):
        assert ri.delete(rid) == 0
    ri.sort_lists()
    g = ri.export()
    q = uniform(6, 64, 12)
    r = ri.search(q, 5, 50)
    o = oracle.graph(g).search(q, 5, 50)
    assert np.array_equal(o[0], r[0]) and np.array_equal(o[1], r[1]) and np.array_equal(o[2], r[2])
    assert not (set(o[0].ravel()) & set(range(5, 300, 11)))
    from sqlite_muninn_b200.graph import FlatGraph
    e = FlatGraph(4, 0, 8, 50, np.empty(0, np.int64), np.empty(0, np.int32), np.empty(0, np.uint8),
                  np.empty((0, 4), np.float32), np.zeros(1, np.int64), np.empty(0, np.int32), -1, -1)
    ids, d, c, _, _ = oracle.graph(e).search(np.zeros((2, 4), np.float32), 3, 10)
    assert (c == 0).all() and (ids == -1).all()


def test_heap_restatement(oracle):
    """priority_queue.c order: ascending pops (test/test_priority_queue.c:15-38, :54-70)."""
    rng = np.random.default_rng(3)
    keys = rng.integers(0, 50, 400).astype(np.float32)
    ops = np.ones(400, np.int32)
    slots, ks = oracle.heap_sequence(keys, ops)
    assert np.array_equal(ks, np.sort(keys)) and sorted(slots) == list(range(400))


def test_exact_knn_rule(oracle, reference):
    """test/test_hnsw_algo.c:10-34: ascending distance, ties to the lower index."""
    base = np.round(uniform(9, 500, 6) * 2).astype(np.float32)  # many ties
    q = np.round(uniform(10, 20, 6) * 2).astype(np.float32)
    idx, d, c, _ = oracle.exact_knn(base, q, 12, L2)
    for i in range(20):
        dist = np.array([float(reference.distance(L2, q[i], base[j])) for j in range(500)], np.float32)
        order = np.lexsort((np.arange(500), dist))[:12]
        assert np.array_equal(idx[i], order) and np.array_equal(d[i], dist[order])
