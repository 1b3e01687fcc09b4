"""CPU: the pieces of bench.py the reported numbers rest on — the algorithmic-bytes formula of SURVEY.md §8(d),
recall as `benchmarks/harness/treatments/vss.py:96-102` computes it, the workload generators' determinism, and
that the oracle's per-query counters (what the GPU must reproduce) feed the formula."""
import numpy as np

import bench
from tests.conftest import uniform


def test_algorithmic_bytes_formula():
    stats = np.zeros(3, dtype=[("n_dist", np.int64), ("n_expand", np.int32), ("n_greedy_steps", np.int32)])
    stats["n_dist"] = [1650, 1600, 10]
    stats["n_expand"] = [66, 64, 1]
    stats["n_greedy_steps"] = [5, 4, 0]
    dim, M, k = 128, 16, 10
    want = sum(nd * dim * 4 + ne * 2 * M * 4 + ng * M * 4 + dim * 4 + k * 12
               for nd, ne, ng in zip(stats["n_dist"], stats["n_expand"], stats["n_greedy_steps"]))
    assert bench.algorithmic_bytes(stats, dim, M, k) == float(want)
    # the SURVEY anchor: ef=64 at 10k×128, M=16 → 1,650·512 + 66·128 ≈ 0.85 MB per query
    one = bench.algorithmic_bytes(stats[:1], dim, M, k)
    assert abs(one - (1650 * 512 + 66 * 128)) < 2048


def test_recall_at_k():
    ids = np.array([[1, 2, 3, 4], [5, 6, 7, -1]])
    truth = np.array([[4, 3, 9, 1], [5, 8, 6, 7]])
    assert bench.recall_at_k(ids, truth) == (3 / 4 + 3 / 4) / 2


def test_workloads_are_deterministic_and_shaped():
    for kind in ("latent16", "latent16n", "spectrum"):
        a = bench.make_vectors(500, 96, 7, kind=kind)
        b = bench.make_vectors(500, 96, 7, kind=kind)
        assert a.dtype == np.float32 and a.shape == (500, 96) and a.tobytes() == b.tobytes()
        assert bench.make_vectors(500, 96, 8, kind=kind).tobytes() != a.tobytes()
    n = np.linalg.norm(bench.make_vectors(300, 96, 1, kind="latent16n"), axis=1)
    assert np.allclose(n, 1.0, atol=1e-5)  # config 4 data is L2-normalised (inner product)


def test_oracle_counters_feed_the_formula(oracle):
    from sqlite_muninn_b200.builder import build_graph
    base, q = uniform(1, 800, 16), uniform(2, 20, 16)
    g = build_graph(base, 0, 8, 60)
    _, _, cnt, stats, _ = oracle.graph(g).search(q, 5, 40)
    assert (cnt == 5).all()
    b = bench.algorithmic_bytes(stats, 16, 8, 5)
    assert b > 20 * 40 * 16 * 4  # at least ef distances per query
