"""Shared fixtures. `-m "not gpu"`: oracle vs reference/golden, host logic, ABI surface.
`-m gpu`: parity of the CUDA path (through the C ABI) against the oracle."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Everything native is built once per session (no-op when the .so files are current)."""
    import __graft_entry__ as ge
    ge.build()


@pytest.fixture(scope="session")
def oracle():
    from oracle.bindings import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    from oracle.bindings import Reference, have_reference
    if not have_reference():
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    return Reference()


def uniform(seed: int, n: int, dim: int) -> np.ndarray:
    return np.random.default_rng(seed).uniform(-1.0, 1.0, (n, dim)).astype(np.float32)


def lcg_vectors(n: int, dim: int) -> np.ndarray:
    """The generator of the reference's recall test (test/test_hnsw_algo.c:104-112)."""
    out = np.empty((n, dim), np.float32)
    seed = 12345
    for i in range(n):
        for d in range(dim):
            seed = (seed * 1103515245 + 12345) & 0xFFFFFFFF
            out[i, d] = np.float32(np.float32(seed % 10000) / np.float32(10000.0))
    return out


def assert_results_match(ids_a, dist_a, cnt_a, ids_b, dist_b, cnt_b, what=""):
    """ID-exact up to permutations inside groups of bit-equal distances (heap order among
    equal keys is unspecified in the reference, test/test_priority_queue.c:85-99)."""
    assert np.array_equal(cnt_a, cnt_b), f"{what}: result counts differ"
    assert np.array_equal(dist_a, dist_b), f"{what}: distances differ"
    if np.array_equal(ids_a, ids_b):
        return
    bad = np.nonzero((ids_a != ids_b).any(axis=1))[0]
    for q in bad:
        c = int(cnt_a[q])
        d = dist_a[q, :c]
        for v in np.unique(d):
            m = d == v
            assert sorted(ids_a[q, :c][m]) == sorted(ids_b[q, :c][m]), f"{what}: query {q} differs beyond a tie group"
