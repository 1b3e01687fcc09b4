"""The CPU oracle replayed against outputs of the unmodified reference (tests/golden/*.npz,
made by tests/golden/make_golden.py): vtab queries after close/reopen, and direct hnsw_search."""
import numpy as np
import pytest

from tests.conftest import assert_results_match
from tests.helpers import GOLDEN, load_golden


def test_fixtures_present():
    assert len(GOLDEN) >= 8


@pytest.mark.parametrize("name", GOLDEN)
def test_oracle_reproduces_reference_outputs(oracle, name):
    g, queries, k, cases = load_golden(name)
    og = oracle.graph(g)
    for c in cases:
        ids, dist, cnt, st, _ = og.search(queries, k, c["ef"])
        # the vtab returns (double)float (src/hnsw_vtab.c:673): exact after the cast
        assert np.array_equal(cnt, c["cnt"]), name
        assert np.array_equal(ids, c["ids"]), name
        assert np.array_equal(dist.astype(np.float64), c["dist"].astype(np.float64)), name
        if c["ndist"] is not None:
            assert np.array_equal(st["n_dist"], c["ndist"]), name


@pytest.mark.parametrize("name", GOLDEN)
def test_canonical_order_equals_reference_order(oracle, name):
    """The (distance, slot) list the CUDA kernel follows gives the reference's results on every
    fixture (ties only inside bit-equal distance groups)."""
    g, queries, k, cases = load_golden(name)
    og = oracle.graph(g)
    for c in cases:
        ids, dist, cnt, st, _ = og.search(queries, k, c["ef"])
        ids2, dist2, cnt2, st2, _ = og.search(queries, k, c["ef"], canonical=True)
        assert_results_match(ids, dist, cnt, ids2, dist2, cnt2, name)
        assert np.array_equal(st["n_dist"], st2["n_dist"]) and np.array_equal(st["n_expand"], st2["n_expand"])
        assert np.array_equal(st["term"] == 2, st2["term"] == 2)
