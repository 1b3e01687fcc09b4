"""`-m gpu`: muninn_cuda_index_patch — after the unmodified reference's hnsw_insert / hnsw_delete changed the
host index, sending only the records those calls touched (the new node and the nodes it was linked to; the
deleted node and its former neighbours: /root/reference/src/hnsw_algo.c:520-666, :717-805) leaves the device
copy answering exactly as after a full upload of the new state — ids, distances and work counters."""
import numpy as np
import pytest

from tests.conftest import assert_results_match, uniform

pytestmark = pytest.mark.gpu


def lists_of(g, s):
    first = int(g.first_list()[s])
    return [tuple(g.ids[g.targets[g.list_offsets[first + l]:g.list_offsets[first + l + 1]]]) for l in range(int(g.levels[s]) + 1)]


def changed_slots(g_old, g_new):
    """slots of g_new whose record differs from g_old's (new nodes included): what the reference's call touched"""
    out = []
    for s in range(g_new.n):
        if s >= g_old.n or g_old.deleted[s] != g_new.deleted[s] or lists_of(g_old, s) != lists_of(g_new, s):
            out.append(s)
    return out


@pytest.mark.parametrize("metric,M,dim", [(0, 8, 24), (1, 16, 128), (2, 4, 10)])
def test_patch_equals_full_upload(reference, metric, M, dim):
    from sqlite_muninn_b200.cuda import HnswIndexCuda
    n0, steps, burst = 600, 8, 10
    base = uniform(100 + dim, n0 + burst * steps, dim)
    q = uniform(101 + dim, 200, dim)
    ref = reference.build(base[:n0], ids=np.arange(n0, dtype=np.int64) * 3 + 5, metric=metric, M=M, efc=60)
    g = ref.export()
    patched = HnswIndexCuda.from_graph(g)
    next_id, rng, total_patched = int(g.ids[-1]) + 1, np.random.default_rng(7), 0
    for step in range(steps):
        # a burst of INSERTs (ids above every stored id, as AUTOINCREMENT rowids are) and a few DELETEs
        lo = n0 + burst * step
        new_ids = next_id + 2 * np.arange(burst, dtype=np.int64)
        next_id = int(new_ids[-1]) + 7
        assert ref.insert(new_ids, base[lo:lo + burst]) == 0
        live = g.ids[g.deleted == 0]
        for rid in rng.choice(live, 3, replace=False):
            assert ref.delete(int(rid)) == 0
        g_new = ref.export()
        touched = changed_slots(g, g_new)
        assert len(touched) < g_new.n // 2  # a patch, not a re-upload
        total_patched += len(touched)
        patched.patch(g_new, touched)
        with HnswIndexCuda.from_graph(g_new) as fresh:
            for k, ef in ((10, 64), (5, 5), (20, 200)):
                a = patched.search(q, k, ef, want_stats=True)
                b = fresh.search(q, k, ef, want_stats=True)
                for x, y in zip(a[:3], b[:3]):
                    assert np.array_equal(x, y), (step, k, ef)
                for f in ("n_dist", "n_expand", "n_greedy_steps", "entry0"):
                    assert np.array_equal(a[3][f], b[3][f]), (step, f)
            # and the unmodified reference, searching its own mutated index, agrees
            r_ids, r_dist, r_cnt, _ = ref.search(q, 10, 64)
            p_ids, p_dist, p_cnt = patched.search(q, 10, 64)
            assert_results_match(p_ids, p_dist, p_cnt, r_ids, r_dist, r_cnt, f"step {step} vs reference")
            ea, eb = patched.exact_knn(q, 10), fresh.exact_knn(q, 10)
            assert all(np.array_equal(x, y) for x, y in zip(ea, eb))
        g = g_new
    patched.close()
    assert total_patched > 0


def test_patch_rejects_what_it_cannot_express(reference):
    from sqlite_muninn_b200.cuda import HnswIndexCuda, MuninnCudaError
    base = uniform(5, 120, 8)
    ref = reference.build(base[:100], ids=np.arange(100, dtype=np.int64) + 10, metric=0, M=4, efc=20)
    g = ref.export()
    with HnswIndexCuda.from_graph(g) as idx:
        assert ref.insert(np.array([3], np.int64), base[100:101]) == 0  # an id BELOW the stored ones: slots shift
        g2 = ref.export()
        with pytest.raises(MuninnCudaError):
            idx.patch(g2, [0])
        with pytest.raises(MuninnCudaError):
            idx.patch(g, [0, 0])  # duplicate slot


def test_rejected_patch_leaves_the_handle_untouched(reference):
    """Validation comes before any mutation: after a patch the library refuses (a list over capacity found in
    the LAST record), searches answer as before and a sound patch of the same
    burst still applies — the device copy then equals a fresh upload."""
    import copy
    from sqlite_muninn_b200.cuda import HnswIndexCuda, MuninnCudaError
    base, q = uniform(6, 330, 12), uniform(7, 64, 12)
    ref = reference.build(base[:300], ids=np.arange(300, dtype=np.int64) + 1, metric=0, M=4, efc=30)
    g = ref.export()
    with HnswIndexCuda.from_graph(g) as idx:
        before = idx.search(q, 10, 50, want_stats=True)
        assert ref.insert(np.arange(301, 331, dtype=np.int64), base[300:]) == 0
        assert ref.delete(17) == 0
        g2 = ref.export()
        touched = changed_slots(g, g2)
        bad = copy.deepcopy(g2)
        assert touched[-1] == bad.n - 1  # the newest node: its record is the last one the library looks at
        # stretch the very last list (the newest node's top level) past any capacity (2M = 8)
        bad.targets = np.concatenate([bad.targets, np.zeros(9, np.int32)])
        bad.list_offsets = bad.list_offsets.copy()
        bad.list_offsets[-1] += 9
        with pytest.raises(MuninnCudaError, match="device rows hold"):
            idx.patch_unchecked(bad, touched)
        after = idx.search(q, 10, 50, want_stats=True)
        for x, y in zip(before[:3], after[:3]):
            assert np.array_equal(x, y)
        assert np.array_equal(before[3]["n_dist"], after[3]["n_dist"])
        idx.patch(g2, touched)
        with HnswIndexCuda.from_graph(g2) as fresh:
            a, b = idx.search(q, 10, 50), fresh.search(q, 10, 50)
            assert all(np.array_equal(x, y) for x, y in zip(a, b))


def test_lists_longer_than_capacity_after_heavy_deletion(reference, oracle):
    """600x8, M=4, ~425 random deletes through the unmodified hnsw_delete leave live nodes with more than M
    upper-layer (or 2M layer-0) neighbours. Upload must take the graph as the reference holds it (row width = longest
    list), a patch that would outgrow the device rows must be refused cleanly (the glue then re-uploads), and search
    must equal the oracle on that graph — ids, distances, counters."""
    from sqlite_muninn_b200.cuda import HnswIndexCuda, MuninnCudaError
    n, dim, M = 600, 8, 4
    base, q = uniform(51, n, dim), uniform(52, 100, dim)
    ref = reference.build(base, ids=np.arange(n, dtype=np.int64) + 1, metric=0, M=M, efc=40)
    g0 = ref.export()
    dev = HnswIndexCuda.from_graph(g0)
    g_dev = g0  # the state the device copy holds
    over, refused = False, 0
    for step, i in enumerate(np.random.default_rng(3).permutation(n)[:450]):
        assert ref.delete(int(i) + 1) == 0
        if step % 50 != 0:
            continue
        g = ref.export()
        first = g.first_list()
        lens0 = np.diff(g.list_offsets)[first[:-1]]
        lens_all = np.diff(g.list_offsets)
        upper = np.ones(len(lens_all), bool)
        upper[first[:-1]] = False
        over = over or bool(lens0.max() > 2 * M or (upper.any() and lens_all[upper].max() > M))
        try:
            dev.patch(g, changed_slots(g_dev, g))  # everything the deletes since the last sync touched
        except MuninnCudaError as e:  # outgrew the device rows: a clean refusal, then the fallback the glue takes
            assert "upload again" in str(e)
            refused += 1
            dev.upload(g)
        g_dev = g
        with HnswIndexCuda.from_graph(g) as fresh:
            for k, ef in ((10, 40), (5, 5)):
                a, b = dev.search(q, k, ef, want_stats=True), fresh.search(q, k, ef, want_stats=True)
                o = oracle.graph(g).search(q, k, ef, canonical=True)
                for x, y, z in zip(a[:3], b[:3], o[:3]):
                    assert np.array_equal(x, y) and np.array_equal(x, z), (step, k, ef)
                assert np.array_equal(a[3]["n_dist"], o[3]["n_dist"])
            r_ids, r_dist, r_cnt, _ = ref.search(q, 10, 40)
            a = dev.search(q, 10, 40)
            assert_results_match(a[0], a[1], a[2], r_ids, r_dist, r_cnt, f"vs reference after {step + 1} deletes")
    assert over, "the reference never went past its nominal list capacity: the test exercised nothing new"
    dev.close()
