"""Device-assisted graph construction: host logic on CPU, the real loop on the GPU.

`build_graph_device` hands the layer-0 candidate searches of each insert group to the device index
(`muninn_cuda_search_batch(k = min(2M, efc), ef = efc)`) and writes the changed records back with
`muninn_cuda_index_patch`. For a level-0 node that search IS the one hnsw_insert runs
(/root/reference/src/hnsw_algo.c:553-579), so the graph must equal the host builder's batch mode bit for bit.

CPU tests put a stand-in in the device index's place that keeps the uploaded/patched records and answers searches
with the CPU oracle: they check the loop, the touched-node log and the patch records. The product path (searches on
the GPU, patches through the C ABI) is tests/test_gpu_zz_device_build.py.
"""
import ctypes as C

import numpy as np
import pytest

from oracle.bindings import Oracle
from sqlite_muninn_b200.builder import build_graph, build_graph_device, build_graph_device_c
from sqlite_muninn_b200.graph import FlatGraph


class OracleBackedIndex:
    """Same three methods as HnswIndexCuda (upload / patch_records / search); state = the records it was sent."""

    def __init__(self, dim, metric, M):
        self.dim, self.metric, self.M = dim, metric, M
        self.orc = Oracle()
        self.patch_sizes = []

    def upload(self, g: FlatGraph):
        g.validate()
        self.ids = g.ids.copy()
        self.levels = g.levels.copy()
        self.vectors = g.vectors.copy()
        first = g.first_list()
        self.lists = [[g.targets[g.list_offsets[first[s] + l]:g.list_offsets[first[s] + l + 1]].copy()
                       for l in range(g.levels[s] + 1)] for s in range(g.n)]
        self.entry, self.max_level = g.entry_slot, g.max_level

    def patch_records(self, n_total, slots, ids, levels, deleted, vectors_ptr, list_offsets, targets, entry_slot,
                      max_level):
        n_old = len(self.ids)
        n_new = n_total - n_old
        assert deleted is None and len(set(slots.tolist())) == len(slots)
        # the library's contract: every new node is in the patch; only their vectors are read, record-indexed
        assert sorted(s for s in slots.tolist() if s >= n_old) == list(range(n_old, n_total))
        self.ids = np.concatenate([self.ids, np.zeros(n_new, np.int64)])
        self.levels = np.concatenate([self.levels, np.zeros(n_new, np.int32)])
        self.vectors = np.concatenate([self.vectors, np.zeros((n_new, self.dim), np.float32)])
        self.lists += [None] * n_new
        li = 0
        for i, s in enumerate(slots.tolist()):
            if s >= n_old:
                row = (C.c_float * self.dim).from_address(vectors_ptr + i * self.dim * 4)
                self.vectors[s] = np.frombuffer(row, np.float32)
                self.ids[s], self.levels[s] = ids[i], levels[i]
            else:
                assert levels[i] == self.levels[s] and ids[i] == self.ids[s]
            self.lists[s] = [np.array(targets[list_offsets[li + l]:list_offsets[li + l + 1]], np.int32)
                             for l in range(levels[i] + 1)]
            li += levels[i] + 1
        assert li + 1 == len(list_offsets)
        self.entry, self.max_level = entry_slot, max_level
        self.patch_sizes.append(len(slots))

    def graph(self) -> FlatGraph:
        flat = [l for node in self.lists for l in node]
        offs = np.zeros(len(flat) + 1, np.int64)
        np.cumsum([len(l) for l in flat], out=offs[1:])
        tg = np.ascontiguousarray(np.concatenate(flat) if flat else np.empty(0), np.int32)
        n = len(self.ids)
        return FlatGraph(self.dim, self.metric, self.M, 0, self.ids, self.levels, np.zeros(n, np.uint8),
                         np.ascontiguousarray(self.vectors), offs, tg, self.entry, self.max_level)

    def search(self, queries, k, ef):
        og = self.orc.graph(self.graph())
        ids, dist, cnt, _, _ = og.search(queries, k, ef, want_stats=False, canonical=True)
        og.close()
        return ids, dist, cnt


CASES = [  # n, dim, metric, M, efc, batch, warm
    (1600, 24, 0, 8, 40, 256, 300),
    (1200, 16, 1, 4, 30, 100, 0),      # warm = 0: the first node is the exact insert that creates the entry point
    (1500, 32, 2, 16, 24, 400, 500),   # efc < 2M: k = efc
]


@pytest.mark.parametrize("n,dim,metric,M,efc,batch,warm", CASES)
def test_group_loop_equals_host_batch_mode(n, dim, metric, M, efc, batch, warm):
    rng = np.random.default_rng(n)
    x = rng.standard_normal((n, dim)).astype(np.float32)
    want = build_graph(x, metric, M, efc, batch=batch, warm=warm, sort_lists=False, threads=4)
    idx = OracleBackedIndex(dim, metric, M)
    st = {}
    got = build_graph_device(x, metric, M, efc, batch=batch, warm=warm, sort_lists=False, threads=4, index=idx,
                             stats=st)
    assert got.same_structure(want)
    # the device copy ends up holding the same graph, lists in the builder's own order
    assert idx.graph().same_structure(want)
    # the device answered for the level-0 nodes of every group, the host for the rest
    grouped = st["device_queries"] + st["host_searches"]  # the rest: level-raising nodes, inserted exactly
    assert n - max(warm, 1) - 2 * st["groups"] - 8 <= grouped <= n - max(warm, 1)
    assert st["device_queries"] > 0.5 * (n - warm)
    # patches carry the touched nodes only
    assert max(idx.patch_sizes) <= batch * (2 * M + 2)


def test_rowids_need_not_be_slots():
    rng = np.random.default_rng(5)
    x = rng.standard_normal((900, 16)).astype(np.float32)
    ids = np.cumsum(rng.integers(1, 9, 900)).astype(np.int64) + 100
    want = build_graph(x, 0, 8, 32, ids=ids, batch=128, warm=200, threads=2)
    got = build_graph_device(x, 0, 8, 32, ids=ids, batch=128, warm=200, threads=2, index=OracleBackedIndex(16, 0, 8))
    assert got.same_structure(want)


def test_link_rejects_slots_from_outside_the_graph():
    from sqlite_muninn_b200.builder import host_lib
    L = host_lib()
    x = np.random.default_rng(1).standard_normal((300, 8)).astype(np.float32)
    h = L.muninn_build_create(8, 0, 4, 20, 300, x)
    try:
        assert L.muninn_build_insert(h, 100, 0, 0, 1) == 0
        g0, gn = C.c_int64(), C.c_int64()
        assert L.muninn_build_group_link(h, None, None, 0, 1) == -1          # no open group
        assert L.muninn_build_group_begin(h, 50, C.byref(g0), C.byref(gn)) == 0 and gn.value > 0
        assert L.muninn_build_group_begin(h, 50, C.byref(g0), C.byref(gn)) == -1  # already open
        assert L.muninn_build_insert(h, 10, 0, 0, 1) == -1                   # not while a group is open
        cand = np.full((gn.value, 8), 5, np.int32)
        cand[0, 0] = g0.value                                                 # a node of the group itself
        cnt = np.full(gn.value, 8, np.int32)
        assert L.muninn_build_group_link(h, cand.ctypes.data, cnt.ctypes.data, 8, 1) == -2
        cand[0, 0] = 7
        assert L.muninn_build_group_link(h, cand.ctypes.data, cnt.ctypes.data, 8, 1) == 0
        assert L.muninn_build_count(h) >= 100 + gn.value
    finally:
        L.muninn_build_destroy(h)


def check_graph_invariants(g: FlatGraph):
    """What every graph of the reference's construction rule satisfies (src/hnsw_algo.c:582-646)."""
    g.validate()
    first = g.first_list()
    assert g.levels[g.entry_slot] == g.max_level == g.levels.max()
    for s in range(g.n):
        for l in range(g.levels[s] + 1):
            lst = g.targets[g.list_offsets[first[s] + l]:g.list_offsets[first[s] + l + 1]]
            assert len(lst) <= (2 * g.M if l == 0 else g.M)
            assert len(set(lst.tolist())) == len(lst) and s not in lst and lst.min(initial=0) >= 0
            assert np.all(g.levels[lst] >= l)


def recall_at_10(g: FlatGraph, queries, ef=64):
    orc = Oracle()
    og = orc.graph(g)
    ids, _, _, _, _ = og.search(queries, 10, ef, want_stats=False, canonical=True)
    og.close()
    truth, _, _, _ = orc.exact_knn(g.vectors, queries, 10, g.metric)
    return np.mean([len(set(a) & set(b)) / 10 for a, b in zip(ids.tolist(), truth.tolist())])


def test_tie_heavy_data_gives_a_sound_graph():
    """Small-integer coordinates: bit-equal distances everywhere. The device's (distance, slot) order and the host's
    heap order may then pick different neighbours (both are graphs of the reference's rule): check the invariants
    and that search quality matches the host-built graph's."""
    rng = np.random.default_rng(9)
    x = rng.integers(0, 4, (1500, 12)).astype(np.float32)
    q = rng.integers(0, 4, (100, 12)).astype(np.float32)
    host = build_graph(x, 0, 8, 40, batch=128, warm=200, threads=2)
    dev = build_graph_device(x, 0, 8, 40, batch=128, warm=200, threads=2, index=OracleBackedIndex(12, 0, 8))
    check_graph_invariants(dev)
    check_graph_invariants(host)
    assert np.array_equal(dev.levels, host.levels) and dev.entry_slot == host.entry_slot
    assert abs(recall_at_10(dev, q) - recall_at_10(host, q)) < 0.1


def test_tie_free_quality_equals_host():
    rng = np.random.default_rng(10)
    x = rng.standard_normal((1500, 12)).astype(np.float32)
    q = rng.standard_normal((100, 12)).astype(np.float32)
    dev = build_graph_device(x, 0, 8, 40, batch=128, warm=200, threads=2, index=OracleBackedIndex(12, 0, 8))
    check_graph_invariants(dev)
    assert recall_at_10(dev, q) >= 0.9



def test_no_device_no_build():
    """Without an index argument the searches go to the GPU or nowhere: there is no host fallback behind this call."""
    from sqlite_muninn_b200 import cuda
    if cuda.device_count() > 0:
        pytest.skip("a CUDA device is present")
    x = np.random.default_rng(2).standard_normal((200, 8)).astype(np.float32)
    with pytest.raises(cuda.MuninnCudaError) as e:
        build_graph_device(x, 0, 4, 20, batch=32, warm=50)
    assert e.value.code == -2  # MUNINN_CUDA_E_NODEVICE


@pytest.mark.parametrize("n,dim,metric,M,efc,batch,warm", CASES)
def test_c_group_loop_equals_host_batch_mode(n, dim, metric, M, efc, batch, warm):
    """muninn_build_run (the loop in C, include/muninn_build.h) driving the stand-in through callbacks."""
    x = np.random.default_rng(n).standard_normal((n, dim)).astype(np.float32)
    want = build_graph(x, metric, M, efc, batch=batch, warm=warm, sort_lists=False, threads=4)
    idx = OracleBackedIndex(dim, metric, M)
    st = {}
    got = build_graph_device_c(x, metric, M, efc, batch=batch, warm=warm, sort_lists=False, threads=4, index=idx, stats=st)
    assert got.same_structure(want)
    assert idx.graph().same_structure(want)
    st_py = {}
    build_graph_device(x, metric, M, efc, batch=batch, warm=warm, sort_lists=False, threads=4,
                       index=OracleBackedIndex(dim, metric, M), stats=st_py)
    for key in ("groups", "device_queries", "host_searches", "patches", "patched_records"):
        assert st[key] == st_py[key], key


def test_c_group_loop_with_rowids_and_device_failure():
    rng = np.random.default_rng(6)
    x = rng.standard_normal((800, 16)).astype(np.float32)
    ids = np.cumsum(rng.integers(1, 9, 800)).astype(np.int64) + 7
    want = build_graph(x, 0, 8, 32, ids=ids, batch=128, warm=200, threads=2)
    got = build_graph_device_c(x, 0, 8, 32, ids=ids, batch=128, warm=200, threads=2, index=OracleBackedIndex(16, 0, 8))
    assert got.same_structure(want)

    class Failing(OracleBackedIndex):
        def search(self, queries, k, ef):
            raise RuntimeError("device lost")

    with pytest.raises(RuntimeError, match="device lost"):  # no host fallback: the build ends
        build_graph_device_c(x, 0, 8, 32, batch=128, warm=200, threads=2, index=Failing(16, 0, 8))

