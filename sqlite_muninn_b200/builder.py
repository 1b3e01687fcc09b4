"""Host-side graph construction in the reference's format (csrc/hnsw_build.c).

`build_graph` follows /root/reference/src/hnsw_algo.c:520-666 (hnsw_insert). With
batch=0 the graph is identical to the reference's; with batch>0 groups of nodes are
searched in parallel against the graph as it stood before the group (benchmark-input
mode, see the header of csrc/hnsw_build.c).
"""
from __future__ import annotations

import ctypes as C
import os
import time

import numpy as np

from .graph import FlatGraph

_HERE = os.path.dirname(os.path.abspath(__file__))
# MUNINN_HOST_SO: alternative build of the same library (tools/sanitize_host.sh: ASan + UBSan)
HOST_SO = os.environ.get("MUNINN_HOST_SO") or os.path.join(_HERE, "csrc", "libmuninn_host.so")

_p = np.ctypeslib.ndpointer
_f32 = _p(np.float32, flags="C_CONTIGUOUS")
_i64 = _p(np.int64, flags="C_CONTIGUOUS")
_i32 = _p(np.int32, flags="C_CONTIGUOUS")

_UPLOAD_T = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                        C.c_int32, C.c_int32)
_PATCH_T = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                       C.c_void_p, C.c_void_p, C.c_int32, C.c_int32)
_SEARCH_T = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p)


class BuildDevice(C.Structure):
    """MuninnBuildDevice (include/muninn_build.h)."""
    _fields_ = [("ctx", C.c_void_p), ("upload", _UPLOAD_T), ("patch", _PATCH_T), ("search", _SEARCH_T)]


class BuildRunStats(C.Structure):
    """MuninnBuildRunStats (include/muninn_build.h)."""
    _fields_ = [("groups", C.c_int64), ("device_queries", C.c_int64), ("host_searches", C.c_int64),
                ("patches", C.c_int64), ("patched_records", C.c_int64), ("warm_s", C.c_double),
                ("search_s", C.c_double), ("link_s", C.c_double), ("patch_s", C.c_double)]


_lib = None


def host_lib():
    global _lib
    if _lib is None:
        if not os.path.exists(HOST_SO):
            raise RuntimeError(f"{HOST_SO} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(or `make -C sqlite_muninn_b200/csrc`)")
        L = C.CDLL(HOST_SO)
        L.muninn_build_create.restype = C.c_void_p
        L.muninn_build_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, _f32]
        L.muninn_build_seed.argtypes = [C.c_void_p, C.c_uint32]
        L.muninn_build_insert.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int64, C.c_int]
        L.muninn_build_sizes.argtypes = [C.c_void_p] + [C.POINTER(C.c_int64)] * 3 + [C.POINTER(C.c_int32)] * 2
        L.muninn_build_export.argtypes = [C.c_void_p, _i32, _i64, _i32, C.c_int]
        L.muninn_build_destroy.argtypes = [C.c_void_p]
        L.muninn_build_set_avx2.argtypes = [C.c_int]
        # group-at-a-time interface (the device-assisted loop below)
        L.muninn_build_log_touched.argtypes = [C.c_void_p, C.c_int]
        L.muninn_build_count.restype = C.c_int64
        L.muninn_build_count.argtypes = [C.c_void_p]
        L.muninn_build_group_begin.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.muninn_build_group_levels.argtypes = [C.c_void_p, _i32]
        L.muninn_build_group_link.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.muninn_build_touched_count.restype = C.c_int64
        L.muninn_build_touched_count.argtypes = [C.c_void_p]
        L.muninn_build_touched_take.restype = C.c_int64
        L.muninn_build_touched_take.argtypes = [C.c_void_p, _i32, C.c_int64]
        L.muninn_build_export_nodes.argtypes = [C.c_void_p, C.c_int64, _i32, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.muninn_build_entry.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.muninn_build_phase_seconds.argtypes = [C.c_void_p, C.POINTER(C.c_double * 3)]
        L.muninn_build_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int, C.POINTER(BuildDevice),
                                       C.POINTER(BuildRunStats)]
        _lib = L
    return _lib


def _export(L, h, n, dim, metric, M, efc, v, ids, sort_lists) -> FlatGraph:
    nn, nl, ne = C.c_int64(), C.c_int64(), C.c_int64()
    es, ml = C.c_int32(), C.c_int32()
    L.muninn_build_sizes(h, C.byref(nn), C.byref(nl), C.byref(ne), C.byref(es), C.byref(ml))
    n = nn.value
    levels = np.empty(n, np.int32)
    offs = np.empty(nl.value + 1, np.int64)
    tg = np.empty(ne.value, np.int32)
    L.muninn_build_export(h, levels, offs, tg, 1 if sort_lists else 0)
    return FlatGraph(dim, metric, M, efc, np.ascontiguousarray(ids[:n]), levels, np.zeros(n, np.uint8), v[:n], offs,
                     tg, es.value, ml.value)


def build_graph(vectors, metric: int = 0, M: int = 16, ef_construction: int = 200, *, ids=None, batch: int = 0,
                warm: int = 4096, threads: int | None = None, sort_lists: bool = True,
                seed: int | None = None) -> FlatGraph:
    """Insert `vectors` in order and return the graph.

    sort_lists=True returns neighbour lists in ascending-target order: the state
    `load_index_from_shadow` (/root/reference/src/hnsw_vtab.c:286-341) rebuilds, which is
    what search parity is pinned on (SURVEY.md finding 5).
    """
    L = host_lib()
    v = np.ascontiguousarray(vectors, np.float32)
    n, dim = v.shape
    threads = threads or (os.cpu_count() or 1)
    h = L.muninn_build_create(dim, metric, M, ef_construction, n, v)
    if not h:
        raise MemoryError("muninn_build_create failed")
    try:
        if seed is not None:
            L.muninn_build_seed(h, seed)
        if L.muninn_build_insert(h, n, batch, warm, threads if batch > 0 else 1) != 0:
            raise RuntimeError("muninn_build_insert failed")
        ids = np.arange(n, dtype=np.int64) if ids is None else np.ascontiguousarray(ids, np.int64)
        return _export(L, h, n, dim, metric, M, ef_construction, v, ids, sort_lists)
    finally:
        L.muninn_build_destroy(h)


def build_graph_device(vectors, metric: int = 0, M: int = 16, ef_construction: int = 200, *, ids=None,
                       batch: int = 1024, warm: int = 4096, threads: int | None = None, sort_lists: bool = True,
                       seed: int | None = None, device: int = 0, index=None, stats: dict | None = None) -> FlatGraph:
    """Batch-mode construction with the layer-0 candidate searches on the GPU.

    Same group-synchronous rule as `build_graph(batch=...)` (header of csrc/hnsw_build.c), so without bit-equal
    candidate distances the graph is identical to it. Per group of `batch` nodes:

      1. the host draws the levels (reference generator, src/hnsw_algo.c:19-30, :240-248);
      2. the level-0 nodes (a fraction 1 - 1/M of them) are searched as ONE `muninn_cuda_search_batch(k = min(2M,
         efc), ef = efc)` on the device copy of the graph as it stood before the group: for a node of level 0,
         hnsw_insert's search (src/hnsw_algo.c:553-579) IS hnsw_search with ef = ef_construction, query = its vector;
      3. the host links the group (upper-level nodes search on the host; reverse edges, MN-RU prune) and logs the
         nodes whose lists changed;
      4. those records go back to the device as one `muninn_cuda_index_patch`.

    `index` is the device index to build into: by default an `HnswIndexCuda` on `device` (no CPU path: without a GPU
    this raises); the tests pass a stand-in with the same three methods (upload / patch_records / search) to check
    the host logic. Neighbour lists on the device are in insertion order, the order the host builder itself searches.
    """
    L = host_lib()
    v = np.ascontiguousarray(vectors, np.float32)
    n, dim = v.shape
    threads = threads or (os.cpu_count() or 1)
    ids = np.arange(n, dtype=np.int64) if ids is None else np.ascontiguousarray(ids, np.int64)
    assert ids.shape == (n,) and (n < 2 or bool(np.all(np.diff(ids) > 0))), "ids must be strictly ascending"
    identity_ids = bool(np.array_equal(ids, np.arange(n, dtype=np.int64)))
    own_index = index is None
    if own_index:
        from .cuda import HnswIndexCuda
        index = HnswIndexCuda(dim, metric, M, device)
    M0 = 2 * M
    k = min(M0, ef_construction)
    st = {"groups": 0, "device_queries": 0, "host_searches": 0, "patched_records": 0, "patches": 0,
          "warm_s": 0.0, "search_s": 0.0, "link_s": 0.0, "patch_s": 0.0}
    clock = time.perf_counter
    h = L.muninn_build_create(dim, metric, M, ef_construction, n, v)
    if not h:
        raise MemoryError("muninn_build_create failed")
    try:
        if seed is not None:
            L.muninn_build_seed(h, seed)
        # the first `warm` nodes: exact sequential inserts on the host, then one full upload
        n_warm = min(n, max(int(warm), 1))
        t = clock()
        if L.muninn_build_insert(h, n_warm, 0, 0, 1) != 0:
            raise RuntimeError("muninn_build_insert failed")
        index.upload(_export(L, h, n_warm, dim, metric, M, ef_construction, v, ids, False))
        st["warm_s"] = clock() - t
        L.muninn_build_log_touched(h, 1)
        n_dev = n_warm  # rows the device copy holds

        def flush():
            """device copy ← every record the host changed since the last flush"""
            nonlocal n_dev
            cnt = L.muninn_build_touched_count(h)
            n_now = L.muninn_build_count(h)
            if cnt == 0 and n_now == n_dev:
                return
            t = clock()
            log = np.empty(max(cnt, 1), np.int32)
            cnt = L.muninn_build_touched_take(h, log, log.shape[0])
            old = log[:cnt]  # each slot once (the builder de-duplicates while logging)
            old = old[old < n_dev]
            slots = np.concatenate([np.arange(n_dev, n_now, dtype=np.int32), old.astype(np.int32)])  # new nodes first
            nl, ne = C.c_int64(), C.c_int64()
            L.muninn_build_export_nodes(h, len(slots), slots, None, None, None, C.byref(nl), C.byref(ne))
            lv = np.empty(len(slots), np.int32)
            offs = np.empty(nl.value + 1, np.int64)
            tg = np.empty(max(ne.value, 1), np.int32)
            L.muninn_build_export_nodes(h, len(slots), slots, lv.ctypes.data, offs.ctypes.data, tg.ctypes.data,
                                        C.byref(nl), C.byref(ne))
            es, ml = C.c_int32(), C.c_int32()
            L.muninn_build_entry(h, C.byref(es), C.byref(ml))
            index.patch_records(n_now, slots, ids[slots], lv, None, v[n_dev:].ctypes.data if n_now > n_dev else
                                v.ctypes.data, offs, tg[:ne.value], es.value, ml.value)
            st["patches"] += 1
            st["patched_records"] += len(slots)
            st["patch_s"] += clock() - t
            n_dev = n_now

        g0, gn = C.c_int64(), C.c_int64()
        while L.muninn_build_count(h) < n:
            if L.muninn_build_group_begin(h, batch, C.byref(g0), C.byref(gn)) != 0:
                raise RuntimeError("muninn_build_group_begin failed")
            flush()  # a level-raising first node was inserted exactly inside group_begin
            if gn.value == 0:
                continue
            a, z = g0.value, g0.value + gn.value
            glev = np.empty(gn.value, np.int32)
            L.muninn_build_group_levels(h, glev)
            on_dev = np.flatnonzero(glev == 0)
            cand = np.full((gn.value, k), -1, np.int32)
            cand_cnt = np.full(gn.value, -1, np.int32)
            t = clock()
            if on_dev.size:
                found, _, cnt = index.search(v[a:z][on_dev], k, ef_construction)
                found = found[:, :k]
                rows = found if identity_ids else np.searchsorted(ids[:a], found)
                cand[on_dev] = np.where(np.arange(k)[None, :] < cnt[:, None], rows, -1).astype(np.int32)
                cand_cnt[on_dev] = cnt
            st["search_s"] += clock() - t
            t = clock()
            rc = L.muninn_build_group_link(h, cand.ctypes.data, cand_cnt.ctypes.data, k, threads)
            st["link_s"] += clock() - t
            if rc != 0:
                raise RuntimeError(f"muninn_build_group_link failed ({rc}): the device returned a slot outside the "
                                   "graph as it stood before the group")
            st["groups"] += 1
            st["device_queries"] += int(on_dev.size)
            st["host_searches"] += int(gn.value - on_dev.size)
            flush()
        ph = (C.c_double * 3)()
        L.muninn_build_phase_seconds(h, C.byref(ph))
        st["link_host_search_s"], st["link_sort_s"], st["link_reverse_prune_s"] = ph[0], ph[1], ph[2]
        if stats is not None:
            stats.update({k_: round(v_, 3) if isinstance(v_, float) else v_ for k_, v_ in st.items()})
        return _export(L, h, n, dim, metric, M, ef_construction, v, ids, sort_lists)
    finally:
        L.muninn_build_destroy(h)
        if own_index:
            index.close()


def _arr(ptr, n, dtype):
    return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(np.ctypeslib.as_ctypes_type(dtype))), shape=(max(int(n), 0),)) \
        if n else np.empty(0, dtype)


def _callbacks(index, dim, metric, M, efc):
    """A MuninnBuildDevice whose three calls land in a Python object with upload / patch_records / search (the test
    stand-in). Exceptions cannot cross the C frame: they are kept and turned into a non-zero return code."""
    box = {"err": None}

    def guard(fn):
        def run(*a):
            try:
                fn(*a)
                return 0
            except BaseException as e:  # noqa: BLE001
                box["err"] = e
                return -100
        return run

    def upload(ctx, n, ids, levels, vectors, offs, tg, entry, max_level):
        lv = _arr(levels, n, np.int32).copy()
        nl = int(lv.astype(np.int64).sum()) + n
        o = _arr(offs, nl + 1, np.int64).copy()
        index.upload(FlatGraph(dim, metric, M, efc, _arr(ids, n, np.int64).copy(), lv, np.zeros(n, np.uint8),
                               _arr(vectors, n * dim, np.float32).reshape(n, dim).copy(), o,
                               _arr(tg, int(o[-1]), np.int32).copy(), entry, max_level))

    def patch(ctx, n_total, n_patch, slots, ids, levels, vectors, offs, tg, entry, max_level):
        lv = _arr(levels, n_patch, np.int32).copy()
        nl = int(lv.astype(np.int64).sum()) + n_patch
        o = _arr(offs, nl + 1, np.int64).copy()
        index.patch_records(n_total, _arr(slots, n_patch, np.int32).copy(), _arr(ids, n_patch, np.int64).copy(), lv,
                            None, vectors, o, _arr(tg, int(o[-1]), np.int32).copy(), entry, max_level)

    def search(ctx, q, nq, k, ef, out_ids, out_dist, out_cnt):
        ids, dist, cnt = index.search(_arr(q, nq * dim, np.float32).reshape(nq, dim).copy(), k, ef)
        _arr(out_ids, nq * k, np.int64)[:] = np.ascontiguousarray(ids[:, :k], np.int64).ravel()
        _arr(out_dist, nq * k, np.float32)[:] = np.ascontiguousarray(dist[:, :k], np.float32).ravel()
        _arr(out_cnt, nq, np.int32)[:] = cnt

    dev = BuildDevice(None, _UPLOAD_T(guard(upload)), _PATCH_T(guard(patch)), _SEARCH_T(guard(search)))
    return dev, box


def build_graph_device_c(vectors, metric: int = 0, M: int = 16, ef_construction: int = 200, *, ids=None,
                         batch: int = 1024, warm: int = 4096, threads: int | None = None, sort_lists: bool = True,
                         seed: int | None = None, device: int = 0, index=None, stats: dict | None = None) -> FlatGraph:
    """`build_graph_device` with the group loop in C (`muninn_build_run`, include/muninn_build.h): the form a C host
    links — libmuninn_host drives libmuninn_cuda through the `MuninnBuildDevice` that `muninn_cuda_build_device`
    fills in; nothing of the loop runs in Python. Same graph as `build_graph_device` and as the host batch mode.
    `index`: an `HnswIndexCuda` (default: a new one on `device`), or for the CPU tests any object with upload /
    patch_records / search, reached through callbacks."""
    L = host_lib()
    v = np.ascontiguousarray(vectors, np.float32)
    n, dim = v.shape
    threads = threads or (os.cpu_count() or 1)
    ids = np.arange(n, dtype=np.int64) if ids is None else np.ascontiguousarray(ids, np.int64)
    assert ids.shape == (n,) and (n < 2 or bool(np.all(np.diff(ids) > 0))), "ids must be strictly ascending"
    from .cuda import HnswIndexCuda, lib as cuda_lib
    own_index = index is None
    if own_index:
        index = HnswIndexCuda(dim, metric, M, device)
    box = {"err": None}
    if isinstance(index, HnswIndexCuda):
        dev = BuildDevice()
        if cuda_lib().muninn_cuda_build_device(index.h, C.byref(dev)) != 0:
            raise RuntimeError("muninn_cuda_build_device failed")
    else:
        dev, box = _callbacks(index, dim, metric, M, ef_construction)
    h = L.muninn_build_create(dim, metric, M, ef_construction, n, v)
    if not h:
        raise MemoryError("muninn_build_create failed")
    try:
        if seed is not None:
            L.muninn_build_seed(h, seed)
        st = BuildRunStats()
        rc = L.muninn_build_run(h, ids.ctypes.data, batch, warm, threads, C.byref(dev), C.byref(st))
        if box["err"] is not None:
            raise box["err"]
        if rc != 0:
            msg = index.L.muninn_cuda_last_error(index.h).decode() if isinstance(index, HnswIndexCuda) else ""
            raise RuntimeError(f"muninn_build_run failed ({rc}) {msg}")
        if isinstance(index, HnswIndexCuda):
            index.n = n
        if stats is not None:
            stats.update({f: (round(getattr(st, f), 3) if f.endswith("_s") else int(getattr(st, f)))
                          for f, _ in BuildRunStats._fields_})
        return _export(L, h, n, dim, metric, M, ef_construction, v, ids, sort_lists)
    finally:
        L.muninn_build_destroy(h)
        if own_index:
            index.close()

