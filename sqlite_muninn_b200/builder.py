"""Host-side graph construction in the reference's format (csrc/hnsw_build.c).

`build_graph` follows /root/reference/src/hnsw_algo.c:520-666 (hnsw_insert). With
batch=0 the graph is identical to the reference's; with batch>0 groups of nodes are
searched in parallel against the graph as it stood before the group (benchmark-input
mode, see the header of csrc/hnsw_build.c).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .graph import FlatGraph

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_SO = os.path.join(_HERE, "csrc", "libmuninn_host.so")

_p = np.ctypeslib.ndpointer
_f32 = _p(np.float32, flags="C_CONTIGUOUS")
_i64 = _p(np.int64, flags="C_CONTIGUOUS")
_i32 = _p(np.int32, flags="C_CONTIGUOUS")

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
        _lib = L
    return _lib


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
        nn, nl, ne = C.c_int64(), C.c_int64(), C.c_int64()
        es, ml = C.c_int32(), C.c_int32()
        L.muninn_build_sizes(h, C.byref(nn), C.byref(nl), C.byref(ne), C.byref(es), C.byref(ml))
        levels = np.empty(n, np.int32)
        offs = np.empty(nl.value + 1, np.int64)
        tg = np.empty(ne.value, np.int32)
        L.muninn_build_export(h, levels, offs, tg, 1 if sort_lists else 0)
    finally:
        L.muninn_build_destroy(h)
    ids = np.arange(n, dtype=np.int64) if ids is None else np.ascontiguousarray(ids, np.int64)
    return FlatGraph(dim, metric, M, ef_construction, ids, levels, np.zeros(n, np.uint8), v, offs, tg, es.value,
                     ml.value)
