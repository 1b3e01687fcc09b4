"""ctypes bindings for the parity checker: oracle/liboracle.so (our CPU restatement) and,
when it was built, oracle/_ref/libmuninn_ref.so (the unmodified reference + ref_harness.c).

TEST INFRASTRUCTURE. Nothing under sqlite_muninn_b200/ imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.environ.get("MUNINN_ORACLE_SO") or os.path.join(HERE, "liboracle.so")  # sanitizer builds
REF_SO = os.path.join(HERE, "_ref", "libmuninn_ref.so")
REF_EXT_SO = os.path.join(HERE, "_ref", "muninn_ref_ext.so")

_p = np.ctypeslib.ndpointer
_f32 = _p(np.float32, flags="C_CONTIGUOUS")
_i64 = _p(np.int64, flags="C_CONTIGUOUS")
_i32 = _p(np.int32, flags="C_CONTIGUOUS")
_u8 = _p(np.uint8, flags="C_CONTIGUOUS")


def build(force: bool = False) -> None:
    """Compile the checker (liboracle.so always; _ref when /root/reference is present)."""
    if force or not os.path.exists(ORACLE_SO) or os.path.exists("/root/reference/src/hnsw_algo.c"):
        subprocess.run(["make", "-s", "-C", HERE], check=True)


class OrcStats(C.Structure):
    _fields_ = [("n_dist", C.c_int64), ("n_dist_upper", C.c_int64), ("n_expand", C.c_int32),
                ("n_greedy_steps", C.c_int32), ("term", C.c_int32), ("entry0", C.c_int32)]


STATS_DTYPE = np.dtype([("n_dist", np.int64), ("n_dist_upper", np.int64), ("n_expand", np.int32),
                        ("n_greedy_steps", np.int32), ("term", np.int32), ("entry0", np.int32)])


class Oracle:
    """Our CPU restatement (oracle/hnsw_oracle.c)."""

    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build()
        L = C.CDLL(ORACLE_SO)
        L.orc_distance.restype = C.c_float
        L.orc_distance.argtypes = [C.c_int, _f32, _f32, C.c_int]
        L.orc_parse_metric.argtypes = [C.c_char_p]
        L.orc_graph_create.restype = C.c_void_p
        L.orc_graph_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int64, _i64, _i32, _u8, _f32, _i64, _i32,
                                       C.c_int32, C.c_int32]
        L.orc_graph_destroy.argtypes = [C.c_void_p]
        L.orc_search.restype = C.c_double
        L.orc_search.argtypes = [C.c_void_p, _f32, C.c_int, C.c_int, C.c_int, C.c_int, _i64, _f32, _i32, C.c_void_p]
        L.orc_search_mode.restype = C.c_double
        L.orc_search_mode.argtypes = [C.c_void_p, _f32, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _i64, _f32, _i32,
                                      C.c_void_p]
        L.orc_exact_knn.restype = C.c_double
        L.orc_exact_knn.argtypes = [_f32, C.c_void_p, C.c_int64, C.c_int, C.c_int, _f32, C.c_int, C.c_int, C.c_int,
                                    _i64, _f32, _i32]
        L.orc_heap_sequence.argtypes = [_f32, _i32, C.c_int, _i32, _f32]
        self.L = L

    def distance(self, metric: int, a, b) -> np.float32:
        a = np.ascontiguousarray(a, np.float32)
        b = np.ascontiguousarray(b, np.float32)
        return np.float32(self.L.orc_distance(metric, a, b, a.shape[0]))

    def parse_metric(self, name: str) -> int:
        return self.L.orc_parse_metric(name.encode())

    def graph(self, g) -> "OracleGraph":
        return OracleGraph(self, g)

    def exact_knn(self, base, queries, k: int, metric: int, deleted=None, threads: int = 1):
        base = np.ascontiguousarray(base, np.float32)
        queries = np.ascontiguousarray(queries, np.float32)
        nq = queries.shape[0]
        idx = np.empty((nq, k), np.int64)
        dist = np.empty((nq, k), np.float32)
        cnt = np.empty(nq, np.int32)
        dp = None if deleted is None else np.ascontiguousarray(deleted, np.uint8).ctypes.data_as(C.c_void_p)
        secs = self.L.orc_exact_knn(base, dp, base.shape[0], base.shape[1], metric, queries, nq, k, threads, idx, dist,
                                    cnt)
        return idx, dist, cnt, secs

    def heap_sequence(self, keys, ops):
        keys = np.ascontiguousarray(keys, np.float32)
        ops = np.ascontiguousarray(ops, np.int32)
        n_out = int(ops.sum())
        out_s = np.empty(n_out, np.int32)
        out_k = np.empty(n_out, np.float32)
        self.L.orc_heap_sequence(keys, ops, len(ops), out_s, out_k)
        return out_s, out_k


class OracleGraph:
    def __init__(self, orc: Oracle, g):
        self.orc, self.g = orc, g
        g.validate()
        self.h = orc.L.orc_graph_create(g.dim, g.metric, g.M, g.n, g.ids, g.levels, g.deleted, g.vectors,
                                        g.list_offsets, g.targets, g.entry_slot, g.max_level)

    def search(self, queries, k: int, ef: int, threads: int = 1, want_stats: bool = True, canonical: bool = False):
        """canonical=False: the reference's two-heap loop (pinned against the real reference);
        canonical=True: the (distance, slot)-ordered list the CUDA kernel is specified to follow."""
        q = np.ascontiguousarray(queries, np.float32)
        nq = q.shape[0]
        ids = np.empty((nq, k), np.int64)
        dist = np.empty((nq, k), np.float32)
        cnt = np.empty(nq, np.int32)
        stats = np.zeros(nq, STATS_DTYPE) if want_stats else None
        sp = stats.ctypes.data_as(C.c_void_p) if want_stats else None
        secs = self.orc.L.orc_search_mode(self.h, q, nq, k, ef, threads, 1 if canonical else 0, ids, dist, cnt, sp)
        return ids, dist, cnt, stats, secs

    def close(self):
        if self.h:
            self.orc.L.orc_graph_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()


def have_reference() -> bool:
    return os.path.exists(REF_SO)


class Reference:
    """The unmodified reference hot path (oracle/_ref/libmuninn_ref.so)."""

    def __init__(self):
        if not have_reference():
            raise RuntimeError("oracle/_ref/libmuninn_ref.so not built (needs /root/reference)")
        L = C.CDLL(REF_SO)
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [C.c_int] * 4
        L.ref_destroy.argtypes = [C.c_void_p]
        L.ref_insert_many.argtypes = [C.c_void_p, C.c_int64, _i64, _f32]
        L.ref_delete.argtypes = [C.c_void_p, C.c_int64]
        L.ref_sort_lists.argtypes = [C.c_void_p]
        L.ref_export_sizes.argtypes = [C.c_void_p] + [C.POINTER(C.c_int64)] * 3
        L.ref_export.argtypes = [C.c_void_p, _i64, _i32, _u8, _f32, _i64, _i32, C.POINTER(C.c_int32),
                                 C.POINTER(C.c_int32)]
        L.ref_import.restype = C.c_void_p
        L.ref_import.argtypes = [C.c_int] * 4 + [C.c_int64, _i64, _i32, _u8, _f32, _i64, _i32, C.c_int32, C.c_int32]
        L.ref_search.argtypes = [C.c_void_p, _f32, C.c_int, C.c_int, C.c_int, _i64, _f32, _i32, C.c_void_p]
        L.ref_search_mt.restype = C.c_double
        L.ref_search_mt.argtypes = [C.c_void_p, _f32, C.c_int, C.c_int, C.c_int, C.c_int, _i64, _f32, _i32]
        L.ref_distance.restype = C.c_float
        L.ref_distance.argtypes = [C.c_int, _f32, _f32, C.c_int]
        L.ref_parse_metric.argtypes = [C.c_char_p]
        self.L = L

    def distance(self, metric, a, b):
        a = np.ascontiguousarray(a, np.float32)
        b = np.ascontiguousarray(b, np.float32)
        return np.float32(self.L.ref_distance(metric, a, b, a.shape[0]))

    def parse_metric(self, name: str) -> int:
        return self.L.ref_parse_metric(name.encode())

    def build(self, vectors, ids=None, metric=0, M=16, efc=200) -> "RefIndex":
        v = np.ascontiguousarray(vectors, np.float32)
        ids = np.arange(v.shape[0], dtype=np.int64) if ids is None else np.ascontiguousarray(ids, np.int64)
        h = self.L.ref_create(v.shape[1], metric, M, efc)
        if self.L.ref_insert_many(h, v.shape[0], ids, v) != 0:
            raise RuntimeError("reference hnsw_insert failed")
        return RefIndex(self, h, v.shape[1], metric, M, efc)

    def load(self, g) -> "RefIndex":
        g.validate()
        h = self.L.ref_import(g.dim, g.metric, g.M, g.ef_construction, g.n, g.ids, g.levels, g.deleted, g.vectors,
                              g.list_offsets, g.targets, g.entry_slot, g.max_level)
        return RefIndex(self, h, g.dim, g.metric, g.M, g.ef_construction)


class RefIndex:
    def __init__(self, ref, h, dim, metric, M, efc):
        self.ref, self.h, self.dim, self.metric, self.M, self.efc = ref, h, dim, metric, M, efc

    def delete(self, rowid: int) -> int:
        return self.ref.L.ref_delete(self.h, rowid)

    def insert(self, ids, vectors) -> int:
        """hnsw_insert for each (id, vector), in order."""
        v = np.ascontiguousarray(np.atleast_2d(vectors), np.float32)
        ids = np.ascontiguousarray(ids, np.int64)
        return self.ref.L.ref_insert_many(self.h, v.shape[0], ids, v)

    def sort_lists(self):
        """Bring neighbour lists into shadow-table (reloaded) order."""
        self.ref.L.ref_sort_lists(self.h)

    def export(self):
        from sqlite_muninn_b200.graph import FlatGraph
        nn, nl, ne = C.c_int64(), C.c_int64(), C.c_int64()
        self.ref.L.ref_export_sizes(self.h, C.byref(nn), C.byref(nl), C.byref(ne))
        n = nn.value
        ids = np.empty(n, np.int64)
        levels = np.empty(n, np.int32)
        deleted = np.empty(n, np.uint8)
        vecs = np.empty((n, self.dim), np.float32)
        offs = np.empty(nl.value + 1, np.int64)
        tg = np.empty(ne.value, np.int32)
        es, ml = C.c_int32(), C.c_int32()
        self.ref.L.ref_export(self.h, ids, levels, deleted, vecs, offs, tg, C.byref(es), C.byref(ml))
        return FlatGraph(self.dim, self.metric, self.M, self.efc, ids, levels, deleted, vecs, offs, tg, es.value,
                         ml.value)

    def search(self, queries, k, ef, want_ndist=True):
        q = np.ascontiguousarray(queries, np.float32)
        nq = q.shape[0]
        ids = np.empty((nq, k), np.int64)
        dist = np.empty((nq, k), np.float32)
        cnt = np.empty(nq, np.int32)
        nd = np.zeros(nq, np.int64) if want_ndist else None
        self.ref.L.ref_search(self.h, q, nq, k, ef, ids, dist, cnt,
                              nd.ctypes.data_as(C.c_void_p) if want_ndist else None)
        return ids, dist, cnt, nd

    def search_mt(self, queries, k, ef, threads):
        q = np.ascontiguousarray(queries, np.float32)
        nq = q.shape[0]
        ids = np.empty((nq, k), np.int64)
        dist = np.empty((nq, k), np.float32)
        cnt = np.empty(nq, np.int32)
        secs = self.ref.L.ref_search_mt(self.h, q, nq, k, ef, threads, ids, dist, cnt)
        return ids, dist, cnt, secs

    def close(self):
        if self.h:
            self.ref.L.ref_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()
