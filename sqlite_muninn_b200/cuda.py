"""ctypes binding of libmuninn_cuda (include/muninn_cuda.h) and the host-side index object.

`HnswIndexCuda` mirrors the reference's in-memory index API for the search path
(/root/reference/src/hnsw_algo.h:56-73): create(dim, metric, M) → upload(graph) →
search(queries, k, ef) with the reference's argument meaning and error behaviour. It is a
thin veneer: every call goes through the C ABI; there is no Python or CPU compute path, and
a missing library or device raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .graph import FlatGraph

_HERE = os.path.dirname(os.path.abspath(__file__))
# MUNINN_CUDA_SO: alternative build of the same library (kernel A/B experiments)
CUDA_SO = os.environ.get("MUNINN_CUDA_SO") or os.path.join(_HERE, "csrc", "libmuninn_cuda.so")

STATS_DTYPE = np.dtype([("n_dist", np.int64), ("n_dist_upper", np.int64), ("n_expand", np.int32),
                        ("n_greedy_steps", np.int32), ("term", np.int32), ("entry0", np.int32)])

# every symbol include/muninn_cuda.h declares (tests check the library exports all of them)
ABI_SYMBOLS = (
    "muninn_cuda_device_count", "muninn_cuda_version", "muninn_cuda_strerror", "muninn_cuda_parse_metric",
    "muninn_cuda_index_create", "muninn_cuda_index_destroy", "muninn_cuda_last_error", "muninn_cuda_index_upload",
    "muninn_cuda_search_batch", "muninn_cuda_search_batch_device", "muninn_cuda_search_batch_device_static",
    "muninn_cuda_static_scratch_limits", "muninn_cuda_set_stream",
    "muninn_cuda_last_kernel_ms", "muninn_cuda_launch_count", "muninn_cuda_set_tuning", "muninn_cuda_exact_knn",
    "muninn_cuda_exact_knn_device", "muninn_cuda_topk_merge_device", "muninn_cuda_distance_batch",
    "muninn_cuda_set_exact_mode", "muninn_cuda_exact_stats", "muninn_cuda_exact_filter_probe",
    "muninn_cuda_gather_ceiling", "muninn_cuda_index_patch", "muninn_cuda_build_device",
    "muninn_cuda_last_team_width", "muninn_cuda_last_kernel_gen", "muninn_cuda_set_option", "muninn_cuda_exact_knn_sharded_device",
    "muninn_cuda_comm_unique_id", "muninn_cuda_comm_create", "muninn_cuda_comm_destroy",
    "muninn_cuda_topk_pack_keys_device", "muninn_cuda_topk_merge_keys_device", "muninn_cuda_index_upload_rows",
)

EXACT_CUDA_CORE, EXACT_TC_3XTF32, EXACT_TC_1XTF32, EXACT_TC_AUTO = 0, 1, 2, 3


class ExactStats(C.Structure):
    """MuninnExactStats (include/muninn_cuda.h)."""
    _fields_ = [("mode", C.c_int32), ("chunks", C.c_int32), ("candidates", C.c_int64), ("overflow_queries", C.c_int32),
                ("filter_ms", C.c_float), ("c_eps", C.c_float)]

_p = np.ctypeslib.ndpointer
_f32 = _p(np.float32, flags="C_CONTIGUOUS")
_i64 = _p(np.int64, flags="C_CONTIGUOUS")
_i32 = _p(np.int32, flags="C_CONTIGUOUS")
_vp = C.c_void_p

_lib = None


class MuninnCudaError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libmuninn_cuda error {code}: {msg}")
        self.code = code


def lib():
    """Load libmuninn_cuda.so; fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(CUDA_SO):
            raise RuntimeError(f"{CUDA_SO} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(or `make -C sqlite_muninn_b200/csrc`). There is no CPU fallback.")
        L = C.CDLL(CUDA_SO)
        L.muninn_cuda_version.restype = C.c_char_p
        L.muninn_cuda_strerror.restype = C.c_char_p
        L.muninn_cuda_strerror.argtypes = [C.c_int]
        L.muninn_cuda_parse_metric.argtypes = [C.c_char_p]
        L.muninn_cuda_index_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(_vp)]
        L.muninn_cuda_index_destroy.argtypes = [_vp]
        L.muninn_cuda_index_destroy.restype = None
        L.muninn_cuda_last_error.restype = C.c_char_p
        L.muninn_cuda_last_error.argtypes = [_vp]
        L.muninn_cuda_index_upload.argtypes = [_vp, C.c_int64, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int32, C.c_int32]
        L.muninn_cuda_search_batch.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp]
        L.muninn_cuda_search_batch_device.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp]
        L.muninn_cuda_search_batch_device_static.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp]
        L.muninn_cuda_static_scratch_limits.argtypes = [C.POINTER(C.c_int), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.muninn_cuda_set_stream.argtypes = [_vp, _vp]
        L.muninn_cuda_last_kernel_ms.restype = C.c_float
        L.muninn_cuda_last_kernel_ms.argtypes = [_vp]
        L.muninn_cuda_launch_count.restype = C.c_int64
        L.muninn_cuda_launch_count.argtypes = [_vp]
        L.muninn_cuda_set_tuning.argtypes = [_vp, C.c_int, C.c_int]
        L.muninn_cuda_exact_knn.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int64, _vp, _vp, _vp]
        L.muninn_cuda_exact_knn_device.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int64, _vp, _vp, _vp]
        L.muninn_cuda_topk_merge_device.argtypes = [_vp, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp]
        L.muninn_cuda_distance_batch.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int64, _f32, _f32, _f32]
        L.muninn_cuda_set_exact_mode.argtypes = [_vp, C.c_int, C.c_int]
        L.muninn_cuda_exact_stats.argtypes = [_vp, C.POINTER(ExactStats)]
        L.muninn_cuda_exact_filter_probe.argtypes = [_vp, _vp, C.c_int, C.c_int64, _vp]
        L.muninn_cuda_index_patch.argtypes = [_vp, C.c_int64, C.c_int32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int32, C.c_int32]
        L.muninn_cuda_gather_ceiling.argtypes = [C.c_int, C.c_int, C.c_int64, C.POINTER(C.c_float)]
        L.muninn_cuda_build_device.argtypes = [_vp, _vp]
        L.muninn_cuda_last_team_width.argtypes = [_vp]
        L.muninn_cuda_last_kernel_gen.argtypes = [_vp]
        L.muninn_cuda_set_option.argtypes = [_vp, C.c_char_p, C.c_double]
        L.muninn_cuda_index_upload_rows.argtypes = [_vp, C.c_int64, _vp]
        L.muninn_cuda_exact_knn_sharded_device.argtypes = [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int64, _vp, _vp, _vp]
        L.muninn_cuda_comm_unique_id.argtypes = [_vp]
        L.muninn_cuda_comm_create.argtypes = [C.c_int, C.c_int, C.c_int, _vp, C.POINTER(_vp)]
        L.muninn_cuda_comm_destroy.argtypes = [_vp]
        L.muninn_cuda_comm_destroy.restype = None
        L.muninn_cuda_topk_pack_keys_device.argtypes = [_vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp]
        L.muninn_cuda_topk_merge_keys_device.argtypes = [_vp, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp]
        _lib = L
    return _lib


def device_count() -> int:
    return lib().muninn_cuda_device_count()


def parse_metric(name: str) -> int:
    return lib().muninn_cuda_parse_metric(name.encode())


def distance_batch(metric: int, a, b, device: int = 0) -> np.ndarray:
    """metric(a[i], b[i]) on the GPU with the reference's summation order."""
    a = np.ascontiguousarray(np.atleast_2d(a), np.float32)
    b = np.ascontiguousarray(np.atleast_2d(b), np.float32)
    assert a.shape == b.shape
    out = np.empty(a.shape[0], np.float32)
    rc = lib().muninn_cuda_distance_batch(device, metric, a.shape[1], a.shape[0], a, b, out)
    if rc != 0:
        raise MuninnCudaError(rc, lib().muninn_cuda_strerror(rc).decode())
    return out


def gather_ceiling(row_bytes: int, slab_bytes: int = 1 << 30, device: int = 0) -> float:
    """GB/s of independent random row gathers on this device (roofline diagnostic)."""
    out = C.c_float(0.0)
    rc = lib().muninn_cuda_gather_ceiling(device, row_bytes, slab_bytes, C.byref(out))
    if rc != 0:
        raise MuninnCudaError(rc, lib().muninn_cuda_strerror(rc).decode())
    return float(out.value)


def static_scratch_limits() -> dict:
    """Limits of HnswIndexCuda.search_device_static (muninn_cuda_static_scratch_limits; no counterpart in the reference)."""
    nq, nodes, nbytes = C.c_int(0), C.c_int64(0), C.c_int64(0)
    lib().muninn_cuda_static_scratch_limits(C.byref(nq), C.byref(nodes), C.byref(nbytes))
    return {"max_nq": nq.value, "max_nodes": nodes.value, "bytes": nbytes.value}


def comm_unique_id() -> bytes:
    """ncclGetUniqueId through the library (rank 0), to be shipped to the other ranks."""
    buf = C.create_string_buffer(128)
    rc = lib().muninn_cuda_comm_unique_id(buf)
    if rc != 0:
        raise MuninnCudaError(rc, lib().muninn_cuda_last_error(None).decode())
    return buf.raw


def comm_create(device: int, world: int, rank: int, uid: bytes) -> int:
    """ncclCommInitRank through the library (collective); returns the communicator as an integer address."""
    out = _vp()
    rc = lib().muninn_cuda_comm_create(device, world, rank, C.create_string_buffer(uid, 128), C.byref(out))
    if rc != 0:
        raise MuninnCudaError(rc, lib().muninn_cuda_last_error(None).decode())
    return out.value


def comm_destroy(comm: int):
    lib().muninn_cuda_comm_destroy(comm)


def _ptr(a: np.ndarray | None):
    return None if a is None else a.ctypes.data_as(_vp)


class HnswIndexCuda:
    """Device-resident hnsw_index graph + batched search (one handle = one GPU)."""

    def __init__(self, dim: int, metric: int = 0, M: int = 16, device: int = 0):
        self.L = lib()
        h = _vp()
        rc = self.L.muninn_cuda_index_create(device, dim, metric, M, C.byref(h))
        if rc != 0:
            raise MuninnCudaError(rc, self.L.muninn_cuda_last_error(None).decode())
        self.h, self.dim, self.metric, self.M, self.device = h, dim, metric, M, device
        self.n = 0

    # -- lifecycle ---------------------------------------------------------------------
    def close(self):
        if getattr(self, "h", None):
            self.L.muninn_cuda_index_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc: int):
        if rc != 0:
            raise MuninnCudaError(rc, self.L.muninn_cuda_last_error(self.h).decode())

    @classmethod
    def from_graph(cls, g: FlatGraph, device: int = 0) -> "HnswIndexCuda":
        idx = cls(g.dim, g.metric, g.M, device)
        idx.upload(g)
        return idx

    def upload(self, g: FlatGraph):
        """load_index_from_shadow → device (/root/reference/src/hnsw_vtab.c:286-341)."""
        g.validate()
        assert g.dim == self.dim and g.metric == self.metric and g.M == self.M
        self._check(self.L.muninn_cuda_index_upload(self.h, g.n, _ptr(g.ids), _ptr(g.levels), _ptr(g.deleted),
                                                    _ptr(g.vectors), _ptr(g.list_offsets), _ptr(g.targets),
                                                    g.entry_slot, g.max_level))
        self.n = g.n

    def upload_rows(self, rows_ptr: int, n: int):
        """Rows only (exact k-NN over a shard): `rows_ptr` is the address of [n × dim] fp32 in host or device memory."""
        self._check(self.L.muninn_cuda_index_upload_rows(self.h, n, rows_ptr))
        self.n = n

    def patch(self, g: FlatGraph, slots):
        """Send only the records of `slots` (every new node included) of the new state `g`: what hnsw_insert /
        hnsw_delete changed (/root/reference/src/hnsw_algo.c:520-666, :717-805)."""
        g.validate()
        self.patch_unchecked(g, slots)

    def patch_unchecked(self, g: FlatGraph, slots):
        """patch() without the host-side FlatGraph.validate(): the library's own checks are the only gate (tests)."""
        slots = np.ascontiguousarray(slots, np.int32)
        first = g.first_list()
        offs, tg = [0], []
        for s_ in slots:
            for l in range(int(g.levels[s_]) + 1):
                li = int(first[s_]) + l
                tg.append(g.targets[g.list_offsets[li]:g.list_offsets[li + 1]])
                offs.append(offs[-1] + len(tg[-1]))
        offs = np.asarray(offs, np.int64)
        tg = np.ascontiguousarray(np.concatenate(tg) if tg else np.empty(0), np.int32)
        ids = np.ascontiguousarray(g.ids[slots]); lv = np.ascontiguousarray(g.levels[slots])
        dl = np.ascontiguousarray(g.deleted[slots]); vec = np.ascontiguousarray(g.vectors[slots])
        self._check(self.L.muninn_cuda_index_patch(self.h, g.n, len(slots), _ptr(slots), _ptr(ids), _ptr(lv), _ptr(dl),
                                                   _ptr(vec), _ptr(offs), _ptr(tg), g.entry_slot, g.max_level))
        self.n = g.n

    def patch_records(self, n_total: int, slots, ids, levels, deleted, vectors_ptr, list_offsets, targets,
                      entry_slot: int, max_level: int):
        """muninn_cuda_index_patch with caller-assembled records (the device-assisted builder). `vectors_ptr` is the
        address of record 0's vector; only the records of NEW nodes are read from it (include/muninn_cuda.h)."""
        slots = np.ascontiguousarray(slots, np.int32)
        ids = np.ascontiguousarray(ids, np.int64); levels = np.ascontiguousarray(levels, np.int32)
        deleted = None if deleted is None else np.ascontiguousarray(deleted, np.uint8)
        list_offsets = np.ascontiguousarray(list_offsets, np.int64); targets = np.ascontiguousarray(targets, np.int32)
        self._check(self.L.muninn_cuda_index_patch(self.h, n_total, len(slots), _ptr(slots), _ptr(ids), _ptr(levels),
                                                   _ptr(deleted), vectors_ptr, _ptr(list_offsets), _ptr(targets),
                                                   entry_slot, max_level))
        self.n = n_total

    # -- search ------------------------------------------------------------------------
    def search(self, queries, k: int, ef: int | None = None, want_stats: bool = False):
        """Batched hnsw_search (/root/reference/src/hnsw_algo.c:670-704). ef defaults to 2k, the
        vtab's default (/root/reference/src/hnsw_vtab.c:591). Returns ids, dist, count[, stats]."""
        q = np.ascontiguousarray(np.atleast_2d(queries), np.float32)
        if q.shape[1] != self.dim:
            raise ValueError(f"hnsw_index: expected {self.dim}-dim vector ({self.dim * 4} bytes), "
                             f"got {q.shape[1] * 4} bytes")  # src/hnsw_vtab.c:595-597
        nq = q.shape[0]
        ef = 2 * k if ef is None else ef
        ids = np.empty((nq, k), np.int64)
        dist = np.empty((nq, k), np.float32)
        cnt = np.empty(nq, np.int32)
        stats = np.zeros(nq, STATS_DTYPE) if want_stats else None
        self._check(self.L.muninn_cuda_search_batch(self.h, _ptr(q), nq, k, ef, _ptr(ids), _ptr(dist), _ptr(cnt),
                                                    _ptr(stats)))
        return (ids, dist, cnt, stats) if want_stats else (ids, dist, cnt)

    def search_into(self, q_ptr: int, nq: int, k: int, ef: int, ids_ptr: int, dist_ptr: int, cnt_ptr: int,
                    stats_ptr: int | None = None):
        """Host-buffer entry with caller-owned (e.g. pinned) memory: raw addresses."""
        self._check(self.L.muninn_cuda_search_batch(self.h, q_ptr, nq, k, ef, ids_ptr, dist_ptr, cnt_ptr, stats_ptr))

    def search_device(self, q_ptr: int, nq: int, k: int, ef: int, ids_ptr: int, dist_ptr: int, cnt_ptr: int,
                      stats_ptr: int | None = None):
        """Device-resident buffers (raw device addresses), asynchronous on the handle's stream."""
        self._check(self.L.muninn_cuda_search_batch_device(self.h, q_ptr, nq, k, ef, ids_ptr, dist_ptr, cnt_ptr,
                                                           stats_ptr))

    def search_device_static(self, q_ptr: int, nq: int, k: int, ef: int, ids_ptr: int, dist_ptr: int, cnt_ptr: int,
                             stats_ptr: int | None = None):
        """search_device on the library's fixed device scratch: allocates nothing (include/muninn_cuda.h,
        muninn_cuda_search_batch_device_static; no counterpart in the reference). Calls on one device must not overlap."""
        self._check(self.L.muninn_cuda_search_batch_device_static(self.h, q_ptr, nq, k, ef, ids_ptr, dist_ptr, cnt_ptr,
                                                                  stats_ptr))

    def exact_knn(self, queries, k: int, index_base: int = 0):
        """Brute force over the uploaded rows, ascending (distance, row index)
        (/root/reference/test/test_hnsw_algo.c:10-34). Returns row indices, not rowids."""
        q = np.ascontiguousarray(np.atleast_2d(queries), np.float32)
        nq = q.shape[0]
        idx = np.empty((nq, k), np.int64)
        dist = np.empty((nq, k), np.float32)
        cnt = np.empty(nq, np.int32)
        self._check(self.L.muninn_cuda_exact_knn(self.h, _ptr(q), nq, k, index_base, _ptr(idx), _ptr(dist), _ptr(cnt)))
        return idx, dist, cnt

    def exact_knn_device(self, q_ptr: int, nq: int, k: int, index_base: int, idx_ptr: int, dist_ptr: int, cnt_ptr: int):
        self._check(self.L.muninn_cuda_exact_knn_device(self.h, q_ptr, nq, k, index_base, idx_ptr, dist_ptr, cnt_ptr))

    def set_exact_mode(self, mode: int = EXACT_TC_AUTO, tile_rows: int = 0):
        """How exact k-NN prunes its n·nq distance evaluations (results are identical in every mode)."""
        self._check(self.L.muninn_cuda_set_exact_mode(self.h, mode, tile_rows))

    def exact_stats(self) -> dict:
        st = ExactStats()
        self._check(self.L.muninn_cuda_exact_stats(self.h, C.byref(st)))
        return {f: getattr(st, f) for f, _ in ExactStats._fields_}

    def exact_filter_probe(self, queries, rows: int) -> np.ndarray:
        """The tensor-core filter's approximate distances for rows [0, rows) (error-bound tests)."""
        q = np.ascontiguousarray(np.atleast_2d(queries), np.float32)
        out = np.empty((q.shape[0], rows), np.float32)
        self._check(self.L.muninn_cuda_exact_filter_probe(self.h, _ptr(q), q.shape[0], rows, _ptr(out)))
        return out

    def topk_merge_device(self, parts: int, nq: int, k: int, idx_parts: int, dist_parts: int, cnt_parts: int,
                          idx_ptr: int, dist_ptr: int, cnt_ptr: int):
        self._check(self.L.muninn_cuda_topk_merge_device(self.h, parts, nq, k, idx_parts, dist_parts, cnt_parts,
                                                         idx_ptr, dist_ptr, cnt_ptr))

    def exact_knn_sharded_device(self, comm_ptr: int, q_ptr: int, nq: int, k: int, index_base: int, idx_ptr: int,
                                 dist_ptr: int, cnt_ptr: int):
        """This rank's shard scan + ONE ncclAllGather of 64-bit keys + merge. Device addresses."""
        self._check(self.L.muninn_cuda_exact_knn_sharded_device(self.h, comm_ptr, q_ptr, nq, k, index_base, idx_ptr,
                                                                dist_ptr, cnt_ptr))

    def topk_pack_keys_device(self, nq: int, k: int, idx_ptr: int, dist_ptr: int, cnt_ptr: int, keys_ptr: int):
        self._check(self.L.muninn_cuda_topk_pack_keys_device(self.h, nq, k, idx_ptr, dist_ptr, cnt_ptr, keys_ptr))

    def topk_merge_keys_device(self, parts: int, nq: int, k: int, keys_ptr: int, idx_ptr: int, dist_ptr: int, cnt_ptr: int):
        self._check(self.L.muninn_cuda_topk_merge_keys_device(self.h, parts, nq, k, keys_ptr, idx_ptr, dist_ptr, cnt_ptr))

    # -- plumbing ----------------------------------------------------------------------
    def set_stream(self, stream_ptr: int | None):
        self._check(self.L.muninn_cuda_set_stream(self.h, stream_ptr))

    def set_tuning(self, warps_per_query: int = 0, teams_per_sm: int = 0):
        self._check(self.L.muninn_cuda_set_tuning(self.h, warps_per_query, teams_per_sm))

    def set_option(self, name: str, value: float):
        self._check(self.L.muninn_cuda_set_option(self.h, name.encode(), float(value)))

    def last_team_width(self) -> int:
        return int(self.L.muninn_cuda_last_team_width(self.h))

    def last_kernel_gen(self) -> int:
        return int(self.L.muninn_cuda_last_kernel_gen(self.h))

    def last_kernel_ms(self) -> float:
        return float(self.L.muninn_cuda_last_kernel_ms(self.h))

    def launch_count(self) -> int:
        return int(self.L.muninn_cuda_launch_count(self.h))
