"""hnsw_index shadow tables ⇄ FlatGraph.

The on-disk format is the reference's (/root/reference/src/hnsw_vtab.c:138-199):
  <name>_config(key TEXT PRIMARY KEY, value TEXT)   dimensions, metric (int), m, ef_construction,
                                                    entry_point, max_level
  <name>_nodes(id INTEGER PRIMARY KEY, vector BLOB, level, deleted)
  <name>_edges(source_id, target_id, level, distance, PRIMARY KEY(source_id, level, target_id)) WITHOUT ROWID

`load_from_shadow` reproduces load_index_from_shadow (src/hnsw_vtab.c:286-341): every stored
node, deleted ones included; an edge is kept only if its source exists and level <= the
source's level (:334); neighbour lists come back in ascending target_id order. A target id with
no node row is kept as -1 (the reference keeps the dangling id and skips it at search time).

`write_shadow` is the bulk writer: it stores a FlatGraph in exactly this
format so that the unmodified reference extension can xConnect to graphs built here.
"""
from __future__ import annotations

import itertools
import sqlite3

import numpy as np

from .graph import FlatGraph

_CHUNK = 262144  # rows per fetchmany


def read_config(conn: sqlite3.Connection, name: str) -> dict:
    rows = conn.execute(f'SELECT key, value FROM "{name}_config"').fetchall()
    return {k: v for k, v in rows}


def load_from_shadow(conn: sqlite3.Connection, name: str) -> FlatGraph:
    cfg = read_config(conn, name)
    dim, metric, M = int(cfg["dimensions"]), int(cfg["metric"]), int(cfg["m"])
    efc = int(cfg["ef_construction"])
    entry_id, max_level = int(cfg.get("entry_point", -1)), int(cfg.get("max_level", -1))
    # rows are streamed in chunks into flat numpy arrays: a fetchall() of 32M edge tuples would hold several GB of
    # Python objects at 1M x 32 edges
    n = int(conn.execute(f'SELECT count(*) FROM "{name}_nodes"').fetchone()[0])
    ids, levels, deleted = np.empty(n, np.int64), np.empty(n, np.int32), np.empty(n, np.uint8)
    vectors = np.empty((n, dim), np.float32)
    cur = conn.execute(f'SELECT id, level, deleted, vector FROM "{name}_nodes" ORDER BY id')
    at = 0
    while True:
        rows = cur.fetchmany(_CHUNK)
        if not rows:
            break
        m = len(rows)
        meta = np.fromiter(itertools.chain.from_iterable((r[0], r[1], 1 if r[2] else 0) for r in rows), np.int64, 3 * m)
        meta = meta.reshape(m, 3)
        ids[at:at + m], levels[at:at + m], deleted[at:at + m] = meta[:, 0], meta[:, 1], meta[:, 2]
        blob = b"".join(r[3] for r in rows)
        if len(blob) != m * dim * 4:
            raise ValueError(f"{name}_nodes: a vector blob is not {dim * 4} bytes")
        vectors[at:at + m] = np.frombuffer(blob, np.float32).reshape(m, dim)
        at += m
    first = np.zeros(n + 1, np.int64)
    np.cumsum(levels.astype(np.int64) + 1, out=first[1:])
    n_lists = int(first[-1])
    n_edges = int(conn.execute(f'SELECT count(*) FROM "{name}_edges"').fetchone()[0])
    edges = np.empty((n_edges, 3), np.int64)
    cur = conn.execute(f'SELECT source_id, level, target_id FROM "{name}_edges" ORDER BY source_id, level, target_id')
    at = 0
    while True:
        rows = cur.fetchmany(_CHUNK)
        if not rows:
            break
        m = len(rows)
        edges[at:at + m] = np.fromiter(itertools.chain.from_iterable(rows), np.int64, 3 * m).reshape(m, 3)
        at += m
    counts = np.zeros(n_lists, np.int64)
    if edges.shape[0]:
        src = np.searchsorted(ids, edges[:, 0])
        src_ok = (src < n) & (ids[np.minimum(src, n - 1)] == edges[:, 0])
        src = np.where(src_ok, src, 0)
        keep = src_ok & (edges[:, 1] <= levels[src]) & (edges[:, 1] >= 0)  # src/hnsw_vtab.c:334
        edges, src = edges[keep], src[keep]
        li = first[src] + edges[:, 1]
        tgt = np.searchsorted(ids, edges[:, 2])
        tgt_ok = (tgt < n) & (ids[np.minimum(tgt, n - 1)] == edges[:, 2])
        targets = np.where(tgt_ok, tgt, -1).astype(np.int32)
        np.add.at(counts, li, 1)  # rows are already ordered by (source, level, target) = list order
    else:
        targets = np.empty(0, np.int32)
    offs = np.zeros(n_lists + 1, np.int64)
    np.cumsum(counts, out=offs[1:])
    entry_slot = -1
    if entry_id != -1 and n:
        p = int(np.searchsorted(ids, entry_id))
        if p < n and ids[p] == entry_id:
            entry_slot = p
    return FlatGraph(dim, metric, M, efc, ids, levels, deleted, vectors, offs, targets, entry_slot, max_level)


def write_shadow(conn: sqlite3.Connection, name: str, g: FlatGraph, with_distance: bool = False) -> None:
    """Create <name>_config/_nodes/_edges holding `g` (bulk, one transaction). The `distance`
    column is informational in the reference (never read back, src/hnsw_vtab.c:322); it is
    written as 0.0 unless with_distance."""
    g.validate()
    cur = conn.cursor()
    cur.execute(f'CREATE TABLE IF NOT EXISTS "{name}_config" (key TEXT PRIMARY KEY, value TEXT NOT NULL)')
    cur.execute(f'CREATE TABLE IF NOT EXISTS "{name}_nodes" (id INTEGER PRIMARY KEY, vector BLOB NOT NULL, '
                f'level INTEGER NOT NULL, deleted INTEGER NOT NULL DEFAULT 0)')
    cur.execute(f'CREATE TABLE IF NOT EXISTS "{name}_edges" (source_id INTEGER NOT NULL, target_id INTEGER NOT NULL, '
                f'level INTEGER NOT NULL, distance REAL NOT NULL, PRIMARY KEY (source_id, level, target_id)) '
                f'WITHOUT ROWID')
    entry_id = int(g.ids[g.entry_slot]) if g.entry_slot >= 0 else -1
    cur.executemany(f'INSERT OR REPLACE INTO "{name}_config" (key, value) VALUES (?, ?)',
                    [("dimensions", str(g.dim)), ("metric", str(g.metric)), ("m", str(g.M)),
                     ("ef_construction", str(g.ef_construction)), ("entry_point", str(entry_id)),
                     ("max_level", str(g.max_level))])
    cur.executemany(f'INSERT INTO "{name}_nodes" (id, vector, level, deleted) VALUES (?, ?, ?, ?)',
                    ((int(g.ids[s]), g.vectors[s].tobytes(), int(g.levels[s]), int(g.deleted[s]))
                     for s in range(g.n)))
    # edge rows in bulk: list index → (source slot, level) with numpy, one executemany over plain Python lists
    # (a per-edge Python loop costs minutes at 1M × 32 edges)
    n_lists = len(g.list_offsets) - 1
    counts = np.diff(g.list_offsets).astype(np.int64)
    first = g.first_list()
    list_src = np.repeat(np.arange(g.n, dtype=np.int64), g.levels.astype(np.int64) + 1)
    list_lvl = np.arange(n_lists, dtype=np.int64) - first[list_src]
    e_src = np.repeat(list_src, counts)
    e_lvl = np.repeat(list_lvl, counts)
    e_tgt = g.targets.astype(np.int64)
    keep = e_tgt >= 0
    e_src, e_lvl, e_tgt = e_src[keep], e_lvl[keep], e_tgt[keep]
    id_src, id_tgt = g.ids[e_src], g.ids[e_tgt]
    for lo in range(0, int(e_src.shape[0]), _CHUNK * 4):  # chunks bound the Python-object footprint (32M edges at 1M x 32)
        hi = min(lo + _CHUNK * 4, int(e_src.shape[0]))
        if with_distance:  # informational: plain numpy sums, not the reference's summation order
            a, b = g.vectors[e_src[lo:hi]], g.vectors[e_tgt[lo:hi]]
            if g.metric == 0:
                dist = np.einsum("ij,ij->i", a - b, a - b)
            elif g.metric == 2:
                dist = -np.einsum("ij,ij->i", a, b)
            else:
                den = np.sqrt(np.einsum("ij,ij->i", a, a)) * np.sqrt(np.einsum("ij,ij->i", b, b))
                dist = np.where(den < 1e-30, 1.0, 1.0 - np.einsum("ij,ij->i", a, b) / np.maximum(den, 1e-30))
            dist = dist.astype(np.float64).tolist()
        else:
            dist = itertools.repeat(0.0, hi - lo)
        cur.executemany(f'INSERT INTO "{name}_edges" (source_id, target_id, level, distance) VALUES (?, ?, ?, ?)',
                        zip(id_src[lo:hi].tolist(), id_tgt[lo:hi].tolist(), e_lvl[lo:hi].tolist(), dist))
    # the reverse index (src/hnsw_vtab.c:177) is built once over the finished table, not row by row
    cur.execute(f'CREATE INDEX IF NOT EXISTS "{name}_edges_rev" ON "{name}_edges"(target_id, level)')
    conn.commit()


def register_vtab(conn: sqlite3.Connection, name: str, g: FlatGraph) -> None:
    """Make the shadow tables written by `write_shadow` a virtual table of this database without running xCreate
    (which would start an empty index): the schema row `CREATE VIRTUAL TABLE <name> USING hnsw_index(...)`
    (arguments as parsed at src/hnsw_vtab.c:80-134) is stored directly, so the next connection that has the
    extension loaded xConnects to the stored graph (src/hnsw_vtab.c:286-341)."""
    from .graph import METRIC_NAMES
    sql = (f"CREATE VIRTUAL TABLE {name} USING hnsw_index(dimensions={g.dim}, metric='{METRIC_NAMES[g.metric]}', "
           f"m={g.M}, ef_construction={g.ef_construction})")
    conn.execute("PRAGMA writable_schema=ON")
    conn.execute("INSERT INTO sqlite_master(type, name, tbl_name, rootpage, sql) VALUES('table', ?, ?, 0, ?)",
                 (name, name, sql))
    conn.execute("PRAGMA writable_schema=OFF")
    conn.commit()



# ───────────────────────── host C reader / writer (libmuninn_host: csrc/shadow_io.c) ─────────────────────────
# The product's bulk I/O (include/muninn_shadow.h): the Python functions above are the readable statement of the same
# format and the cross-check of the tests; tools and bench.py go through these.

import ctypes as _C


class _ShadowGraph(_C.Structure):
    """MuninnShadowGraph (include/muninn_shadow.h)."""
    _fields_ = [("dim", _C.c_int32), ("metric", _C.c_int32), ("M", _C.c_int32), ("ef_construction", _C.c_int32),
                ("n", _C.c_int64), ("n_lists", _C.c_int64), ("n_edges", _C.c_int64),
                ("ids", _C.POINTER(_C.c_int64)), ("levels", _C.POINTER(_C.c_int32)), ("deleted", _C.POINTER(_C.c_uint8)),
                ("vectors", _C.POINTER(_C.c_float)), ("list_offsets", _C.POINTER(_C.c_int64)),
                ("targets", _C.POINTER(_C.c_int32)), ("entry_slot", _C.c_int32), ("max_level", _C.c_int32)]


def _host():
    from .builder import host_lib
    L = host_lib()
    if not getattr(L, "_shadow_bound", False):
        L.muninn_shadow_load_path.argtypes = [_C.c_char_p, _C.c_char_p, _C.POINTER(_ShadowGraph), _C.POINTER(_C.c_char_p)]
        L.muninn_shadow_write_path.argtypes = [_C.c_char_p, _C.c_char_p, _C.POINTER(_ShadowGraph), _C.c_int, _C.POINTER(_C.c_char_p)]
        L.muninn_shadow_free.argtypes = [_C.POINTER(_ShadowGraph)]
        L.muninn_shadow_free_message.argtypes = [_C.c_char_p]
        L._shadow_bound = True
    return L


def _raise(L, rc, err):
    msg = err.value.decode() if err.value else f"muninn_shadow error {rc}"
    raise RuntimeError(msg)


def load_from_shadow_c(db_path: str, name: str) -> FlatGraph:
    """muninn_shadow_load_path: the stored index as a FlatGraph (arrays copied out of the C allocation)."""
    L = _host()
    sg, err = _ShadowGraph(), _C.c_char_p()
    rc = L.muninn_shadow_load_path(db_path.encode(), name.encode(), _C.byref(sg), _C.byref(err))
    if rc != 0:
        _raise(L, rc, err)
    try:
        n, nl, ne, dim = int(sg.n), int(sg.n_lists), int(sg.n_edges), int(sg.dim)
        arr = lambda p, cnt, dt: np.ctypeslib.as_array(p, shape=(max(cnt, 1),))[:cnt].astype(dt, copy=True)
        g = FlatGraph(dim, int(sg.metric), int(sg.M), int(sg.ef_construction), arr(sg.ids, n, np.int64),
                      arr(sg.levels, n, np.int32), arr(sg.deleted, n, np.uint8),
                      arr(sg.vectors, n * dim, np.float32).reshape(n, dim), arr(sg.list_offsets, nl + 1, np.int64),
                      arr(sg.targets, ne, np.int32), int(sg.entry_slot), int(sg.max_level))
    finally:
        L.muninn_shadow_free(_C.byref(sg))
    return g


def write_shadow_c(db_path: str, name: str, g: FlatGraph, register: bool = False) -> None:
    """muninn_shadow_write_path: one transaction, primary-key order; register=True also stores the
    `CREATE VIRTUAL TABLE … USING hnsw_index(…)` schema row so the unmodified extension xConnects to it."""
    g.validate()
    L = _host()
    ids = np.ascontiguousarray(g.ids, np.int64); levels = np.ascontiguousarray(g.levels, np.int32)
    deleted = np.ascontiguousarray(g.deleted, np.uint8); vec = np.ascontiguousarray(g.vectors, np.float32)
    offs = np.ascontiguousarray(g.list_offsets, np.int64); tg = np.ascontiguousarray(g.targets, np.int32)
    p = lambda a, t: a.ctypes.data_as(_C.POINTER(t))
    sg = _ShadowGraph(g.dim, g.metric, g.M, g.ef_construction, g.n, len(offs) - 1, len(tg), p(ids, _C.c_int64),
                      p(levels, _C.c_int32), p(deleted, _C.c_uint8), p(vec, _C.c_float), p(offs, _C.c_int64),
                      p(tg, _C.c_int32), g.entry_slot, g.max_level)
    err = _C.c_char_p()
    rc = L.muninn_shadow_write_path(db_path.encode(), name.encode(), _C.byref(sg), 1 if register else 0, _C.byref(err))
    if rc != 0:
        _raise(L, rc, err)
