#!/usr/bin/env python
"""Measure the drop-in boundary at the SQL level (SURVEY.md §8d CPU baseline (i), §8f N3/N4):

  * per-query `vector MATCH ?` latency: patched extension (one launch per statement) vs the
    UNMODIFIED reference extension on the host (the reference's published methodology);
  * `hnsw_knn_batch` throughput for the same queries in one statement;
  * cost of keeping the device copy current: first search after a burst of SQL INSERT/DELETE
    (incremental muninn_cuda_index_patch) vs the first search after connect (flatten + full upload).

The graph is built by csrc/hnsw_build.c (batch mode), written to shadow tables by
sqlite_muninn_b200.shadow.write_shadow and opened by both extensions through xConnect, i.e. both
search the same persisted index. Prints one JSON line. Not part of bench.py's contract; its output
is kept under profiles/.
"""
import argparse
import ctypes
import json
import os
import sqlite3
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import make_vectors  # noqa: E402
from sqlite_muninn_b200.builder import build_graph  # noqa: E402
from sqlite_muninn_b200.shadow import write_shadow  # noqa: E402

CUDA_EXT = os.path.join(ROOT, "integration", "_build", "muninn_cuda_ext.so")
REF_EXT = os.path.join(ROOT, "oracle", "_ref", "muninn_ref_ext.so")


def connect(path, ext):
    c = sqlite3.connect(path)
    c.enable_load_extension(True)
    c.load_extension(ext, entrypoint="sqlite3_muninn_init")
    return c


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=200_000)
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--nq", type=int, default=10_000)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--ef", type=int, default=64)
    ap.add_argument("--ref-queries", type=int, default=300)
    ap.add_argument("--burst", type=int, default=100)
    ap.add_argument("--db", default="/tmp/bench_vtab.db")
    a = ap.parse_args()
    M, efc = 16, 200
    base = make_vectors(a.n + a.burst, a.dim, 11)
    q = make_vectors(a.nq, a.dim, 12)
    t0 = time.time()
    g = build_graph(base[:a.n], 0, M, efc, batch=1024)
    t_build = time.time() - t0
    g.ids[:] = np.arange(1, a.n + 1)
    if os.path.exists(a.db):
        os.remove(a.db)
    conn = sqlite3.connect(a.db)
    t0 = time.time()
    write_shadow(conn, "t", g)
    conn.execute("PRAGMA writable_schema=ON")
    conn.execute("INSERT INTO sqlite_master(type,name,tbl_name,rootpage,sql) VALUES('table','t','t',0,"
                 f"'CREATE VIRTUAL TABLE t USING hnsw_index(dimensions={a.dim}, metric=''l2'', m={M}, ef_construction={efc})')")
    conn.commit()
    conn.close()
    t_write = time.time() - t0
    out = {"workload": f"{a.n}x{a.dim} fp32 L2, M={M}, k={a.k}, ef={a.ef}, {a.nq} queries; shadow tables written in bulk, "
                       f"opened through xConnect by both extensions", "build_s": round(t_build, 2), "write_shadow_s": round(t_write, 2)}
    sql = "SELECT rowid, distance FROM t WHERE vector MATCH ? AND k = ? AND ef_search = ?"

    # unmodified reference on the host, one query per statement
    if os.path.exists(REF_EXT):
        import shutil
        ref_db = a.db + ".ref"
        shutil.copy(a.db, ref_db)
        t0 = time.time()
        rc = connect(ref_db, REF_EXT)
        rc.execute(sql, (q[0].tobytes(), a.k, a.ef)).fetchall()  # xConnect: load_index_from_shadow
        out["reference_connect_s"] = round(time.time() - t0, 2)
        t0 = time.time()
        ref_rows = [rc.execute(sql, (q[i].tobytes(), a.k, a.ef)).fetchall() for i in range(a.ref_queries)]
        dt = time.time() - t0
        out["reference_vtab_us_per_query"] = round(dt / a.ref_queries * 1e6, 1)
        out["reference_vtab_qps"] = round(a.ref_queries / dt, 1)
        rc.close()
    else:
        ref_rows = None

    glue = ctypes.CDLL(CUDA_EXT)
    glue.muninn_glue_total_patches.restype = glue.muninn_glue_total_uploads.restype = ctypes.c_int64
    t0 = time.time()
    c = connect(a.db, CUDA_EXT)
    c.execute("SELECT count(*) FROM t_config").fetchall()
    first = c.execute(sql, (q[0].tobytes(), a.k, a.ef)).fetchall()  # xConnect + flatten + upload + search
    out["patched_connect_plus_upload_s"] = round(time.time() - t0, 2)
    for i in range(50):
        c.execute(sql, (q[i].tobytes(), a.k, a.ef)).fetchall()
    t0 = time.time()
    rows = [c.execute(sql, (q[i].tobytes(), a.k, a.ef)).fetchall() for i in range(a.nq)]
    dt = time.time() - t0
    out["patched_vtab_us_per_query"] = round(dt / a.nq * 1e6, 1)
    out["patched_vtab_qps"] = round(a.nq / dt, 1)
    if ref_rows is not None:
        out["vtab_rows_identical_to_reference"] = all(rows[i] == ref_rows[i] for i in range(a.ref_queries))

    bsql = "SELECT query, id, distance FROM hnsw_knn_batch('t', ?, ?, ?)"
    blob = q.tobytes()
    c.execute(bsql, (blob, a.k, a.ef)).fetchall()
    best = 1e9
    for _ in range(5):
        t0 = time.time()
        brows = c.execute(bsql, (blob, a.k, a.ef)).fetchall()
        best = min(best, time.time() - t0)
    out["batch_tvf_ms"] = round(best * 1e3, 2)
    out["batch_tvf_qps"] = round(a.nq / best, 1)
    # the statement's cost without Python's row materialisation: count(*) still steps every row
    best = 1e9
    for _ in range(5):
        t0 = time.time()
        c.execute("SELECT count(*), sum(id) FROM hnsw_knn_batch('t', ?, ?, ?)", (blob, a.k, a.ef)).fetchall()
        best = min(best, time.time() - t0)
    out["batch_tvf_aggregate_ms"] = round(best * 1e3, 2)
    out["batch_tvf_aggregate_qps"] = round(a.nq / best, 1)
    flat = [(r[1], r[2]) for r in brows]
    out["batch_rows_identical_to_vtab"] = flat == [r for rr in rows for r in rr]

    # incremental update: a burst of INSERTs and DELETEs, then the next search
    p0, u0 = glue.muninn_glue_total_patches(), glue.muninn_glue_total_uploads()
    t0 = time.time()
    for i in range(a.burst):
        c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (a.n + 1 + i, base[a.n + i].tobytes()))
    for i in range(a.burst // 10):
        c.execute("DELETE FROM t WHERE rowid = ?", (7 * i + 3,))
    out["burst"] = f"{a.burst} INSERT + {a.burst // 10} DELETE"
    out["burst_sql_s"] = round(time.time() - t0, 3)
    t0 = time.time()
    c.execute(sql, (q[0].tobytes(), a.k, a.ef)).fetchall()
    out["first_search_after_burst_ms"] = round((time.time() - t0) * 1e3, 2)
    out["patches"], out["uploads"] = glue.muninn_glue_total_patches() - p0, glue.muninn_glue_total_uploads() - u0
    # the same burst answered by a full upload: a rowid below the maximum forces one
    c.execute("DELETE FROM t WHERE rowid = 5")
    c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (a.n + 1 + a.burst + 7, base[0].tobytes()))
    c.execute(sql, (q[0].tobytes(), a.k, a.ef)).fetchall()
    c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (a.n + 1 + a.burst + 3, base[1].tobytes()))
    t0 = time.time()
    c.execute(sql, (q[0].tobytes(), a.k, a.ef)).fetchall()
    out["first_search_after_full_upload_ms"] = round((time.time() - t0) * 1e3, 2)
    out["uploads_total"] = glue.muninn_glue_total_uploads() - u0
    c.rollback()
    c.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
