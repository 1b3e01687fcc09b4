#!/usr/bin/env python
"""Shadow-table bulk I/O in host C at size: the benchmark graph (bench.py config 2 by
default) goes through `<t>_config/_nodes/_edges` once — muninn_shadow_write_path → muninn_shadow_load_path
(libmuninn_host, csrc/shadow_io.c) — and comes back identical; optionally the UNMODIFIED reference extension opens the
same file (xConnect → its own load_index_from_shadow) and answers a MATCH. CPU only. One JSON line.

  python tools/bench_shadow_io.py [--config 2] [--n 1000000] [--reference]"""
import argparse
import json
import os
import sqlite3
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from sqlite_muninn_b200.shadow import load_from_shadow_c, write_shadow_c  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=2)
    ap.add_argument("--n", type=int, default=None)
    ap.add_argument("--db", default="/tmp/bench_shadow_io.db")
    ap.add_argument("--reference", action="store_true", help="also let the unmodified extension open the file and answer one MATCH")
    a = ap.parse_args()
    c = bench.cfg_of(a.config)
    if a.n:
        c.n = a.n
    g = bench.get_graph(c, 0, 1, lambda: None)
    g.ids[:] = np.arange(1, g.n + 1)
    if os.path.exists(a.db):
        os.remove(a.db)
    t0 = time.perf_counter()
    write_shadow_c(a.db, "t", g, register=True)
    t_w = time.perf_counter() - t0
    t0 = time.perf_counter()
    g2 = load_from_shadow_c(a.db, "t")
    t_r = time.perf_counter() - t0
    same = bool(g.same_structure(g2) and np.array_equal(g.vectors, g2.vectors) and np.array_equal(g.ids, g2.ids))
    out = {"workload": f"{c.name} {g.n}x{c.dim}, M={c.M}: graph of bench.py through the hnsw_index shadow tables (host C, SQLite C API)",
           "nodes": g.n, "edges": int((g.targets >= 0).sum()), "db_bytes": os.path.getsize(a.db),
           "muninn_shadow_write_s": round(t_w, 2), "muninn_shadow_load_s": round(t_r, 2),
           "write_rows_per_s": round((g.n + int((g.targets >= 0).sum())) / t_w), "round_trip_identical": same,
           "host_threads_used": 1}
    if a.reference:
        ext = os.path.join(ROOT, "oracle", "_ref", "muninn_ref_ext.so")
        conn = sqlite3.connect(a.db)
        conn.enable_load_extension(True)
        conn.load_extension(ext, entrypoint="sqlite3_muninn_init")
        q = bench.make_vectors(4, c.dim, c.seeds[1], kind=c.data)
        t0 = time.perf_counter()
        rows = conn.execute("SELECT rowid, distance FROM t WHERE vector MATCH ? AND k = 10 AND ef_search = 64", (q[0].tobytes(),)).fetchall()
        out["reference_xconnect_plus_first_match_s"] = round(time.perf_counter() - t0, 2)
        from oracle.bindings import Oracle
        o_ids, o_d, o_c, _, _ = Oracle().graph(g).search(q[:1], 10, 64)
        out["reference_rows_equal_oracle_on_same_graph"] = [r[0] for r in rows] == list(o_ids[0, :o_c[0]])
        conn.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
