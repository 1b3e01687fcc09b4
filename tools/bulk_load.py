"""Bulk loader (SURVEY.md §8f N1 + N2): vectors → an hnsw_index the reference extension opens.

    python tools/bulk_load.py vectors.npy out.db my_index --metric l2 --m 16 --ef-construction 200

The reference builds an index one `INSERT` at a time (≈ 1 000 row operations each, /root/reference/src/hnsw_vtab.c:237-283,
and a graph search whose visited set is a linear scan): hours at 1M rows. This builds the graph with the reference's
construction rule in group-synchronous form (csrc/hnsw_build.c), candidate searches on the GPU
(`builder.build_graph_device`; `--host` keeps them on the host threads), writes the three shadow tables in one
transaction (`shadow.write_shadow`) and registers the virtual table, so that

    SELECT rowid, distance FROM my_index WHERE vector MATCH ? AND k = 10

works on the next connection with the extension loaded — the unmodified one or the GPU-patched one.
`--exact` inserts sequentially instead and gives the graph the reference's own INSERT loop would give, bit for bit.
"""
import argparse
import json
import os
import sqlite3
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sqlite_muninn_b200.builder import build_graph, build_graph_device  # noqa: E402
from sqlite_muninn_b200.graph import METRICS  # noqa: E402
from sqlite_muninn_b200.shadow import register_vtab, write_shadow  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("vectors", help=".npy file, float32 [n, dim]")
    ap.add_argument("db")
    ap.add_argument("name", help="virtual table name")
    ap.add_argument("--ids", help=".npy file of int64 rowids, strictly ascending (default 1..n)")
    ap.add_argument("--metric", default="cosine", choices=sorted(METRICS))  # reference default, src/hnsw_vtab.c:68-73
    ap.add_argument("--m", type=int, default=16)
    ap.add_argument("--ef-construction", type=int, default=200)
    ap.add_argument("--batch", type=int, default=1024, help="insert group size")
    ap.add_argument("--host", action="store_true", help="candidate searches on the host threads instead of the GPU")
    ap.add_argument("--exact", action="store_true", help="sequential inserts: the reference's own graph (slow)")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--with-distance", action="store_true", help="fill _edges.distance (never read back by the reference)")
    a = ap.parse_args(argv)
    x = np.ascontiguousarray(np.load(a.vectors), np.float32)
    n = x.shape[0]
    ids = np.load(a.ids).astype(np.int64) if a.ids else np.arange(1, n + 1, dtype=np.int64)
    kw = dict(metric=METRICS[a.metric], M=a.m, ef_construction=a.ef_construction, ids=ids)
    t0 = time.perf_counter()
    st = {}
    if a.exact:
        g, how = build_graph(x, **kw), "exact sequential inserts (host)"
    elif a.host:
        g, how = build_graph(x, batch=a.batch, **kw), f"groups of {a.batch}, host searches"
    else:
        g, how = build_graph_device(x, batch=a.batch, device=a.device, stats=st, **kw), f"groups of {a.batch}, GPU searches"
    t1 = time.perf_counter()
    conn = sqlite3.connect(a.db)
    write_shadow(conn, a.name, g, with_distance=a.with_distance)
    register_vtab(conn, a.name, g)
    conn.close()
    t2 = time.perf_counter()
    print(json.dumps({"n": n, "dim": int(x.shape[1]), "metric": a.metric, "m": a.m, "ef_construction": a.ef_construction,
                      "build": how, "build_s": round(t1 - t0, 2), "write_s": round(t2 - t1, 2),
                      "max_level": g.max_level, "entry_point": int(g.ids[g.entry_slot]), **st}))


if __name__ == "__main__":
    main()
