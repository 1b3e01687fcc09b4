#!/usr/bin/env python
"""BASELINE.json configs[0] — the reference's own published methodology (benchmarks/harness/treatments/vss.py:288-319,
constants benchmarks/harness/common.py:42-46), on this box:

  10 000 x 128 fp32, i.i.d. uniform[-1, 1) (seed 1), L2, M = 16, ef_construction = 200; the graph is built by the
  UNMODIFIED reference through its own vtab (one `INSERT INTO t(rowid, vector)` per row); 1 000 queries (seed 2),
  k = 10, one `vector MATCH ? AND k = ? [AND ef_search = ?]` statement per query, ef = 20 (the vtab default 2k,
  src/hnsw_vtab.c:591) and ef = 64.

Arms, all on the SAME database file: the unmodified reference extension on the host CPU (oracle/_ref/muninn_ref_ext.so),
the patched extension with the KNN plan routed through libmuninn_cuda (integration/_build/muninn_cuda_ext.so), and
`hnsw_knn_batch` (one statement for all queries) in the patched extension. Reports ms/query, QPS, recall@10 against
exact k-NN (numpy, as the reference's harness computes it), and that the rows are identical. One JSON line.
"""
import argparse
import json
import os
import shutil
import sqlite3
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CUDA_EXT = os.path.join(ROOT, "integration", "_build", "muninn_cuda_ext.so")
REF_EXT = os.path.join(ROOT, "oracle", "_ref", "muninn_ref_ext.so")


def connect(path, ext):
    c = sqlite3.connect(path)
    c.enable_load_extension(True)
    c.load_extension(ext, entrypoint="sqlite3_muninn_init")
    return c


def same_rows(a, b):
    """identical up to the order inside groups of bit-equal distances (unspecified in the reference's heaps)"""
    return sorted(a, key=lambda r: (r[1], r[0])) == sorted(b, key=lambda r: (r[1], r[0]))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=10_000)
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--nq", type=int, default=1_000)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--db", default="/tmp/bench_config1.db")
    ap.add_argument("--repeats", type=int, default=5)
    a = ap.parse_args()
    M, efc = 16, 200
    base = np.random.default_rng(1).uniform(-1.0, 1.0, (a.n, a.dim)).astype(np.float32)
    q = np.random.default_rng(2).uniform(-1.0, 1.0, (a.nq, a.dim)).astype(np.float32)
    out = {"workload": f"{a.n}x{a.dim} fp32 uniform[-1,1) seeds 1/2, L2, M={M}, ef_construction={efc}, {a.nq} queries, k={a.k} "
                       "(BASELINE.json configs[0]); graph built by the unmodified reference through its vtab",
           "methodology": "one MATCH statement per query on one connection, wall clock over all queries (benchmarks/harness/treatments/vss.py:288-319); "
                          f"median of {a.repeats} repeats"}
    for p in (a.db, a.db + ".cuda"):
        if os.path.exists(p):
            os.remove(p)
    # ── the reference builds and persists the index (its own INSERT path: hnsw_insert + shadow-table upserts) ──
    rc = connect(a.db, REF_EXT)
    rc.execute(f"CREATE VIRTUAL TABLE t USING hnsw_index(dimensions={a.dim}, metric='l2', m={M}, ef_construction={efc})")
    t0 = time.perf_counter()
    for i in range(a.n):
        rc.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (i + 1, base[i].tobytes()))
    rc.commit()
    out["reference_insert_rate_vps"] = round(a.n / (time.perf_counter() - t0), 1)
    rc.close()
    shutil.copy(a.db, a.db + ".cuda")
    # exact ground truth as the reference's harness computes it (numpy argsort, 1-based rowids)
    d2 = (q * q).sum(1)[:, None] + (base * base).sum(1)[None, :] - 2.0 * (q.astype(np.float64) @ base.T.astype(np.float64))
    truth = [set((np.argsort(r)[: a.k] + 1).tolist()) for r in d2]
    rc = connect(a.db, REF_EXT)        # reopened: xConnect → load_index_from_shadow (the state parity is pinned on)
    cc = connect(a.db + ".cuda", CUDA_EXT)
    arms = {}
    for ef in (None, 64):
        tag = "ef=default(2k=20)" if ef is None else f"ef={ef}"
        sql = "SELECT rowid, distance FROM t WHERE vector MATCH ? AND k = ?" + ("" if ef is None else " AND ef_search = ?")
        args = (lambda v: (v.tobytes(), a.k)) if ef is None else (lambda v: (v.tobytes(), a.k, ef))
        res = {}
        for name, conn in (("reference_cpu", rc), ("patched_cuda", cc)):
            for i in range(20):
                conn.execute(sql, args(q[i])).fetchall()
            times, rows = [], None
            for _ in range(a.repeats):
                t0 = time.perf_counter()
                rows = [conn.execute(sql, args(q[i])).fetchall() for i in range(a.nq)]
                times.append(time.perf_counter() - t0)
            med = float(np.median(times))
            rec = float(np.mean([len({r[0] for r in rr} & t) / a.k for rr, t in zip(rows, truth)]))
            res[name] = {"ms_per_query": round(med / a.nq * 1e3, 4), "qps": round(a.nq / med, 1), "best_qps": round(a.nq / min(times), 1),
                         "recall_at_10": round(rec, 4)}
            res[name + "_rows"] = rows
        res["rows_identical"] = all(same_rows(x, y) for x, y in zip(res.pop("reference_cpu_rows"), res["patched_cuda_rows"]))
        # the batched form of the same plan: one statement, one launch
        blob = q.tobytes()
        bsql = "SELECT query, id, distance FROM hnsw_knn_batch('t', ?, ?, ?)"
        ef_eff = 2 * a.k if ef is None else ef
        cc.execute(bsql, (blob, a.k, ef_eff)).fetchall()
        times, brows = [], None
        for _ in range(a.repeats):
            t0 = time.perf_counter()
            brows = cc.execute(bsql, (blob, a.k, ef_eff)).fetchall()
            times.append(time.perf_counter() - t0)
        med = float(np.median(times))
        res["patched_cuda_knn_batch"] = {"ms_per_query": round(med / a.nq * 1e3, 5), "qps": round(a.nq / med, 1), "statement_ms": round(med * 1e3, 3),
                                         "rows_identical_to_per_query_statements": [(r[1], r[2]) for r in brows] == [r for rr in res["patched_cuda_rows"] for r in rr]}
        del res["patched_cuda_rows"]
        res["speedup_per_statement"] = round(res["patched_cuda"]["qps"] / res["reference_cpu"]["qps"], 2)
        res["speedup_batched"] = round(res["patched_cuda_knn_batch"]["qps"] / res["reference_cpu"]["qps"], 1)
        arms[tag] = res
    out["arms"] = arms
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                out["cpu_model"] = line.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    rc.close()
    cc.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
