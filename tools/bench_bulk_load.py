"""Insert throughput on the host: the unmodified reference's INSERT loop vs tools/bulk_load.py, BASELINE.json config 1
(10k x 128 uniform[-1,1), L2, M=16, ef_construction=200; seeds 1/2), and the recall@10 of each resulting table measured
through the UNMODIFIED extension (ef_search=64, 1 000 queries). CPU only: the GPU variant of the loader builds the same
graph as `--host`. The reference publishes 148-227 vectors/s for this step.

    python tools/bench_bulk_load.py [n] > profiles/r1_bench/bulk_load_cpu.json
"""
import json
import os
import sqlite3
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bulk_load  # noqa: E402

REF_EXT_SO = os.path.join(ROOT, "oracle", "_ref", "muninn_ref_ext.so")  # the unmodified reference extension: the arm timed

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000
dim, M, efc, nq = 128, 16, 200, 1000
base = np.random.default_rng(1).uniform(-1, 1, (n, dim)).astype(np.float32)
queries = np.random.default_rng(2).uniform(-1, 1, (nq, dim)).astype(np.float32)
# ground truth as the reference's harness computes it (numpy brute force, benchmarks/harness/treatments/vss.py:70-93)
d2 = (queries ** 2).sum(1)[:, None] - 2.0 * queries.astype(np.float64) @ base.astype(np.float64).T + (base.astype(np.float64) ** 2).sum(1)[None, :]
truth = np.argsort(d2, axis=1, kind="stable")[:, :10] + 1  # rowids 1..n


def connect(path):
    c = sqlite3.connect(path)
    c.enable_load_extension(True)
    c.load_extension(REF_EXT_SO, entrypoint="sqlite3_muninn_init")
    return c


def recall(path):
    c = connect(path)
    t0 = time.perf_counter()
    got = [[r[0] for r in c.execute("SELECT rowid FROM t WHERE vector MATCH ? AND k = 10 AND ef_search = 64",
                                    (q.tobytes(),))] for q in queries]
    dt = time.perf_counter() - t0
    c.close()
    return float(np.mean([len(set(a) & set(b)) / 10 for a, b in zip(got, truth.tolist())])), dt / nq * 1e3


out = {"workload": f"{n}x{dim} uniform[-1,1) L2, M={M}, ef_construction={efc}", "host_threads": os.cpu_count(), "arms": []}
with tempfile.TemporaryDirectory() as tmp:
    np.save(os.path.join(tmp, "v.npy"), base)
    # arm 1: the reference's own path, one INSERT per vector inside one transaction
    p = os.path.join(tmp, "ref.db")
    c = connect(p)
    c.execute(f"CREATE VIRTUAL TABLE t USING hnsw_index(dimensions={dim}, metric='l2', m={M}, ef_construction={efc})")
    t0 = time.perf_counter()
    c.executemany("INSERT INTO t(rowid, vector) VALUES (?, ?)", ((i + 1, base[i].tobytes()) for i in range(n)))
    c.commit()
    dt = time.perf_counter() - t0
    c.close()
    r, ms = recall(p)
    out["arms"].append({"arm": "unmodified reference, INSERT per vector", "seconds": round(dt, 2),
                        "vectors_per_s": round(n / dt, 1), "recall_at_10_ef64": round(r, 4), "ms_per_query_vtab": round(ms, 3)})
    for label, flags in (("bulk_load --exact (the reference's graph, bit for bit)", ["--exact"]),
                         ("bulk_load --host --batch 1024 (group-synchronous rule)", ["--host", "--batch", "1024"])):
        p = os.path.join(tmp, f"bulk{len(out['arms'])}.db")
        t0 = time.perf_counter()
        bulk_load.main([os.path.join(tmp, "v.npy"), p, "t", "--metric", "l2", "--m", str(M), "--ef-construction", str(efc)] + flags)
        dt = time.perf_counter() - t0
        r, ms = recall(p)
        out["arms"].append({"arm": label, "seconds": round(dt, 2), "vectors_per_s": round(n / dt, 1),
                            "recall_at_10_ef64": round(r, 4), "ms_per_query_vtab": round(ms, 3)})
print(json.dumps(out))
