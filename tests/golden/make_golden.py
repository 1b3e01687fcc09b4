"""Generate tests/golden/*.npz from the UNMODIFIED reference (run in the build container only).

Two sources, both compiled by oracle/Makefile from /root/reference where it lies:
  oracle/_ref/muninn_ref_ext.so   the reference's SQLite extension, driven through stdlib sqlite3:
                                  CREATE VIRTUAL TABLE … USING hnsw_index, INSERT, close, reopen
                                  (xConnect → load_index_from_shadow), `vector MATCH ? AND k = ?`
  oracle/_ref/libmuninn_ref.so    hnsw_insert / hnsw_search called directly (ref_harness.c)

Each fixture stores the graph (as read back from the shadow tables, or exported from the live
index), the queries and the reference's own outputs. The reference has no golden search
vectors of its own, so these are the pinned outputs the CPU suite replays.

    python tests/golden/make_golden.py
"""
import os
import sqlite3
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle.bindings import REF_EXT_SO, Reference  # noqa: E402
from sqlite_muninn_b200.shadow import load_from_shadow  # noqa: E402
from tests.conftest import lcg_vectors  # noqa: E402

METRIC_SQL = {0: "l2", 1: "cosine", 2: "inner_product"}


def graph_fields(g, prefix="g_"):
    return {prefix + "dim": g.dim, prefix + "metric": g.metric, prefix + "M": g.M, prefix + "efc": g.ef_construction,
            prefix + "ids": g.ids, prefix + "levels": g.levels, prefix + "deleted": g.deleted,
            prefix + "vectors": g.vectors, prefix + "list_offsets": g.list_offsets, prefix + "targets": g.targets,
            prefix + "entry_slot": g.entry_slot, prefix + "max_level": g.max_level}


def connect(path):
    conn = sqlite3.connect(path)
    conn.enable_load_extension(True)
    conn.load_extension(REF_EXT_SO, entrypoint="sqlite3_muninn_init")
    return conn


def vtab_query(conn, q, k, ef):
    if ef is None:
        rows = conn.execute("SELECT rowid, distance FROM t WHERE vector MATCH ? AND k = ?", (q.tobytes(), k)).fetchall()
    else:
        rows = conn.execute("SELECT rowid, distance FROM t WHERE vector MATCH ? AND k = ? AND ef_search = ?",
                            (q.tobytes(), k, ef)).fetchall()
    return rows


def vtab_case(name, base, ids, queries, metric, M, efc, k, efs, delete=()):
    """Build through the vtab, close, reopen, query; read the shadow tables back."""
    tmp = tempfile.mkdtemp()
    path = os.path.join(tmp, "g.db")
    conn = connect(path)
    conn.execute(f"CREATE VIRTUAL TABLE t USING hnsw_index(dimensions={base.shape[1]}, metric='{METRIC_SQL[metric]}', "
                 f"m={M}, ef_construction={efc})")
    for i, v in zip(ids, base):
        conn.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (int(i), v.tobytes()))
    for d in delete:
        conn.execute("DELETE FROM t WHERE rowid = ?", (int(d),))
    conn.commit()
    conn.close()
    conn = connect(path)  # xConnect → load_index_from_shadow: the state parity is pinned on
    out = {}
    for ef in efs:
        tag = "def" if ef is None else str(ef)
        o_ids = np.full((len(queries), k), -1, np.int64)
        o_d = np.zeros((len(queries), k), np.float64)
        o_c = np.zeros(len(queries), np.int32)
        for qi, q in enumerate(queries):
            rows = vtab_query(conn, q, k, ef)
            o_c[qi] = len(rows)
            for j, (rid, dist) in enumerate(rows):
                o_ids[qi, j], o_d[qi, j] = rid, dist
        out[f"ids_{tag}"], out[f"dist_{tag}"], out[f"cnt_{tag}"] = o_ids, o_d, o_c
    g = load_from_shadow(conn, "t")
    conn.close()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), queries=queries, k=k,
                        efs=np.array([-1 if e is None else e for e in efs]), **graph_fields(g), **out)
    print(name, "n =", g.n, "max_level =", g.max_level)


def direct_case(name, base, ids, queries, metric, M, efc, k, efs, sort, delete=()):
    """hnsw_insert / hnsw_search called directly; sort=False keeps the freshly built list order."""
    ref = Reference()
    ri = ref.build(base, ids=ids, metric=metric, M=M, efc=efc)
    for d in delete:
        assert ri.delete(int(d)) == 0
    if sort:
        ri.sort_lists()
    g = ri.export()
    out = {}
    for ef in efs:
        o_ids, o_d, o_c, nd = ri.search(queries, k, ef)
        out[f"ids_{ef}"], out[f"dist_{ef}"], out[f"cnt_{ef}"], out[f"ndist_{ef}"] = o_ids, o_d, o_c, nd
    np.savez_compressed(os.path.join(HERE, name + ".npz"), queries=queries, k=k, efs=np.array(efs), **graph_fields(g),
                        **out)
    print(name, "n =", g.n, "max_level =", g.max_level)


def main():
    rng = np.random.default_rng(2024)
    # 1. vtab, L2, uniform — the published methodology (M, ef) at toy size
    base = rng.uniform(-1, 1, (400, 16)).astype(np.float32)
    q = rng.uniform(-1, 1, (48, 16)).astype(np.float32)
    vtab_case("vtab_l2_400x16", base, np.arange(1, 401), q, 0, 8, 64, 5, [None, 32, 100])
    # 2. vtab, cosine (default metric), gauss, sparse rowids
    base = rng.standard_normal((300, 24)).astype(np.float32)
    q = rng.standard_normal((32, 24)).astype(np.float32)
    vtab_case("vtab_cosine_300x24", base, np.arange(300) * 7 + 3, q, 1, 16, 100, 10, [None, 64])
    # 3. vtab, inner product on unit vectors, dim not a multiple of 4 (scalar tail of vec_math.c:91-94)
    base = rng.standard_normal((256, 10)).astype(np.float32)
    base /= np.linalg.norm(base, axis=1, keepdims=True)
    q = rng.standard_normal((32, 10)).astype(np.float32)
    vtab_case("vtab_ip_256x10", base, np.arange(256), q, 2, 6, 40, 4, [None, 40])
    # 4. vtab with deletes, then reopen (pytests/test_hnsw_vtab.py:195-222 at larger size)
    base = rng.uniform(-1, 1, (200, 8)).astype(np.float32)
    q = rng.uniform(-1, 1, (32, 8)).astype(np.float32)
    vtab_case("vtab_l2_delete_200x8", base, np.arange(200), q, 0, 8, 50, 5, [None, 64], delete=range(0, 200, 9))
    # 5. the reference's own recall test data: 50×8 LCG, M=8, efc=50, ef=64, k=5 (test/test_hnsw_algo.c:96-145)
    base = lcg_vectors(50, 8)
    direct_case("direct_lcg_50x8_fresh", base, np.arange(50), base[:20].copy(), 0, 8, 50, 5, [5, 64], sort=False)
    direct_case("direct_lcg_50x8_reloaded", base, np.arange(50), base[:20].copy(), 0, 8, 50, 5, [5, 64], sort=True)
    # 6. a graph with four levels, dim 128 (the headline row width), ef sweep incl. patience-prone small ef
    base = rng.uniform(-1, 1, (1500, 128)).astype(np.float32)
    q = rng.uniform(-1, 1, (40, 128)).astype(np.float32)
    direct_case("direct_l2_1500x128", base, np.arange(1500), q, 0, 4, 40, 10, [10, 20, 64, 128], sort=True)
    # 7. direct, after hnsw_delete with reconnection (src/hnsw_algo.c:717-805), live state
    base = rng.uniform(-1, 1, (300, 12)).astype(np.float32)
    q = rng.uniform(-1, 1, (32, 12)).astype(np.float32)
    direct_case("direct_l2_delete_300x12", base, np.arange(300), q, 0, 8, 60, 5, [10, 50], sort=True,
                delete=range(5, 300, 11))


if __name__ == "__main__":
    main()
