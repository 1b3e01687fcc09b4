"""Host-side logic that needs no GPU: the graph builder against the real reference, the
shadow-table reader/writer, and the C-ABI surface of libmuninn_cuda."""
import ctypes
import os
import re
import sqlite3
import sys

import numpy as np
import pytest

from sqlite_muninn_b200 import cuda
from sqlite_muninn_b200.builder import build_graph
from sqlite_muninn_b200.shadow import load_from_shadow, write_shadow
from tests.conftest import ROOT, uniform
from tests.helpers import load_golden


@pytest.mark.parametrize("metric,M,dim,n,efc", [(0, 8, 16, 1500, 60), (1, 16, 24, 700, 100), (2, 6, 10, 600, 40),
                                                (0, 4, 128, 900, 40)])
def test_builder_exact_mode_equals_reference(reference, metric, M, dim, n, efc):
    """csrc/hnsw_build.c (batch=0) builds the graph hnsw_insert builds: entry point, levels and
    ordered neighbour lists (src/hnsw_algo.c:520-666)."""
    base = uniform(n + 7, n, dim)
    if metric == 2:
        base /= np.linalg.norm(base, axis=1, keepdims=True)
    ri = reference.build(base, metric=metric, M=M, efc=efc)
    assert build_graph(base, metric, M, efc, sort_lists=False).same_structure(ri.export())
    ri.sort_lists()
    assert build_graph(base, metric, M, efc, sort_lists=True).same_structure(ri.export())


def test_builder_exact_mode_equals_reference_at_50k():
    """the exact mode of csrc/hnsw_build.c against the UNMODIFIED reference's hnsw_insert at 50 000 x 128,
    M=16, ef_construction=200 (bench.py's config-2 distribution). The reference build takes five minutes, so its graph
    is pinned as a digest (entry point, max level, every level, every neighbour list in insertion order and in
    shadow-table order) by tests/golden/make_build_fixture.py; this test rebuilds with the product's builder."""
    import json
    import bench
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    from make_build_fixture import graph_digest
    fx = json.load(open(os.path.join(ROOT, "tests", "golden", "build_50kx128_M16_efc200.json")))
    base = bench.make_vectors(fx["n"], fx["dim"], 11)
    g = build_graph(base, 0, fx["M"], fx["ef_construction"], sort_lists=False)
    assert (g.entry_slot, g.max_level, int(g.targets.shape[0])) == (fx["entry_slot"], fx["max_level"], fx["n_edges"])
    assert np.bincount(g.levels).tolist() == fx["level_histogram"]
    for s_, lst in fx["level0_lists_insertion_order"].items():
        assert [int(t) for t in g.neighbours(int(s_), 0)] == lst, s_
    assert graph_digest(g) == fx["digest_insertion_order"]
    gs = build_graph(base, 0, fx["M"], fx["ef_construction"], sort_lists=True)
    assert graph_digest(gs) == fx["digest_sorted_lists"]


def test_builder_batch_mode_is_deterministic_and_searchable(oracle):
    base = uniform(11, 6000, 32)
    g1 = build_graph(base, 0, 8, 40, batch=256, warm=512, threads=4)
    g2 = build_graph(base, 0, 8, 40, batch=256, warm=512, threads=2)
    assert g1.same_structure(g2)
    q = uniform(12, 50, 32)
    ids, _, _, _, _ = oracle.graph(g1).search(q, 10, 200)
    truth, _, _, _ = oracle.exact_knn(base, q, 10, 0)
    recall = np.mean([len(set(a) & set(b)) / 10 for a, b in zip(ids, truth)])
    assert recall > 0.8


def test_shadow_roundtrip(tmp_path):
    g, _, _, _ = load_golden("vtab_cosine_300x24")
    conn = sqlite3.connect(str(tmp_path / "x.db"))
    write_shadow(conn, "t", g)
    g2 = load_from_shadow(conn, "t")
    assert g.same_structure(g2) and np.array_equal(g.vectors, g2.vectors) and np.array_equal(g.deleted, g2.deleted)
    assert (g2.dim, g2.metric, g2.M, g2.ef_construction) == (g.dim, g.metric, g.M, g.ef_construction)


def test_shadow_host_c_reader_and_writer(tmp_path, reference):
    """csrc/shadow_io.c (libmuninn_host, against the SQLite C API): (1) it reads what the Python statement of the
    format writes and vice versa — same graph, same bytes; (2) it reads what the UNMODIFIED reference persisted through
    its own INSERT / DELETE statements exactly as the reference reloads it; (3) the unmodified extension opens, queries
    and extends a table the C writer stored and registered."""
    from oracle.bindings import REF_EXT_SO
    from sqlite_muninn_b200.shadow import load_from_shadow_c, write_shadow_c
    g, _, _, _ = load_golden("vtab_l2_delete_200x8")  # deleted rows, dangling-free lists
    # (1) C writer → Python reader, Python writer → C reader
    pa, pb = str(tmp_path / "a.db"), str(tmp_path / "b.db")
    write_shadow_c(pa, "t", g)
    ga = load_from_shadow(sqlite3.connect(pa), "t")
    conn = sqlite3.connect(pb)
    write_shadow(conn, "t", g)
    conn.close()
    gb = load_from_shadow_c(pb, "t")
    for x in (ga, gb):
        assert g.same_structure(x) and np.array_equal(g.vectors, x.vectors) and np.array_equal(g.deleted, x.deleted)
        assert np.array_equal(g.ids, x.ids) and (x.dim, x.metric, x.M, x.ef_construction) == (g.dim, g.metric, g.M, g.ef_construction)
    rows = lambda p, tbl, by: sqlite3.connect(p).execute(f"SELECT * FROM {tbl} ORDER BY {by}").fetchall()
    for tbl, by in (("t_config", "1"), ("t_nodes", "1"), ("t_edges", "1, 3, 2")):
        assert rows(pa, tbl, by) == rows(pb, tbl, by), tbl
    if not os.path.exists(REF_EXT_SO):
        pytest.skip("reference extension not built")
    # (2) what the reference itself persisted (INSERTs + DELETEs through its vtab), read back by the C loader
    pc = str(tmp_path / "c.db")
    conn = sqlite3.connect(pc)
    conn.enable_load_extension(True)
    conn.load_extension(REF_EXT_SO, entrypoint="sqlite3_muninn_init")
    conn.execute("CREATE VIRTUAL TABLE t USING hnsw_index(dimensions=12, metric='cosine', m=6, ef_construction=40)")
    base = uniform(61, 300, 12)
    for i, v in enumerate(base):
        conn.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (3 * i + 2, v.tobytes()))
    for i in range(0, 300, 9):
        conn.execute("DELETE FROM t WHERE rowid = ?", (3 * i + 2,))
    conn.commit()
    gp = load_from_shadow(conn, "t")
    conn.close()
    gc = load_from_shadow_c(pc, "t")
    assert gp.same_structure(gc) and np.array_equal(gp.vectors, gc.vectors) and np.array_equal(gp.deleted, gc.deleted)
    assert np.array_equal(gp.ids, gc.ids) and gc.deleted.sum() == len(range(0, 300, 9))
    # (3) the unmodified extension on a table the C writer stored and registered
    pd = str(tmp_path / "d.db")
    gb2 = build_graph(uniform(62, 400, 12), 0, 6, 40, ids=np.arange(400, dtype=np.int64) * 2 + 1)
    write_shadow_c(pd, "v", gb2, register=True)
    conn = sqlite3.connect(pd)
    conn.enable_load_extension(True)
    conn.load_extension(REF_EXT_SO, entrypoint="sqlite3_muninn_init")
    q = uniform(63, 12, 12)
    got = [conn.execute("SELECT rowid FROM v WHERE vector MATCH ? AND k = 5 AND ef_search = 40", (x.tobytes(),)).fetchall() for x in q]
    from oracle.bindings import Oracle
    o_ids, _, o_cnt, _, _ = Oracle().graph(gb2).search(q, 5, 40)
    assert [[r[0] for r in rows_] for rows_ in got] == [list(o_ids[i, :o_cnt[i]]) for i in range(len(q))]
    conn.execute("INSERT INTO v (rowid, vector) VALUES (?, ?)", (100001, q[0].tobytes()))  # the reference extends it
    assert conn.execute("SELECT rowid FROM v WHERE vector MATCH ? AND k = 1", (q[0].tobytes(),)).fetchall() == [(100001,)]
    conn.close()


def test_reference_vtab_opens_our_shadow_tables(tmp_path, oracle):
    """A graph built and written here is a valid hnsw_index for the UNMODIFIED reference
    extension, and its MATCH results equal the oracle's on that graph."""
    from oracle.bindings import REF_EXT_SO
    if not os.path.exists(REF_EXT_SO):
        pytest.skip("reference extension not built")
    base = uniform(21, 500, 16)
    g = build_graph(base, 0, 8, 64)
    g.ids[:] = np.arange(500) * 3 + 1
    path = str(tmp_path / "ours.db")
    conn = sqlite3.connect(path)
    write_shadow(conn, "t", g)
    conn.close()
    conn = sqlite3.connect(path)
    conn.enable_load_extension(True)
    conn.load_extension(REF_EXT_SO, entrypoint="sqlite3_muninn_init")
    # the module table in sqlite_master is what makes the shadow tables a vtab
    conn.execute("PRAGMA writable_schema=ON")
    conn.execute("INSERT INTO sqlite_master(type,name,tbl_name,rootpage,sql) VALUES('table','t','t',0,"
                 "'CREATE VIRTUAL TABLE t USING hnsw_index(dimensions=16, metric=''l2'', m=8, ef_construction=64)')")
    conn.commit()
    conn.close()
    conn = sqlite3.connect(path)
    conn.enable_load_extension(True)
    conn.load_extension(REF_EXT_SO, entrypoint="sqlite3_muninn_init")
    q = uniform(22, 20, 16)
    og = oracle.graph(g)
    ids, dist, cnt, _, _ = og.search(q, 5, 30)
    for i in range(20):
        rows = conn.execute("SELECT rowid, distance FROM t WHERE vector MATCH ? AND k = 5 AND ef_search = 30",
                            (q[i].tobytes(),)).fetchall()
        assert [r[0] for r in rows] == list(ids[i, :cnt[i]])
        assert [r[1] for r in rows] == [float(x) for x in dist[i, :cnt[i]]]


def test_abi_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "muninn_cuda.h")).read()
    declared = set(re.findall(r"\b(muninn_cuda_[a-z0-9_]+)\s*\(", header))
    assert declared == set(cuda.ABI_SYMBOLS)
    L = ctypes.CDLL(cuda.CUDA_SO)
    for sym in declared:
        assert hasattr(L, sym), sym


def test_host_library_exports_every_declared_symbol():
    from sqlite_muninn_b200.builder import HOST_SO
    header = open(os.path.join(ROOT, "include", "muninn_build.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)  # the usage sketch in the comment is not a declaration
    declared = set(re.findall(r"\b(muninn_build_[a-z0-9_]+)\s*\(", header))
    assert len(declared) == 18
    L = ctypes.CDLL(HOST_SO)
    for sym in declared:
        assert hasattr(L, sym), sym
    exported = set(re.findall(rb"\bmuninn_build_[a-z0-9_]+", open(HOST_SO, "rb").read()))
    assert {e.decode() for e in exported} <= declared | {"muninn_build.h"}, "an exported entry point is missing from the header"


def test_abi_host_only_calls():
    L = cuda.lib()
    assert L.muninn_cuda_version().startswith(b"muninn-b200")
    assert cuda.parse_metric("l2") == 0 and cuda.parse_metric("cosine") == 1
    assert cuda.parse_metric("inner_product") == 2 and cuda.parse_metric("hamming") == -1
    assert L.muninn_cuda_strerror(-2) == b"no CUDA device"
    # limits of the allocation-free search entry point: pure host code, as documented in include/muninn_cuda.h
    lim = cuda.static_scratch_limits()
    assert lim == {"max_nq": 65536, "max_nodes": 2097152 - 256, "bytes": 17301760}
    hdr = open(os.path.join(ROOT, "include", "muninn_cuda.h")).read()
    assert "17 301 760 bytes" in hdr and "65 536 queries" in hdr and "2 097 119 nodes" in hdr


def test_no_cpu_fallback_without_device():
    if cuda.device_count() > 0:
        pytest.skip("a device is present")
    with pytest.raises(cuda.MuninnCudaError) as e:
        cuda.HnswIndexCuda(8, 0, 8)
    assert e.value.code == -2


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "sqlite_muninn_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".c", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("the CPU oracle", "").replace("CPU oracle's", ""), f


def test_patched_and_reference_extensions_stay_distinct_in_one_process(tmp_path):
    """SQLite dlopens extensions RTLD_GLOBAL, so two builds of the same sources would interpose on
    each other (the second one's init registering the first one's module): both are linked
    -Bsymbolic. Proof, without a GPU: the patched extension's KNN plan fails loudly (no device, no
    CPU path) while the unmodified reference, loaded after it, still answers on the CPU."""
    from oracle.bindings import REF_EXT_SO
    cuda_ext = os.path.join(ROOT, "integration", "_build", "muninn_cuda_ext.so")
    if not (os.path.exists(cuda_ext) and os.path.exists(REF_EXT_SO)):
        pytest.skip("extensions not built")
    assert b"muninn_glue_search" in open(cuda_ext, "rb").read()
    assert b"muninn_glue_search" not in open(REF_EXT_SO, "rb").read()
    if cuda.device_count() > 0:
        pytest.skip("a device is present: covered by tests/test_gpu_vtab.py")
    conns = []
    for i, ext in enumerate((cuda_ext, REF_EXT_SO)):
        c = sqlite3.connect(str(tmp_path / f"{i}.db"))
        c.enable_load_extension(True)
        c.load_extension(ext, entrypoint="sqlite3_muninn_init")
        c.execute("CREATE VIRTUAL TABLE t USING hnsw_index(dimensions=4, metric='l2')")
        for r in range(1, 20):
            c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (r, np.float32([r, 0, 0, 1]).tobytes()))
        conns.append(c)
    q = np.float32([3.2, 0, 0, 1]).tobytes()
    sql = "SELECT rowid FROM t WHERE vector MATCH ? AND k = 2"
    assert [r[0] for r in conns[1].execute(sql, (q,)).fetchall()] == [3, 4]
    with pytest.raises(sqlite3.OperationalError, match="no CUDA device"):
        conns[0].execute(sql, (q,)).fetchall()


def test_batch_tvf_argument_errors_need_no_device(tmp_path):
    """hnsw_knn_batch (integration/muninn_batch_tvf.c): argument checking and the empty-index case run on
    the host; a non-empty index without a device fails loudly."""
    cuda_ext = os.path.join(ROOT, "integration", "_build", "muninn_cuda_ext.so")
    if not os.path.exists(cuda_ext):
        pytest.skip("integration/_build/muninn_cuda_ext.so not built")
    c = sqlite3.connect(str(tmp_path / "b.db"))
    c.enable_load_extension(True)
    c.load_extension(cuda_ext, entrypoint="sqlite3_muninn_init")
    c.execute("CREATE VIRTUAL TABLE t USING hnsw_index(dimensions=4, metric='l2')")
    q = np.float32([[1, 0, 0, 1], [2, 0, 0, 1]]).tobytes()
    assert c.execute("SELECT * FROM hnsw_knn_batch('t', ?, 3)", (q,)).fetchall() == []  # src/hnsw_algo.c:671
    assert c.execute("SELECT * FROM hnsw_knn_batch('t', ?, 3, 50)", (b"",)).fetchall() == []
    c.execute("INSERT INTO t (rowid, vector) VALUES (1, ?)", (np.float32([1, 0, 0, 1]).tobytes(),))
    c.execute("CREATE TABLE plain(vector, k, ef_search)")
    for sql, args, msg in (("SELECT * FROM hnsw_knn_batch('t', ?, 3)", (q[:-4],), "whole number of 4-dim vectors"),
                           ("SELECT * FROM hnsw_knn_batch('nope', ?, 3)", (q,), "no such table"),
                           ("SELECT * FROM hnsw_knn_batch('plain', ?, 3)", (q,), "not an hnsw_index table"),
                           ("SELECT * FROM hnsw_knn_batch('t', ?)", (q,), "missing argument"),
                           ("SELECT * FROM hnsw_knn_batch('t', ?, 0)", (q,), "k >= 1"),
                           ("SELECT rowid FROM t WHERE vector MATCH ? AND k = 2", (None,), "expected 4-dim vector")):
        with pytest.raises(sqlite3.OperationalError, match=msg):
            c.execute(sql, args).fetchall()
    if cuda.device_count() == 0:
        with pytest.raises(sqlite3.OperationalError, match="no CUDA device"):
            c.execute("SELECT * FROM hnsw_knn_batch('t', ?, 3)", (q,)).fetchall()


def test_bulk_loader_output_is_a_table_of_the_unmodified_reference(tmp_path, oracle):
    """tools/bulk_load.py (host searches here; the GPU variant builds the same graph): the next connection with the
    unmodified extension loaded sees a populated hnsw_index and answers MATCH like the oracle on that graph."""
    from oracle.bindings import REF_EXT_SO
    if not os.path.exists(REF_EXT_SO):
        pytest.skip("reference extension not built")
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bulk_load
    base = uniform(31, 700, 12)
    np.save(tmp_path / "v.npy", base)
    db = str(tmp_path / "bulk.db")
    bulk_load.main([str(tmp_path / "v.npy"), db, "docs", "--metric", "cosine", "--m", "8", "--ef-construction", "48",
                    "--batch", "64", "--host", "--with-distance"])
    g = build_graph(base, 1, 8, 48, ids=np.arange(1, 701, dtype=np.int64), batch=64)
    conn = sqlite3.connect(db)
    conn.enable_load_extension(True)
    conn.load_extension(REF_EXT_SO, entrypoint="sqlite3_muninn_init")
    assert conn.execute("SELECT count(*) FROM docs_nodes").fetchone()[0] == 700
    q = uniform(32, 10, 12)
    ids, dist, cnt, _, _ = oracle.graph(g).search(q, 5, 40)
    for i in range(10):
        rows = conn.execute("SELECT rowid, distance FROM docs WHERE vector MATCH ? AND k = 5 AND ef_search = 40",
                            (q[i].tobytes(),)).fetchall()
        assert [r[0] for r in rows] == list(ids[i, :cnt[i]])
        assert [r[1] for r in rows] == [float(x) for x in dist[i, :cnt[i]]]
    # and it is a live table: the reference's own INSERT extends it
    conn.execute("INSERT INTO docs(rowid, vector) VALUES (?, ?)", (1000, q[0].tobytes()))
    row = conn.execute("SELECT rowid FROM docs WHERE vector MATCH ? AND k = 1", (q[0].tobytes(),)).fetchone()
    assert row[0] == 1000


@pytest.mark.parametrize("dim", [1, 3, 7, 10, 131])
@pytest.mark.parametrize("metric", [0, 1, 2])
def test_builder_avx2_path_changes_nothing(metric, dim):
    """The builder's two-rows-per-call AVX2 distances run the reference's 4-lane SSE chain in each register half:
    same graphs as the scalar-pair path, row tails (dim % 4 != 0) included, in exact and in batch mode."""
    from sqlite_muninn_b200.builder import host_lib
    L = host_lib()
    if L.muninn_build_set_avx2(1) != 1:
        L.muninn_build_set_avx2(-1)
        pytest.skip("host CPU has no AVX2")
    try:
        x = np.random.default_rng(dim * 3 + metric).standard_normal((900, dim)).astype(np.float32)
        got = [build_graph(x, metric, 8, 40, batch=b, warm=150, sort_lists=False, threads=2) for b in (0, 96)]
        L.muninn_build_set_avx2(0)
        want = [build_graph(x, metric, 8, 40, batch=b, warm=150, sort_lists=False, threads=2) for b in (0, 96)]
        assert got[0].same_structure(want[0]) and got[1].same_structure(want[1])
    finally:
        L.muninn_build_set_avx2(-1)


def test_c_bulk_build_example_compiles_and_refuses_without_a_device(tmp_path):
    """integration/bulk_build_example.c: the two public headers are valid C together, the call signatures match the
    libraries, and without a GPU the program stops at index creation (no host fallback)."""
    import subprocess
    exe = str(tmp_path / "bulk_build_example")
    csrc = os.path.join(ROOT, "sqlite_muninn_b200", "csrc")
    cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
    subprocess.run([cc, "-O2", "-std=c11", "-D_POSIX_C_SOURCE=200809L", "-Wall", "-Wextra", "-Werror",
                    "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "integration", "bulk_build_example.c"),
                    "-L" + csrc, "-lmuninn_host", "-lmuninn_cuda", "-Wl,-rpath," + csrc, "-lm", "-o", exe], check=True)
    if cuda.device_count() > 0:
        pytest.skip("a device is present: the example would run a full build")
    out = subprocess.run([exe, "1000", "16"], capture_output=True, text=True)
    assert out.returncode == 2 and "no CUDA device" in out.stderr
