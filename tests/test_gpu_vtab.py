"""`-m gpu`: the drop-in boundary. The reference's extension with its KNN plan routed through
libmuninn_cuda (integration/_build/muninn_cuda_ext.so: unchanged muninn.c, unchanged SQL
surface, hnsw_vtab.c with the one-line call-site edit of INTEGRATION.md) must return what the
UNMODIFIED reference extension returns for the same SQL, in every index state — freshly built,
after deletes, after close/reopen — and reproduce the committed vtab fixtures.
Mirrors /root/reference/pytests/test_hnsw_vtab.py."""
import os
import sqlite3

import numpy as np
import pytest

from tests.conftest import ROOT, uniform
from tests.helpers import load_golden

pytestmark = pytest.mark.gpu
CUDA_EXT = os.path.join(ROOT, "integration", "_build", "muninn_cuda_ext.so")
METRIC_SQL = {0: "l2", 1: "cosine", 2: "inner_product"}


def connect(path, ext):
    conn = sqlite3.connect(path)
    conn.enable_load_extension(True)
    conn.load_extension(ext, entrypoint="sqlite3_muninn_init")
    return conn


@pytest.fixture(scope="module")
def exts():
    from oracle.bindings import REF_EXT_SO
    if not os.path.exists(CUDA_EXT):
        pytest.skip("integration/_build/muninn_cuda_ext.so not built (needs /root/reference at build time)")
    return CUDA_EXT, (REF_EXT_SO if os.path.exists(REF_EXT_SO) else None)


def knn(conn, q, k, ef=None):
    if ef is None:
        return conn.execute("SELECT rowid, distance FROM t WHERE vector MATCH ? AND k = ?", (q.tobytes(), k)).fetchall()
    return conn.execute("SELECT rowid, distance FROM t WHERE vector MATCH ? AND k = ? AND ef_search = ?",
                        (q.tobytes(), k, ef)).fetchall()


def same_rows(a, b):
    """Equal up to the order inside groups of bit-equal distances: the reference's two binary heaps leave
    that order unspecified (test/test_priority_queue.c:85-99); the device orders such groups by slot
    (canonical tie order)."""
    assert [r[1] for r in a] == sorted(r[1] for r in a) and [r[1] for r in b] == sorted(r[1] for r in b)
    return sorted(a, key=lambda r: (r[1], r[0])) == sorted(b, key=lambda r: (r[1], r[0]))


@pytest.mark.parametrize("name", ["vtab_l2_400x16", "vtab_cosine_300x24", "vtab_ip_256x10", "vtab_l2_delete_200x8"])
def test_vtab_reproduces_reference_fixture(exts, tmp_path, name):
    """Rebuild the fixture's table with SQL INSERTs through the patched extension, reopen, query:
    rowids and (double) distances equal those the unmodified reference returned."""
    g, queries, k, cases = load_golden(name)
    path = str(tmp_path / "g.db")
    conn = connect(path, exts[0])
    conn.execute(f"CREATE VIRTUAL TABLE t USING hnsw_index(dimensions={g.dim}, metric='{METRIC_SQL[g.metric]}', "
                 f"m={g.M}, ef_construction={g.ef_construction})")
    for i in range(g.n):
        conn.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (int(g.ids[i]), g.vectors[i].tobytes()))
    for i in np.nonzero(g.deleted)[0]:
        conn.execute("DELETE FROM t WHERE rowid = ?", (int(g.ids[i]),))
    conn.commit()
    conn.close()
    conn = connect(path, exts[0])
    efs = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))["efs"]
    for ef_raw, c in zip(efs, cases):
        for qi, q in enumerate(queries):
            rows = knn(conn, q, k, None if ef_raw < 0 else int(ef_raw))
            n = int(c["cnt"][qi])
            assert [r[0] for r in rows] == list(c["ids"][qi, :n]), (name, qi)
            assert [r[1] for r in rows] == list(c["dist"][qi, :n]), (name, qi)
    conn.close()


def test_vtab_equals_unmodified_reference_in_every_state(exts, tmp_path):
    if exts[1] is None:
        pytest.skip("reference extension not built")
    base, q = uniform(31, 600, 12), uniform(32, 40, 12)
    conns = [connect(str(tmp_path / f"{i}.db"), e) for i, e in enumerate(exts)]
    for c in conns:
        c.execute("CREATE VIRTUAL TABLE t USING hnsw_index(dimensions=12, metric='l2', m=8, ef_construction=60)")
        assert knn(c, q[0], 3) == []  # empty table
        for i, v in enumerate(base[:400]):
            c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (i * 2 + 1, v.tobytes()))

    def same(tag):
        for qi in range(40):
            for k, ef in ((1, None), (5, None), (10, 64), (450, 500)):
                a, b = knn(conns[0], q[qi], k, ef), knn(conns[1], q[qi], k, ef)
                assert same_rows(a, b), (tag, qi, k, ef)

    same("fresh")  # list order = insertion order
    for c in conns:
        for i in range(0, 400, 7):
            c.execute("DELETE FROM t WHERE rowid = ?", (i * 2 + 1,))
    same("after delete")  # dirty flag → re-upload
    for c in conns:
        for i, v in enumerate(base[400:]):
            c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (10_000 + i, v.tobytes()))
    same("after more inserts")
    for i, c in enumerate(conns):
        c.commit()
        c.close()
    conns = [connect(str(tmp_path / f"{i}.db"), e) for i, e in enumerate(exts)]
    same("reopened")  # xConnect → load_index_from_shadow → upload
    # point lookup and vector column are served by the host copy, unchanged
    for c in conns:
        row = c.execute("SELECT vector FROM t WHERE rowid = 3").fetchone()
        assert np.array_equal(np.frombuffer(row[0], np.float32), base[1])
    for c in conns:
        c.close()


def test_vtab_heavy_deletion_keeps_answering(exts, tmp_path):
    """Delete ~75 % of a small-M table: hnsw_delete's reconnection step (src/hnsw_algo.c:780-781) appends to
    neighbour lists without a capacity check, so the reference ends up holding lists longer than 2M / M — and keeps
    answering from them. The device tables are sized from the longest stored list, a patch that outgrows them falls
    back to an upload, and every MATCH keeps returning what the unmodified extension returns."""
    if exts[1] is None:
        pytest.skip("reference extension not built")
    n, dim = 600, 8
    base, q = uniform(51, n, dim), uniform(42, 25, dim)  # the inputs of tests/test_gpu_patch.py's direct reproduction
    conns = [connect(str(tmp_path / f"hd{i}.db"), e) for i, e in enumerate(exts)]
    for c in conns:
        c.execute(f"CREATE VIRTUAL TABLE t USING hnsw_index(dimensions={dim}, metric='l2', m=4, ef_construction=40)")
        for i, v in enumerate(base):
            c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (i + 1, v.tobytes()))
    order = np.random.default_rng(3).permutation(n)[: int(0.75 * n)]  # 400 of these deletes leave a 5-entry level-1 list (M = 4)
    for step, i in enumerate(order):
        for c in conns:
            c.execute("DELETE FROM t WHERE rowid = ?", (int(i) + 1,))
        if step % 15 == 0 or step == len(order) - 1:  # searches interleaved with the deletes: patches, then uploads
            for qi in range(len(q)):
                for k, ef in ((5, None), (10, 40)):
                    a, b = knn(conns[0], q[qi], k, ef), knn(conns[1], q[qi], k, ef)
                    assert same_rows(a, b), (step, qi, k, ef)
    # (that the in-memory index really goes past its nominal capacity on exactly these inputs is asserted where it can be
    # seen — tests/test_gpu_patch.py::test_lists_longer_than_capacity_after_heavy_deletion exports the reference's index)
    for c in conns:
        c.commit()
        c.close()
    conns = [connect(str(tmp_path / f"hd{i}.db"), e) for i, e in enumerate(exts)]
    for qi in range(len(q)):  # reopened: load_index_from_shadow → upload of the over-full lists
        assert same_rows(knn(conns[0], q[qi], 10, 40), knn(conns[1], q[qi], 10, 40)), qi
    for c in conns:
        c.close()


def test_vtab_interleaved_writes_use_incremental_patches(exts, tmp_path):
    """INSERT / DELETE between searches: the glue sends muninn_cuda_index_patch records for the touched
    nodes (src/hnsw_algo.c:hnsw_insert / hnsw_delete) instead of re-uploading, falls back to one full
    upload when a rowid lands below the stored maximum, and the SQL results stay those of the
    unmodified reference extension after every step."""
    import ctypes
    if exts[1] is None:
        pytest.skip("reference extension not built")
    rng = np.random.default_rng(77)
    base, q = uniform(41, 900, 16), uniform(42, 12, 16)
    conns = [connect(str(tmp_path / f"i{i}.db"), e) for i, e in enumerate(exts)]
    glue = ctypes.CDLL(exts[0])
    for f in (glue.muninn_glue_total_patches, glue.muninn_glue_total_uploads):
        f.restype = ctypes.c_int64
    for c in conns:
        c.execute("CREATE VIRTUAL TABLE t USING hnsw_index(dimensions=16, metric='cosine', m=6, ef_construction=40)")
        for i, v in enumerate(base[:500]):
            c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (i + 1, v.tobytes()))

    def same(tag):
        for qi in range(len(q)):
            for k, ef in ((1, None), (10, None), (30, 80)):
                a, b = knn(conns[0], q[qi], k, ef), knn(conns[1], q[qi], k, ef)
                assert same_rows(a, b), (tag, qi, k, ef)

    same("fresh")
    p0, u0 = glue.muninn_glue_total_patches(), glue.muninn_glue_total_uploads()
    live, nxt = list(range(1, 501)), 501
    for step in range(12):
        dels = [live.pop(int(rng.integers(len(live)))) for _ in range(int(rng.integers(0, 4)))]
        ins = list(range(nxt, nxt + int(rng.integers(1, 9))))
        for c in conns:
            for r in dels:
                c.execute("DELETE FROM t WHERE rowid = ?", (r,))
            for r in ins:
                c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (r, base[r - 1].tobytes()))
        if step % 4 == 3:  # insert-then-delete inside one burst: a new node that is already a tombstone
            for c in conns:
                c.execute("DELETE FROM t WHERE rowid = ?", (ins[-1],))
            ins = ins[:-1]
        live += ins
        nxt += len(ins) + (step % 4 == 3)
        same(f"step {step}")
    assert glue.muninn_glue_total_patches() - p0 == 12, "every burst must reach the device as one patch"
    assert glue.muninn_glue_total_uploads() == u0, "no full upload while rowids only grow"
    # a rowid below the maximum moves slots: the glue must start over with one full upload
    for c in conns:
        c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (100_000, base[899].tobytes()))
        c.execute("DELETE FROM t WHERE rowid = ?", (live[3],))
    same("before the low rowid")
    for c in conns:
        c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (50_000, base[898].tobytes()))
    same("low rowid")
    assert glue.muninn_glue_total_uploads() == u0 + 1
    for c in conns:
        c.close()


def batch_rows(conn, queries, k, ef=None):
    """hnsw_knn_batch → per-query lists of (rowid, distance), the shape knn() returns"""
    if ef is None:
        rows = conn.execute("SELECT query, id, distance, rank FROM hnsw_knn_batch('t', ?, ?)", (queries.tobytes(), k)).fetchall()
    else:
        rows = conn.execute("SELECT query, id, distance, rank FROM hnsw_knn_batch('t', ?, ?, ?)",
                            (queries.tobytes(), k, ef)).fetchall()
    out = [[] for _ in range(len(queries))]
    for qi, rid, d, rank in rows:
        assert rank == len(out[qi])  # rows arrive grouped by query, in rank order
        out[qi].append((rid, d))
    return out


@pytest.mark.parametrize("metric,dim", [("l2", 16), ("cosine", 24), ("inner_product", 10)])
def test_batch_tvf_equals_per_query_sql_of_the_unmodified_reference(exts, tmp_path, metric, dim):
    """hnsw_knn_batch('t', <nq vectors>, k[, ef]) answers in one launch what the
    reference answers with one `vector MATCH ?` statement per query (src/llama_er.c:237-284)."""
    if exts[1] is None:
        pytest.skip("reference extension not built")
    base, q = uniform(51, 700, dim), uniform(52, 64, dim)
    conns = [connect(str(tmp_path / f"b{i}.db"), e) for i, e in enumerate(exts)]
    for c in conns:
        c.execute(f"CREATE VIRTUAL TABLE t USING hnsw_index(dimensions={dim}, metric='{metric}', m=8, ef_construction=50)")
    assert batch_rows(conns[0], q, 5) == [[] for _ in q]  # empty index: no rows
    for c in conns:
        for i, v in enumerate(base[:600]):
            c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (3 * i + 2, v.tobytes()))
        for i in range(0, 600, 11):
            c.execute("DELETE FROM t WHERE rowid = ?", (3 * i + 2,))

    def same(tag):
        for k, ef in ((1, None), (10, None), (10, 100), (40, 30), (700, 800)):
            got = batch_rows(conns[0], q, k, ef)
            for qi in range(len(q)):
                assert same_rows(got[qi], knn(conns[1], q[qi], k, ef)), (tag, qi, k, ef)
            assert got[5] == knn(conns[0], q[5], k, ef)  # and exactly the patched vtab's own single-query answer

    same("built")
    for c in conns:  # writes between batches reach the device as a patch
        for i, v in enumerate(base[600:]):
            c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (5000 + i, v.tobytes()))
        c.execute("DELETE FROM t WHERE rowid = ?", (5003,))
    same("after writes")
    # one query, and a blob SQLite hands over unaligned (sliced out of a larger one)
    assert same_rows(batch_rows(conns[0], q[:1], 7)[0], knn(conns[1], q[0], 7))
    odd = conns[0].execute("SELECT id FROM hnsw_knn_batch('t', substr(?, 2), 3)", (b"\x00" + q[:3].tobytes(),)).fetchall()
    assert [r[0] for r in odd] == [r[0] for qi in range(3) for r in knn(conns[0], q[qi], 3)]
    with pytest.raises(sqlite3.OperationalError, match="whole number of"):
        conns[0].execute("SELECT * FROM hnsw_knn_batch('t', ?, 3)", (q.tobytes()[:-4],)).fetchall()
    for c in conns:
        c.close()


def test_er_blocking_stage_through_the_batch_tvf(exts, tmp_path):
    """The ER blocking loop of src/llama_er.c:237-284 (per entity: KNN, drop self, distance threshold,
    de-duplicated (lo, hi) pairs keeping the smaller distance) written once with the reference's
    per-entity statement on the unmodified extension and once as ONE SQL statement over
    hnsw_knn_batch: the candidate pairs and their distances are identical."""
    if exts[1] is None:
        pytest.skip("reference extension not built")
    rng = np.random.default_rng(9)
    proto = uniform(61, 80, 32)
    ent = (proto[rng.integers(0, 80, 400)] + np.float32(0.15) * uniform(62, 400, 32)).astype(np.float32)
    conns = [connect(str(tmp_path / f"er{i}.db"), e) for i, e in enumerate(exts)]
    for c in conns:
        c.execute("CREATE VIRTUAL TABLE t USING hnsw_index(dimensions=32, metric='cosine', m=16, ef_construction=100)")
        for i, v in enumerate(ent):
            c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (i + 1, v.tobytes()))
    k, thr = 10, 0.05
    want = {}
    for i in range(len(ent)):  # the reference's loop
        for nid, dist in knn(conns[1], ent[i], k):
            if nid == i + 1 or dist > thr:
                continue
            key = (min(i + 1, nid), max(i + 1, nid))
            want[key] = min(dist, want.get(key, np.inf))
    got = conns[0].execute(
        "SELECT min(query + 1, id), max(query + 1, id), min(distance) FROM hnsw_knn_batch('t', ?, ?) "
        "WHERE id <> query + 1 AND distance <= ? GROUP BY 1, 2", (ent.tobytes(), k, thr)).fetchall()
    assert len(want) > 300
    assert {(a, b): d for a, b, d in got} == want
    for c in conns:
        c.close()


def test_vtab_connections_on_several_threads(exts, tmp_path):
    """Each thread owns a connection and a table (the reference's supported threading model); the glue's handle
    table is shared by all of them. Concurrent create / insert / search / drop must give every thread the answers
    a single-threaded run gives."""
    import threading
    base, q = uniform(71, 4 * 150, 8), uniform(72, 30, 8)

    def work(t, ext, out):
        c = connect(str(tmp_path / f"th{t}_{os.path.basename(ext)}.db"), ext)
        c.execute("CREATE VIRTUAL TABLE t USING hnsw_index(dimensions=8, metric='l2', m=6, ef_construction=40)")
        rows = []
        for i, v in enumerate(base[t * 150:(t + 1) * 150]):
            c.execute("INSERT INTO t (rowid, vector) VALUES (?, ?)", (i + 1, v.tobytes()))
            if i % 10 == 9:
                rows.append(knn(c, q[i % 30], 5))
        rows += [knn(c, qq, 8, 40) for qq in q]
        c.execute("DROP TABLE t")
        c.close()
        out[t] = rows

    got, want = {}, {}
    threads = [threading.Thread(target=work, args=(t, exts[0], got)) for t in range(4)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    for t in range(4):
        work(t, exts[1] if exts[1] else exts[0], want)
    assert set(got) == {0, 1, 2, 3}
    for t in range(4):
        assert len(got[t]) == len(want[t])
        assert all(same_rows(a, b) for a, b in zip(got[t], want[t])), t


def test_vtab_errors_and_drop(exts, tmp_path):
    conn = connect(str(tmp_path / "e.db"), exts[0])
    conn.execute("CREATE VIRTUAL TABLE t USING hnsw_index(dimensions=3, metric='l2')")
    conn.execute("INSERT INTO t (rowid, vector) VALUES (1, ?)", (np.float32([0, 0, 1]).tobytes(),))
    with pytest.raises(sqlite3.OperationalError, match="expected 3-dim"):  # src/hnsw_vtab.c:595-597
        knn(conn, np.float32([1, 2]), 1)
    assert knn(conn, np.float32([0, 0, 0.9]), 5)[0][0] == 1  # k > n
    with pytest.raises(sqlite3.OperationalError, match="unknown metric"):
        conn.execute("CREATE VIRTUAL TABLE bad USING hnsw_index(dimensions=4, metric='hamming')")
    conn.execute("DROP TABLE t")  # xDestroy releases the device handle
    tables = {r[0] for r in conn.execute("SELECT name FROM sqlite_master").fetchall()}
    assert not any(n.startswith("t_") for n in tables)
    conn.close()
