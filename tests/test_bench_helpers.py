"""CPU: the pieces of bench.py the reported numbers rest on — the algorithmic-bytes formula of SURVEY.md §8(d),
recall as `benchmarks/harness/treatments/vss.py:96-102` computes it, the workload generators' determinism, and
that the oracle's per-query counters (what the GPU must reproduce) feed the formula."""
import os

import numpy as np
import pytest

import bench
from tests.conftest import ROOT, uniform


def test_algorithmic_bytes_formula():
    stats = np.zeros(3, dtype=[("n_dist", np.int64), ("n_expand", np.int32), ("n_greedy_steps", np.int32)])
    stats["n_dist"] = [1650, 1600, 10]
    stats["n_expand"] = [66, 64, 1]
    stats["n_greedy_steps"] = [5, 4, 0]
    dim, M, k = 128, 16, 10
    want = sum(nd * dim * 4 + ne * 2 * M * 4 + ng * M * 4 + dim * 4 + k * 12
               for nd, ne, ng in zip(stats["n_dist"], stats["n_expand"], stats["n_greedy_steps"]))
    assert bench.algorithmic_bytes(stats, dim, M, k) == float(want)
    # the SURVEY anchor: ef=64 at 10k×128, M=16 → 1,650·512 + 66·128 ≈ 0.85 MB per query
    one = bench.algorithmic_bytes(stats[:1], dim, M, k)
    assert abs(one - (1650 * 512 + 66 * 128)) < 2048


def test_recall_at_k():
    ids = np.array([[1, 2, 3, 4], [5, 6, 7, -1]])
    truth = np.array([[4, 3, 9, 1], [5, 8, 6, 7]])
    assert bench.recall_at_k(ids, truth) == (3 / 4 + 3 / 4) / 2


def test_workloads_are_deterministic_and_shaped():
    for kind in ("latent16", "latent16n", "spectrum"):
        a = bench.make_vectors(500, 96, 7, kind=kind)
        b = bench.make_vectors(500, 96, 7, kind=kind)
        assert a.dtype == np.float32 and a.shape == (500, 96) and a.tobytes() == b.tobytes()
        assert bench.make_vectors(500, 96, 8, kind=kind).tobytes() != a.tobytes()
    n = np.linalg.norm(bench.make_vectors(300, 96, 1, kind="latent16n"), axis=1)
    assert np.allclose(n, 1.0, atol=1e-5)  # config 4 data is L2-normalised (inner product)


def test_oracle_counters_feed_the_formula(oracle):
    from sqlite_muninn_b200.builder import build_graph
    base, q = uniform(1, 800, 16), uniform(2, 20, 16)
    g = build_graph(base, 0, 8, 60)
    _, _, cnt, stats, _ = oracle.graph(g).search(q, 5, 40)
    assert (cnt == 5).all()
    b = bench.algorithmic_bytes(stats, 16, 8, 5)
    assert b > 20 * 40 * 16 * 4  # at least ef distances per query


def test_reference_arm_prints_the_contract_line(tmp_path):
    """`bench.py --impl reference` (the unmodified reference on the host cores) at a reduced size: one JSON line with
    the keys the driver reads. Needs oracle/_ref (built where /root/reference is present; prebuilt on the GPU box)."""
    import json
    import subprocess
    import sys
    from oracle.bindings import have_reference
    if not have_reference():
        pytest.skip("oracle/_ref not built")
    env = dict(os.environ, MUNINN_BENCH_CACHE=str(tmp_path))
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--n", "6000", "--nq", "200",
                          "--efc", "40", "--steps", "2", "--warmup", "1", "--ef", "32"], capture_output=True, text=True,
                         env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"].startswith("HNSW QPS") and d["unit"] == "queries/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0 and "workload" in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["sample"] and cb["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0
    assert d["scaling"] == "strong" and d["config"]["recall_at_10"] is not None and cb["cpu_model"]


def test_reference_arm_never_loads_the_cuda_library(tmp_path):
    """The reference arm is the CPU path only: libmuninn_cuda.so is built but must not be mapped into the process."""
    import subprocess
    import sys
    code = ("import sys, os; sys.argv=['bench.py','--impl','reference','--n','3000','--nq','64','--efc','30','--steps','1',"
            "'--warmup','0','--ef','16']; import runpy; runpy.run_path(os.path.join(%r,'bench.py'), run_name='__main__');"
            "maps=open('/proc/self/maps').read(); print('CUDA_LOADED' if 'libmuninn_cuda' in maps else 'CUDA_NOT_LOADED')" % ROOT)
    env = dict(os.environ, MUNINN_BENCH_CACHE=str(tmp_path))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "CUDA_NOT_LOADED" in out.stdout


def test_committed_dram_traffic_matches_the_kernel_sources_as_shipped():
    """`roofline.traffic` comes from ncu captures keyed by the hash of the search-kernel sources: a kernel edit without a
    fresh capture would silently turn it into null. The tree as committed must carry records for its own sources, for both
    launches the bench reports (the 10k batch and the 80k-query job)."""
    import json
    sha = bench.kernel_source_sha()
    recs = json.load(open(os.path.join(ROOT, "profiles", "search_traffic.json")))
    mine = {r["nq"]: r for r in recs if r["kernel_src_sha"] == sha and r["config"] == 2 and r["ef"] == 64}
    assert set(mine) >= {10000, 80000}, f"no ncu traffic record for kernel sources {sha}: run tools/capture_final.sh + tools/ncu_traffic.py"
    for r in mine.values():
        assert r["dram_bytes"] == r["dram_bytes_read"] + r["dram_bytes_write"] > 0
        assert os.path.exists(os.path.join(ROOT, r["source"])), r["source"]
    assert bench.ncu_traffic(2, 64, 10000) == mine[10000]["dram_bytes"]


def test_documents_cite_files_that_exist():
    """Every profile / tool / source file that DESIGN.md, INTEGRATION.md and the READMEs name in backticks is in the tree
    (paths of the reference — src/, test/, benchmarks/, pytests/ — are not ours to check)."""
    import glob
    import re
    where = ["", "profiles/", "profiles/r2_bench/", "profiles/r2_ncu/", "profiles/r1_bench/", "sqlite_muninn_b200/",
             "sqlite_muninn_b200/csrc/", "tools/", "integration/", "include/", "tests/", "tests/golden/", "oracle/"]
    reference_files = {"hnsw_algo.c", "priority_queue.c", "vec_math.c", "hnsw_vtab.c", "muninn.c"}
    missing = []
    for doc in ("DESIGN.md", "INTEGRATION.md", "README.md", os.path.join("profiles", "README.md")):
        text = open(os.path.join(ROOT, doc)).read()
        for m in re.findall(r"`([A-Za-z0-9_./*-]+\.(?:json|jsonl|csv|log|py|sh|cuh|cu|c|h|md))`", text):
            if m.startswith(("src/", "test/", "benchmarks/", "pytests/")) or m in reference_files or "/_" in m or m.startswith("r1_"):
                continue
            if not any(glob.glob(os.path.join(ROOT, w + m)) for w in where):
                missing.append((doc, m))
    assert not missing, missing
