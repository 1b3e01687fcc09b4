"""Host-side cost of the group-synchronous graph build, phase by phase (CPU only).

Runs `build_graph` (host batch mode) and the same loop through the group interface with a stand-in index that
declines every search (so the host searches everything and the graph is identical), which exposes the per-phase
seconds `build_graph_device` reports: what the device-assisted build keeps on the host (warm-up inserts, request sort,
reverse edges + prune, record export) and what it hands to the GPU (the level-0 candidate searches).

    python tools/bench_host_build.py [n] [dim] > profiles/r1_bench/host_build_breakdown.json
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sqlite_muninn_b200.builder import build_graph, build_graph_device  # noqa: E402


class DecliningIndex:
    """upload / patch_records are no-ops; search answers count = -1 for every query: 'search it yourself'."""

    def upload(self, g):
        pass

    def patch_records(self, *a):
        pass

    def search(self, q, k, ef):
        nq = len(q)
        return np.zeros((nq, k), np.int64), np.zeros((nq, k), np.float32), np.full(nq, -1, np.int32)


n = int(sys.argv[1]) if len(sys.argv) > 1 else 150_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 128
rng = np.random.default_rng(1)
z = rng.standard_normal((n, 16)).astype(np.float32)
x = (z @ rng.standard_normal((16, dim)).astype(np.float32) + 0.05 * rng.standard_normal((n, dim))).astype(np.float32)
t = time.perf_counter()
g = build_graph(x, 0, 16, 200, batch=1024, warm=8192, sort_lists=False)
t_host = time.perf_counter() - t
st = {}
t = time.perf_counter()
g2 = build_graph_device(x, 0, 16, 200, batch=1024, warm=8192, sort_lists=False, index=DecliningIndex(), stats=st)
t_loop = time.perf_counter() - t
print(json.dumps({"workload": f"{n}x{dim} rank-16 latent + noise, L2, M=16, ef_construction=200, groups of 1024, warm 8192",
                  "host_threads": os.cpu_count(), "build_graph_s": round(t_host, 2),
                  "vectors_per_s": round(n / t_host), "group_loop_s": round(t_loop, 2),
                  "identical_graphs": bool(g.same_structure(g2)), "phases": st}))
