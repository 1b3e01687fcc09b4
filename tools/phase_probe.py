#!/usr/bin/env python
"""Where a lone query's time goes: leader-warp cycles per phase of hnsw_beam_kernel for one query at a time (config 2
data, 200k and 10k nodes, team widths 4 / 2 / 1, ef 64 / 20), from the instrumented build of tools/build_phase_variant.py:
  python tools/build_phase_variant.py && MUNINN_CUDA_SO=sqlite_muninn_b200/csrc/libmuninn_cuda_phase.so python tools/phase_probe.py"""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from sqlite_muninn_b200.cuda import HnswIndexCuda
names = ["setup", "descent", "pop", "list", "visited", "handover", "issue", "rows+dist", "wait_team", "merge", "total"]
for n in (200000, 10000):
    c = bench.cfg_of(2); c.n = n
    g = bench.get_graph(c, 0, 1, lambda: None)
    q = bench.make_vectors(64, 128, c.seeds[1], kind=c.data)
    with HnswIndexCuda.from_graph(g) as idx:
        for W in (4, 2, 1):
            idx.set_option("kernel_gen", 8); idx.set_tuning(W, 0)
            for ef in (64, 20):
                acc = np.zeros(11)
                for rep in range(6):
                    for i in range(8):
                        ids, dist, cnt, st = idx.search(q[i:i+1], 11, ef, want_stats=True)
                        if rep: acc += dist[0].astype(np.float64)
                acc /= 40
                ms = idx.last_kernel_ms()
                print(json.dumps({"n": n, "W": W, "ef": ef, "n_expand": float(st["n_expand"][0]), "kernel_ms_last": round(ms, 4),
                                  "cycles": {k: int(v) for k, v in zip(names, acc)}}), flush=True)
