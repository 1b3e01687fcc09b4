#!/usr/bin/env python
"""Cost of the global-memory visited test-and-set inside a lone query: cycles of the bucket gather and of match.any per
loop iteration, from the instrumented build with PHASE_VISITED=1 (tools/build_phase_variant.py; force the global table with
MUNINN_VIS_SMEM=0). Measured: 76 iterations for 65 calls, ≈ 800 cycles per gather, ≈ 450 per match.any."""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from sqlite_muninn_b200.cuda import HnswIndexCuda
c = bench.cfg_of(2); c.n = 200000
g = bench.get_graph(c, 0, 1, lambda: None)
q = bench.make_vectors(8, 128, c.seeds[1], kind=c.data)
with HnswIndexCuda.from_graph(g) as idx:
    for W in (4, 1):
        idx.set_option("kernel_gen", 8); idx.set_tuning(W, 0)
        prev = None
        for rep in range(4):
            ids, dist, cnt, st = idx.search(q[:1], 15, 64, want_stats=True)
            cur = dist[0].astype(np.float64)
            if prev is not None:
                d = cur[11:15] - prev[11:15]
                print(json.dumps({"W": W, "n_expand": float(st["n_expand"][0]), "total": cur[10], "visited_phase": cur[4], "list_phase": cur[3],
                                  "vis_load_cycles": d[0], "vis_match_cycles": d[1], "vis_iterations": d[3]}))
            prev = cur
