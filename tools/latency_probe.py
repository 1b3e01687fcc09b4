#!/usr/bin/env python
"""Small-batch latency of muninn_cuda_search_batch (host buffers, the call the vtab glue makes):
wall time per call and the search kernels' own event time, for nq = 1 … 1024. One JSON line per size."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import make_vectors  # noqa: E402
from sqlite_muninn_b200.builder import build_graph  # noqa: E402
from sqlite_muninn_b200.cuda import HnswIndexCuda  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
ef = int(sys.argv[2]) if len(sys.argv) > 2 else 64
base, q = make_vectors(n, 128, 11), make_vectors(4096, 128, 12)
g = build_graph(base, 0, 16, 200, batch=1024)
with HnswIndexCuda.from_graph(g) as idx:
    for nq in (1, 2, 8, 32, 148, 1024, 4096):
        reps = 300 if nq <= 148 else 50
        for i in range(20):
            idx.search(q[i:i + nq] if nq == 1 else q[:nq], 10, ef)
        kms = []
        t0 = time.perf_counter()
        for i in range(reps):
            idx.search(q[i % 1000:i % 1000 + nq] if nq == 1 else q[:nq], 10, ef)
            kms.append(idx.last_kernel_ms())
        dt = (time.perf_counter() - t0) / reps
        print(json.dumps({"n": n, "ef": ef, "nq": nq, "wall_us_per_call": round(dt * 1e6, 1),
                          "kernel_us": round(float(np.median(kms)) * 1e3, 1), "us_per_query": round(dt * 1e6 / nq, 2)}))
