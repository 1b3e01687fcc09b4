"""Stand-alone check of build_graph_device on cuda:0 (no pytest, no torch): equality with the host batch mode and timing.
usage: python tools/check_device_build.py [n] [dim] [M] [efc] [batch]      (MUNINN_BUILD_LOOP=c: the loop in C, muninn_build_run)"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sqlite_muninn_b200.builder import build_graph, build_graph_device, build_graph_device_c  # noqa: E402

n, dim, M, efc, batch = (int(a) for a in (sys.argv[1:6] + ["8000", "64", "16", "64", "512"][len(sys.argv) - 1:]))
rng = np.random.default_rng(1)
z = rng.standard_normal((n, 16)).astype(np.float32)
x = (z @ rng.standard_normal((16, dim)).astype(np.float32) + 0.05 * rng.standard_normal((n, dim))).astype(np.float32)
t0 = time.perf_counter()
host = build_graph(x, 0, M, efc, batch=batch, warm=4096, sort_lists=False)
t1 = time.perf_counter()
st = {}
loop = build_graph_device_c if os.environ.get("MUNINN_BUILD_LOOP") == "c" else build_graph_device
dev = loop(x, 0, M, efc, batch=batch, warm=4096, sort_lists=False, stats=st)
t2 = time.perf_counter()
first = host.first_list()
same_lists = float(np.mean([np.array_equal(host.neighbours(s, 0), dev.neighbours(s, 0)) for s in range(0, n, max(1, n // 2000))]))
print(json.dumps({"n": n, "dim": dim, "M": M, "efc": efc, "batch": batch, "host_batch_s": round(t1 - t0, 3),
                  "device_assisted_s": round(t2 - t1, 3), "identical": bool(dev.same_structure(host)),
                  "sampled_identical_layer0_lists": same_lists, "host_threads": os.cpu_count(), "loop": loop.__name__, **st}))
