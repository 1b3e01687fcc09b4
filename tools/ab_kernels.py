#!/usr/bin/env python
"""A/B timing of search-kernel shapes on a bench workload, one process per library build:

  MUNINN_CUDA_SO=…/libmuninn_cuda_x.so python tools/ab_kernels.py --config 2 --ef 64,256 --nq 10000,1250,1 \
        --set 8:1:2:0 --set 8:2:2:0 --set 7:1:2:0 [--n 200000]

--set gen:W:vis_factor:teams_per_sm[:vis_in_smem]. For every (set, ef, nq): median / min kernel time over --reps launches
(CUDA events around the launch, muninn_cuda_last_kernel_ms), algorithmic GB/s from the kernel's own counters,
and a checksum of the returned ids so that every shape can be seen to return the same answer. JSON lines.
"""
import argparse
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from sqlite_muninn_b200.cuda import CUDA_SO, HnswIndexCuda  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=2)
    ap.add_argument("--n", type=int, default=None)
    ap.add_argument("--efc", type=int, default=None)
    ap.add_argument("--ef", default="64")
    ap.add_argument("--nq", default="10000")
    ap.add_argument("--set", action="append", default=[])
    ap.add_argument("--reps", type=int, default=15)
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    c = bench.cfg_of(a.config)
    if a.n:
        c.n = a.n
    if a.efc:
        c.efc = a.efc
    cfg = vars(c)
    g = bench.get_graph(c, 0, 1, lambda: None)
    nq_all = max(int(x) for x in a.nq.split(","))
    queries = bench.make_vectors(nq_all, cfg["dim"], cfg["seeds"][1], kind=cfg["data"])
    k = cfg["k"]
    sets = a.set or ["8:0:2:0"]
    with HnswIndexCuda.from_graph(g) as idx:
        for s in sets:
            gen, W, vf, tps, vs = (s.split(":") + ["1"])[:5]
            idx.set_option("vis_in_smem", int(vs))
            idx.set_option("kernel_gen", int(gen))
            idx.set_option("vis_factor", float(vf))
            idx.set_tuning(1 if int(gen) == 7 else int(W), int(tps))
            for ef in (int(x) for x in a.ef.split(",")):
                for nq in (int(x) for x in a.nq.split(",")):
                    q = queries[:nq]
                    ids, dist, cnt, st = idx.search(q, k, ef, want_stats=True)
                    for _ in range(3):
                        idx.search(q, k, ef)
                    ms = []
                    for _ in range(a.reps):
                        idx.search(q, k, ef)
                        ms.append(idx.last_kernel_ms())
                    ab = bench.algorithmic_bytes(st, cfg["dim"], cfg["M"], k)
                    med = float(np.median(ms))
                    print(json.dumps({"so": os.path.basename(CUDA_SO), "tag": a.tag, "config": a.config, "n": cfg["n"], "set": s,
                                      "W_run": idx.last_team_width(), "gen_run": idx.last_kernel_gen(), "ef": ef, "nq": nq, "kernel_ms_med": round(med, 4),
                                      "kernel_ms_min": round(float(np.min(ms)), 4), "algo_GBps": round(ab / med / 1e6, 1),
                                      "us_per_expansion_team": None,
                                      "n_dist_mean": float(st["n_dist"].mean()), "n_expand_mean": float(st["n_expand"].mean()),
                                      "ids_sha": hashlib.sha1(ids.tobytes() + dist.tobytes()).hexdigest()[:12]}), flush=True)


if __name__ == "__main__":
    main()
