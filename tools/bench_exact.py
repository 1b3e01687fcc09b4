#!/usr/bin/env python
"""Exact k-NN timing on one GPU: every evaluation mode on the same inputs, results cross-checked.

  python tools/bench_exact.py [--n 1000000] [--dim 128] [--nq 10000] [--k 10] [--metric 0] [--reps 3]

Prints one JSON line per (mode, tile_rows): wall time of the whole call (CUDA events on the handle's
stream), the tensor-core filter's own time, survivors per query, and the tensor roofline:
algorithmic flops 2·nq·n·dim (3xTF32 issues 3 MMAs per product, so the issued rate is
3× that) against the measured dense bf16 peak / 2 (TF32 runs at half the bf16 rate)."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--nq", type=int, default=10_000)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--metric", type=int, default=0)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--modes", default="1:256,1:128,2:256,0:0")
    args = ap.parse_args()
    import torch
    import __graft_entry__ as ge
    ge.build()
    from bench import make_vectors
    from sqlite_muninn_b200.cuda import HnswIndexCuda
    from sqlite_muninn_b200.graph import FlatGraph
    base = make_vectors(args.n, args.dim, 41)
    q = make_vectors(args.nq, args.dim, 42)
    if args.metric == 2:
        base /= np.linalg.norm(base, axis=1, keepdims=True)
    n = args.n
    g = FlatGraph(args.dim, args.metric, 8, 10, np.arange(n, dtype=np.int64), np.zeros(n, np.int32), np.zeros(n, np.uint8),
                  base, np.zeros(n + 1, np.int64), np.empty(0, np.int32), 0, 0)
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"] / 2.0
        src = "measured bf16 burst / 2"
    except Exception:
        peak, src = 1100.0, "fallback (nominal dense TF32)"
    idx = HnswIndexCuda.from_graph(g)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    idx.set_stream(stream.cuda_stream)
    dq = torch.from_numpy(q).cuda()
    outs = {}
    for spec in args.modes.split(","):
        mode, tile = (int(x) for x in spec.split(":"))
        idx.set_exact_mode(mode, tile)
        i = torch.empty((args.nq, args.k), dtype=torch.int64, device="cuda")
        d = torch.empty((args.nq, args.k), dtype=torch.float32, device="cuda")
        c = torch.empty(args.nq, dtype=torch.int32, device="cuda")
        ms, fms = [], []
        for _ in range(args.reps + 1):
            idx.exact_knn_device(dq.data_ptr(), args.nq, args.k, 0, i.data_ptr(), d.data_ptr(), c.data_ptr())
            torch.cuda.synchronize()
            ms.append(idx.last_kernel_ms())
            fms.append(idx.exact_stats()["filter_ms"])
        st = idx.exact_stats()
        outs[spec] = (i.cpu().numpy(), d.cpu().numpy())
        flops = 2.0 * args.nq * n * args.dim
        best, fbest = min(ms[1:]), min(fms[1:])
        line = {"workload": f"exact k-NN {n}x{args.dim} metric={args.metric} nq={args.nq} k={args.k}", "mode": mode,
                "tile_rows": tile, "ms": round(best, 3), "filter_ms": round(fbest, 3), "chunks": st["chunks"],
                "survivors_per_query": round(st["candidates"] / args.nq, 1), "overflow_queries": st["overflow_queries"],
                "algorithmic_TFLOPs": round(flops / best / 1e9, 1)}
        if st["mode"] != 0:
            issued = flops * (3 if st["mode"] == 1 else 1)
            line["mode_run"] = st["mode"]
            line["roofline"] = {"bound": "tensor", "achieved": round(issued / fbest / 1e9, 1), "peak": peak, "unit": "TFLOP/s",
                                "frac": round(issued / fbest / 1e9 / peak, 4), "peak_source": src,
                                "note": "issued TF32 MMA flops of the filter kernel / its CUDA-event time"}
        print(json.dumps(line), flush=True)
    ref = outs[list(outs)[-1]]
    for spec, (i, d) in outs.items():
        assert np.array_equal(i, ref[0]) and d.tobytes() == ref[1].tobytes(), f"mode {spec} differs from {list(outs)[-1]}"
    print(json.dumps({"all_modes_identical": True, "modes": list(outs)}))
    idx.close()


if __name__ == "__main__":
    main()
