#!/usr/bin/env python
"""Fixture for the builder's exact mode at size ("diff a 50k build against the reference before trusting
it at 1M"): builds N x 128 (bench.py's config-2 distribution, M=16, ef_construction=200) with the UNMODIFIED reference's
hnsw_insert (oracle/_ref, src/hnsw_algo.c:520-666) and stores a digest of the whole graph — entry point, max level,
every node's level and every neighbour list in insertion order and in shadow-table (sorted) order — plus a few lists
in clear. tests/test_host.py rebuilds the same input with csrc/hnsw_build.c (batch=0) and compares.

  python tests/golden/make_build_fixture.py [N=50000]      (about 7 minutes of reference time at 50k)
"""
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def graph_digest(g) -> str:
    h = hashlib.sha256()
    h.update(np.int64([g.n, g.entry_slot, g.max_level]).tobytes())
    h.update(np.ascontiguousarray(g.levels, np.int32).tobytes())
    h.update(np.ascontiguousarray(g.list_offsets, np.int64).tobytes())
    h.update(np.ascontiguousarray(g.targets, np.int32).tobytes())
    return h.hexdigest()


def main():
    from bench import make_vectors
    from oracle.bindings import Reference
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
    dim, M, efc, seed = 128, 16, 200, 11
    base = make_vectors(n, dim, seed)
    t0 = time.time()
    ri = Reference().build(base, metric=0, M=M, efc=efc)
    secs = time.time() - t0
    g = ri.export()
    d_ins = graph_digest(g)
    ri.sort_lists()
    gs = ri.export()
    d_sorted = graph_digest(gs)
    sample = {str(s): [int(t) for t in g.neighbours(s, 0)] for s in (0, 1, n // 2, n - 1)}
    out = {"n": n, "dim": dim, "M": M, "ef_construction": efc, "data": "bench.make_vectors(n, 128, 11, kind='latent16')",
           "reference_build_seconds": round(secs, 1), "entry_slot": int(g.entry_slot), "max_level": int(g.max_level),
           "n_edges": int(g.targets.shape[0]), "level_histogram": np.bincount(g.levels).tolist(),
           "digest_insertion_order": d_ins, "digest_sorted_lists": d_sorted, "level0_lists_insertion_order": sample}
    path = os.path.join(ROOT, "tests", "golden", f"build_{n // 1000}kx{dim}_M{M}_efc{efc}.json")
    json.dump(out, open(path, "w"), indent=1)
    print(json.dumps(out)[:400], "->", path)


if __name__ == "__main__":
    main()
