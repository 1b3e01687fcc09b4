#!/usr/bin/env python
"""profiles/search_traffic.json ← `ncu --page raw --csv` dumps of search-kernel captures.

  python tools/ncu_traffic.py --config 2 --ef 64 --nq 10000 --raw gpurun_out/…/name.raw.csv --sha <kernel_src_sha> [--source profiles/…]

Each record: (config, ef, nq) → dram__bytes_read.sum + dram__bytes_write.sum of ONE level-0 launch, the kernel name and
duration ncu saw, and the hash of the kernel sources the capture was taken with (bench.kernel_source_sha(), computed
on the box inside the capture script). bench.py reports `roofline.traffic` only when that hash equals the hash of the
sources it runs, so a stale figure can never be quoted for a changed kernel."""
import argparse
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, required=True)
    ap.add_argument("--ef", type=int, required=True)
    ap.add_argument("--nq", type=int, required=True)
    ap.add_argument("--raw", required=True)
    ap.add_argument("--sha", required=True)
    ap.add_argument("--source", default="")
    a = ap.parse_args()
    rows = list(csv.reader(open(a.raw)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    col = {h: i for i, h in enumerate(hdr)}

    def get(name):
        i = col[name]
        return float(vals[i]) * UNIT.get(units[i], 1.0)
    rec = {"config": a.config, "ef": a.ef, "nq": a.nq, "kernel_src_sha": a.sha,
           "dram_bytes": get("dram__bytes_read.sum") + get("dram__bytes_write.sum"),
           "dram_bytes_read": get("dram__bytes_read.sum"), "dram_bytes_write": get("dram__bytes_write.sum"),
           "kernel": vals[col["Kernel Name"]], "ncu_duration_ms": float(vals[col["gpu__time_duration.sum"]]) * {"ms": 1.0, "us": 1e-3, "s": 1e3}.get(units[col["gpu__time_duration.sum"]], 1.0),
           "source": a.source or os.path.relpath(a.raw, ROOT)}
    path = os.path.join(ROOT, "profiles", "search_traffic.json")
    recs = json.load(open(path)) if os.path.exists(path) else []
    recs = [r for r in recs if (r["config"], r["ef"], r["nq"], r["kernel_src_sha"]) != (a.config, a.ef, a.nq, a.sha)] + [rec]
    json.dump(recs, open(path, "w"), indent=1)
    print(json.dumps(rec))


if __name__ == "__main__":
    sys.exit(main())
