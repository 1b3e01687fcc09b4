#!/usr/bin/env python
"""Per-kernel SASS instruction histogram of libmuninn_cuda.so (cuobjdump -sass), the evidence that the kernels are
Blackwell-native: UTCHMMA (tcgen05.mma), LDTM (tcgen05.ld), UTMALDG (TMA tensor loads), UBLKCP (TMA bulk copies),
UTCBAR (tcgen05.commit), FADD2 (packed f32x2), … Writes a text table to stdout (kept under profiles/)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "sqlite_muninn_b200", "csrc", "libmuninn_cuda.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
KEY = ("UTCHMMA", "UTCHMMA.2CTA", "LDTM", "UTMALDG", "UTMALDG.2D.2CTA", "UBLKCP", "UTCBAR", "UTCATOMSWS", "SYNCS", "FADD2", "FFMA", "FADD", "FMUL",
       "HMMA", "LDGSTS", "MATCH", "VOTE", "SHFL", "LDS", "STS", "LDG", "STG", "ATOMG", "BAR", "CCTL", "UCGABAR_ARV", "UCGABAR_WAIT")
kern, hist, total = None, {}, collections.Counter()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = re.sub(r"\(.*", "", kern)
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(2)
        hist[kern][op.split(".")[0]] += 1
        if op.startswith("UTCHMMA.2CTA"):
            hist[kern]["UTCHMMA.2CTA"] += 1
        if op.startswith("UTMALDG") and "2CTA" in op:
            hist[kern]["UTMALDG.2D.2CTA"] += 1
        total[kern] += 1
print(f"SASS summary of {os.path.relpath(so, ROOT)} (cuobjdump -sass; counts of static instructions per kernel)\n")
for k in sorted(hist, key=lambda k: -total[k]):
    if total[k] < 50:
        continue
    cols = " ".join(f"{op}={hist[k][op]}" for op in KEY if hist[k].get(op))
    print(f"{k}\n    total={total[k]}  {cols}\n")
