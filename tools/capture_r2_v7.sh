#!/bin/bash
# ncu --set full captures of a search kernel where it is weakest: ef=256, 384-byte rows, 3 KB rows, a lone query.
# Each capture follows a plain run of the same command (B200_PROFILING.md). The reports are summarised ON THE BOX
# (gpurun_out/ is capped at 64 MiB): metrics + per-source-line table, the .ncu-rep itself is dropped.
#   usage: capture_r2_v7.sh <gen> <kernel regex> [so suffix]
cd "$(dirname "$0")/.."
GEN=${1:-7}; RX=${2:-hnsw_search_kernel}; SFX=${3:-}
export MUNINN_CUDA_SO=$PWD/sqlite_muninn_b200/csrc/libmuninn_cuda$SFX.so
O=gpurun_out/r2_g$GEN$SFX; mkdir -p $O
cap() { # name, skip, args…
  name=$1; skip=$2; shift 2
  python tools/ab_kernels.py "$@" > $O/$name.plain.jsonl 2> $O/$name.plain.err &&
  ncu --set full --clock-control none --import-source on -k regex:$RX -s $skip -c 1 -f -o /tmp/$name python tools/ab_kernels.py "$@" > $O/$name.ncu.log 2>&1
  echo "$name rc=$?"
  python profiles/summarize_ncu.py /tmp/$name.ncu-rep 45 > $O/$name.summary.txt 2>&1
  ncu -i /tmp/$name.ncu-rep --page raw --csv > $O/$name.raw.csv 2>/dev/null
  ncu -i /tmp/$name.ncu-rep --page source --csv --print-source cuda 2>/dev/null | gzip -9 > $O/$name.source.csv.gz
  rm -f /tmp/$name.ncu-rep
}
cap c2_ef256 9 --config 2 --ef 256 --nq 10000 --set $GEN:1:2:0 --reps 2
cap c2_ef64_nq1 9 --config 2 --ef 64 --nq 1 --set $GEN:1:2:0 --reps 2
cap c4p_ef64 9 --config 4 --n 2000000 --ef 64 --nq 10000 --set $GEN:1:2:0 --reps 2
cap c3_ef128 9 --config 3 --ef 128 --nq 10000 --set $GEN:1:2:0 --reps 2
du -sh gpurun_out; ls -la $O
