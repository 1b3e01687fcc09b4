#!/bin/bash
# ncu --set full captures of the round-1 kernel (generation 7) where it is weakest: ef=256, 384-byte rows, 3 KB rows,
# and a lone query. Each capture follows a plain run of the same command (B200_PROFILING.md).
cd "$(dirname "$0")/.."
O=gpurun_out/r2_v7; mkdir -p $O
cap() { # name, kernel regex, skip, args…
  name=$1; rx=$2; skip=$3; shift 3
  python tools/ab_kernels.py "$@" > $O/$name.plain.jsonl 2> $O/$name.plain.err &&
  ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -f -o $O/$name python tools/ab_kernels.py "$@" > $O/$name.ncu.log 2>&1
  echo "$name rc=$?"
}
cap c2_ef256 hnsw_search_kernel 9 --config 2 --ef 256 --nq 10000 --set 7:1:2:0 --reps 2
cap c2_ef64_nq1 hnsw_search_kernel 9 --config 2 --ef 64 --nq 1 --set 7:1:2:0 --reps 2
cap c4p_ef64 hnsw_search_kernel 9 --config 4 --n 2000000 --ef 64 --nq 10000 --set 7:1:2:0 --reps 2
cap c3_ef128 hnsw_search_kernel 9 --config 3 --ef 128 --nq 10000 --set 7:1:2:0 --reps 2
ls -la $O
