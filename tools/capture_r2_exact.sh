#!/bin/bash
# ncu --set full of the exact k-NN filter at 1M x 128, nq = 10 000, k = 10 (the largest row chunk of a call):
# the streaming one-product filter (before) and the query-stationary one (after). Summaries made on the box.
cd "$(dirname "$0")/.."
O=gpurun_out/r2_exact; mkdir -p $O
cap() { # name, env, skip
  name=$1; envs=$2; skip=$3
  env $envs python tools/bench_exact.py --n 1000000 --nq 10000 --k 10 --modes 3:0 --reps 1 > $O/$name.plain.jsonl 2> $O/$name.plain.err &&
  env $envs ncu --set full --clock-control none --import-source on -k regex:exact_tc_filter -s $skip -c 1 -f -o /tmp/$name python tools/bench_exact.py --n 1000000 --nq 10000 --k 10 --modes 3:0 --reps 1 > $O/$name.ncu.log 2>&1
  echo "$name rc=$?"
  python profiles/summarize_ncu.py /tmp/$name.ncu-rep 40 > $O/$name.summary.txt 2>&1
  ncu -i /tmp/$name.ncu-rep --page raw --csv > $O/$name.raw.csv 2>/dev/null
  ncu -i /tmp/$name.ncu-rep --page source --csv --print-source cuda 2>/dev/null | gzip -9 > $O/$name.source.csv.gz
  rm -f /tmp/$name.ncu-rep
}
cap stream1x "MUNINN_EXACT_QS=0" 13
cap qs1x "MUNINN_EXACT_QS=1" 13
grep -h -E "gpu__time_duration.sum|sm__pipe_tensor|sm__inst_executed_pipe_tensor|lts__throughput|l1tex__data_bank" $O/*.raw.csv | head -0
du -sh gpurun_out
