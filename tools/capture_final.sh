#!/bin/bash
# Round-end evidence for the shipped kernels: (1) launch list of the default bench command (headline only), (2) ncu --set
# full of the headline search launches (80k-query job and the 10k batch) → DRAM traffic keyed by the kernel-source hash,
# (3) the query-stationary exact filter. Summaries are made on the box; the reports themselves are dropped (64 MiB cap).
cd "$(dirname "$0")/.."
O=gpurun_out/final; mkdir -p $O
python -c "import bench; print(bench.kernel_source_sha())" > $O/kernel_src_sha.txt
CMD="python bench.py --steps 2 --warmup 3 --extras none --no-cpu-baseline"
$CMD > $O/bench_plain.json 2> $O/bench_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench_headline.csv $CMD > $O/ncu_list.log 2>&1
echo "launch list rc=$?"
cap() { # name, kernel regex, skip, args…
  name=$1; rx=$2; skip=$3; shift 3
  python tools/ab_kernels.py "$@" > $O/$name.plain.jsonl 2> $O/$name.plain.err &&
  ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -f -o /tmp/$name python tools/ab_kernels.py "$@" > $O/$name.ncu.log 2>&1
  echo "$name rc=$?"
  python profiles/summarize_ncu.py /tmp/$name.ncu-rep 45 > $O/$name.summary.txt 2>&1
  ncu -i /tmp/$name.ncu-rep --page raw --csv > $O/$name.raw.csv 2>/dev/null
  rm -f /tmp/$name.ncu-rep
}
cap c2_ef64_nq10000 hnsw_search_kernel 9 --config 2 --ef 64 --nq 10000 --set 0:0:2:0 --reps 2
cap c2_ef64_nq80000 hnsw_search_kernel 9 --config 2 --ef 64 --nq 80000 --set 0:0:2:0 --reps 2
cap c2_ef256_gen8 hnsw_beam_kernel 9 --config 2 --ef 256 --nq 10000 --set 0:0:2:0 --reps 2
cap c2_ef64_nq1250_w2 hnsw_beam_kernel 9 --config 2 --ef 64 --nq 1250 --set 0:0:2:0 --reps 2
python tools/bench_exact.py --n 1000000 --nq 10000 --k 10 --modes 3:0 --reps 1 > $O/exact_qs.plain.jsonl 2> $O/exact_qs.plain.err &&
ncu --set full --clock-control none --import-source on -k regex:exact_tc_filter_qs -s 13 -c 1 -f -o /tmp/exact_qs python tools/bench_exact.py --n 1000000 --nq 10000 --k 10 --modes 3:0 --reps 1 > $O/exact_qs.ncu.log 2>&1
python profiles/summarize_ncu.py /tmp/exact_qs.ncu-rep 40 > $O/exact_qs.summary.txt 2>&1
ncu -i /tmp/exact_qs.ncu-rep --page raw --csv > $O/exact_qs.raw.csv 2>/dev/null; rm -f /tmp/exact_qs.ncu-rep
du -sh gpurun_out; ls $O
