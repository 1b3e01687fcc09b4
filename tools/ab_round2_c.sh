#!/bin/bash
# round-2 sweep #3: full GPU suite on the by-shape kernel selection, exact filter with the 8-warp epilogue, bench dry run
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > $O/tests_c.log 2>&1; echo "tests rc=$?" >> $O/tests_c.log
tail -4 $O/tests_c.log
: > $O/exact_c.jsonl
python tools/bench_exact.py --n 1000000 --nq 10000 --k 10 --modes 3:0 >> $O/exact_c.jsonl 2>> $O/exact_c.err
python tools/bench_exact.py --n 6250000 --nq 10000 --k 100 --modes 3:0 --reps 2 >> $O/exact_c.jsonl 2>> $O/exact_c.err
python tools/bench_exact.py --n 1000000 --dim 96 --metric 2 --nq 10000 --k 10 --modes 3:0 --reps 2 >> $O/exact_c.jsonl 2>> $O/exact_c.err
python tools/bench_exact.py --n 6250000 --nq 100 --k 100 --modes 3:0 --reps 3 >> $O/exact_c.jsonl 2>> $O/exact_c.err
cat $O/exact_c.jsonl | cut -c1-330
python tools/ab_kernels.py --config 2 --ef 64,256 --nq 10000,1250,600,148,1 --set 0:0:2:0 --set 7:1:2:0 --reps 20 --tag auto > $O/ab_c.jsonl 2> $O/ab_c.err
timeout 1500 python bench.py --steps 20 --warmup 5 --extras split10k,3,5 > $O/bench_c.json 2> $O/bench_c.err; echo "bench rc=$?"
tail -c 1500 $O/bench_c.json; tail -5 $O/bench_c.err
