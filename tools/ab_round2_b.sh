#!/bin/bash
# round-2 sweep #2: exact k-NN query-stationary filter (tests + timing), ring geometry without speculation, merge fast path
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
D=sqlite_muninn_b200/csrc
python -m pytest tests/test_gpu_exact_tc.py tests/test_multi.py tests/test_gpu_parity.py -x -q -m gpu > $O/tests_b.log 2>&1; echo "tests rc=$?" >> $O/tests_b.log
tail -4 $O/tests_b.log
: > $O/exact_b.jsonl
python tools/bench_exact.py --n 1000000 --nq 10000 --k 10 --modes 3:0,1:256 >> $O/exact_b.jsonl 2>> $O/exact_b.err
MUNINN_EXACT_QS=0 python tools/bench_exact.py --n 1000000 --nq 10000 --k 10 --modes 2:256 >> $O/exact_b.jsonl 2>> $O/exact_b.err
for sb in 8 16 64; do MUNINN_EXACT_SB_MB=$sb python tools/bench_exact.py --n 1000000 --nq 10000 --k 10 --modes 3:0 >> $O/exact_b.jsonl 2>> $O/exact_b.err; done
python tools/bench_exact.py --n 6250000 --nq 10000 --k 100 --modes 3:0,1:256 --reps 2 >> $O/exact_b.jsonl 2>> $O/exact_b.err
python tools/bench_exact.py --n 1000000 --dim 96 --metric 2 --nq 10000 --k 10 --modes 3:0,1:256 --reps 2 >> $O/exact_b.jsonl 2>> $O/exact_b.err
OUT=$O/ab_b.jsonl; : > $OUT
for v in "" _g1 _g2 _g3 _g4; do
  MUNINN_CUDA_SO=$PWD/$D/libmuninn_cuda$v.so python tools/ab_kernels.py --config 3 --n 200000 --ef 128,384 --nq 10000 --set 8:1:2:0 --set 7:1:2:0 --reps 8 --tag c3p >> $OUT 2>> $O/ab_b.err
  MUNINN_CUDA_SO=$PWD/$D/libmuninn_cuda$v.so python tools/ab_kernels.py --config 4 --n 1000000 --ef 64 --nq 10000 --set 8:1:2:0 --set 7:1:2:0 --tag c4p >> $OUT 2>> $O/ab_b.err
  MUNINN_CUDA_SO=$PWD/$D/libmuninn_cuda$v.so python tools/ab_kernels.py --config 2 --ef 64,256 --nq 10000 --set 8:1:2:0 --set 7:1:2:0 --tag c2 >> $OUT 2>> $O/ab_b.err
done
python tools/ab_kernels.py --config 2 --ef 64 --nq 1250,600,148,1 --set 8:1:2:0 --set 8:2:2:0 --set 8:4:2:0 --set 7:1:2:0 --reps 30 --tag small >> $OUT 2>> $O/ab_b.err
wc -l $OUT $O/exact_b.jsonl
