#!/bin/bash
# round-2 A/B sweep #1: ring geometry × speculation × team width (run under gpurun)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
D=sqlite_muninn_b200/csrc
python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/parity_v8.log 2>&1; echo "parity rc=$?" >> gpurun_out/parity_v8.log
tail -5 gpurun_out/parity_v8.log
OUT=gpurun_out/ab_a.jsonl; : > $OUT
for v in "" _u32 _u32n4 _u128 _nospec _spec1; do
  MUNINN_CUDA_SO=$PWD/$D/libmuninn_cuda$v.so python tools/ab_kernels.py --config 2 --ef 64,256 --nq 10000 --set 8:1:2:0 --tag big >> $OUT 2>> gpurun_out/ab_a.err
done
python tools/ab_kernels.py --config 2 --ef 64,128,256 --nq 10000 --set 7:1:2:0 --set 8:1:1:0 --tag big >> $OUT 2>> gpurun_out/ab_a.err
for v in "" _nospec _u32; do
  MUNINN_CUDA_SO=$PWD/$D/libmuninn_cuda$v.so python tools/ab_kernels.py --config 2 --ef 64 --nq 1250,148,1 --set 8:1:2:0 --set 8:2:2:0 --set 8:4:2:0 --set 8:0:2:0 --set 7:1:2:0 --reps 30 --tag small >> $OUT 2>> gpurun_out/ab_a.err
done
for v in "" _u32 _u32n4 _u128 _nospec; do
  MUNINN_CUDA_SO=$PWD/$D/libmuninn_cuda$v.so python tools/ab_kernels.py --config 4 --n 2000000 --ef 64 --nq 10000 --set 8:1:2:0 --set 7:1:2:0 --tag c4proxy >> $OUT 2>> gpurun_out/ab_a.err
done
for v in "" _u32 _u32n4 _u128 _nospec; do
  MUNINN_CUDA_SO=$PWD/$D/libmuninn_cuda$v.so python tools/ab_kernels.py --config 3 --ef 128,384 --nq 10000 --set 8:1:2:0 --set 7:1:2:0 --reps 8 --tag c3 >> $OUT 2>> gpurun_out/ab_a.err
done
wc -l $OUT
