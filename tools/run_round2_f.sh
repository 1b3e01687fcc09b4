#!/bin/bash
# list-ahead prefetch for 64-entry lists (A/B on config 3 / 4 proxies), parity of the touched tests, exact-mode 1M graph headline
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out; D=sqlite_muninn_b200/csrc
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_vtab.py tests/test_gpu_patch.py -x -q -m gpu > $O/tests_f.log 2>&1; echo "tests rc=$?" >> $O/tests_f.log; tail -3 $O/tests_f.log
OUT=$O/ab_f.jsonl; : > $OUT
for v in "" _nola; do
  MUNINN_CUDA_SO=$PWD/$D/libmuninn_cuda$v.so python tools/ab_kernels.py --config 3 --n 200000 --ef 128,384 --nq 10000 --set 7:1:2:0 --reps 8 --tag c3p >> $OUT 2>> $O/ab_f.err
  MUNINN_CUDA_SO=$PWD/$D/libmuninn_cuda$v.so python tools/ab_kernels.py --config 4 --n 2000000 --ef 64,128 --nq 10000,1250 --set 0:0:2:0 --set 7:1:2:0 --tag c4p >> $OUT 2>> $O/ab_f.err
done
cut -c1-230 $OUT
timeout 2400 python bench.py --build-batch 0 --steps 20 --warmup 5 --extras split10k > $O/bench_exactgraph.json 2> $O/bench_exactgraph.err; echo "bench rc=$?"
grep -v absent $O/bench_exactgraph.err | tail -5 | cut -c1-400
