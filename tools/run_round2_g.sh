#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out; D=sqlite_muninn_b200/csrc
: > $O/ab_g.jsonl
for v in "" _occ32 _occ28; do
  MUNINN_CUDA_SO=$PWD/$D/libmuninn_cuda$v.so python tools/ab_kernels.py --config 2 --ef 64,128 --nq 80000,10000 --set 7:1:2:0 --reps 15 --tag occ >> $O/ab_g.jsonl 2>> $O/ab_g.err
done
python -c "
import json
for l in open('$O/ab_g.jsonl'):
    r=json.loads(l); print(r['so'],'ef',r['ef'],'nq',r['nq'],'med',r['kernel_ms_med'],'min',r['kernel_ms_min'],'GB/s',r['algo_GBps'],r['ids_sha'][:6])
"
