#!/bin/bash
# One-warp kernels on batches that fit ONE wave of resident teams: 13-row kernel (gen 7, 24 teams/SM) against the 16-row ring kernel (gen 8).
cd "$(dirname "$0")/.."
O=gpurun_out/wave; mkdir -p $O
S="--set 7:1:2:0 --set 8:1:2:0 --reps 21"
timeout 500 python tools/ab_kernels.py --config 2 --n 200000 --ef 32,128,256 --nq 1500,2500,2960 $S --tag c2_200k > $O/c2.jsonl 2> $O/c2.err; echo "c2 rc=$?"
timeout 700 python tools/ab_kernels.py --config 4 --n 200000 --efc 48 --ef 64 --nq 1500,2500,2960,3552 $S --tag c4_200k > $O/c4.jsonl 2> $O/c4.err; echo "c4 rc=$?"
timeout 700 python tools/ab_kernels.py --config 3 --n 100000 --efc 48 --ef 128 --nq 1000,1500,2000,3000 $S --tag c3_100k > $O/c3.jsonl 2> $O/c3.err; echo "c3 rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/wave/c*.jsonl')):
    for l in open(f):
        d=json.loads(l); print(d['tag'], d['set'], 'ef', d['ef'], 'nq', d['nq'], 'ms', d['kernel_ms_med'], d['kernel_ms_min'], d['ids_sha'])
PY
