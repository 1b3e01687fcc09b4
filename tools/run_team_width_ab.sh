#!/bin/bash
# Automatic team width (4 / 2 / 1 warps per query) against each width forced, config 2 at 200k (the single-statement shape) and at 1M.
cd "$(dirname "$0")/.."
O=gpurun_out/teamw; mkdir -p $O
SETS="--set 0:0:2:0 --set 8:4:2:0 --set 8:2:2:0 --set 8:1:2:0"
timeout 600 python tools/ab_kernels.py --config 2 --n 200000 --ef 64 --nq 1,16,148,500,592,593,1250 $SETS --reps 21 --tag n200k > $O/ab_n200k.jsonl 2> $O/ab_n200k.err; echo "200k rc=$?"
timeout 600 python tools/ab_kernels.py --config 2 --ef 64 --nq 1,148,500,592,593,1250,1480,1481 $SETS --reps 21 --tag n1M > $O/ab_n1M.jsonl 2> $O/ab_n1M.err; echo "1M rc=$?"
cat $O/ab_n200k.jsonl $O/ab_n1M.jsonl > $O/ab_team_width4.jsonl
python - <<'PY'
import json
for l in open('gpurun_out/teamw/ab_team_width4.jsonl'):
    d=json.loads(l); print(d['tag'], d['set'], 'W', d['W_run'], 'nq', d['nq'], 'ms', d['kernel_ms_med'], d['kernel_ms_min'], d['ids_sha'])
PY
