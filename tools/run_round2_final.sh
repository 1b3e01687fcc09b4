#!/bin/bash
# final-state validation: the whole -m gpu suite, smoke(), then the default bench command exactly as the driver runs it
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > $O/tests_final.log 2>&1; echo "tests rc=$?" >> $O/tests_final.log; tail -3 $O/tests_final.log
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > $O/smoke_final.log 2>&1; tail -1 $O/smoke_final.log
python tools/latency_probe.py 200000 64 > $O/latency_final.jsonl 2>/dev/null; cat $O/latency_final.jsonl | head -3
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $O/final_ref.json 2> $O/final_ref.err; echo "ref rc=$?"
for b in 1 0; do MUNINN_BALANCE_ROUNDS=$b python tools/ab_kernels.py --config 2 --ef 64,128 --nq 80000,10000,5000,2500 --set 0:0:2:0 --reps 20 --tag balance$b >> $O/ab_balance.jsonl 2>> $O/ab_balance.err; done
python -c "
import json
for l in open('$O/ab_balance.jsonl'):
    r=json.loads(l); print(r['tag'],'ef',r['ef'],'nq',r['nq'],'med',r['kernel_ms_med'],'min',r['kernel_ms_min'],'GB/s',r['algo_GBps'])
"
( time timeout 3000 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/final_bench.json 2> $O/final_bench.err ) 2> $O/final_bench.time; echo "bench rc=$?"; cat $O/final_bench.time
grep -v absent $O/final_bench.err | grep -i "error\|fail\|Traceback" | head
