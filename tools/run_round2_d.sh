#!/bin/bash
# full GPU suite, config 1 through both extensions, bench with the config-4 block (10M x 96 build time), auto team width check
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > $O/tests_d.log 2>&1; echo "tests rc=$?" >> $O/tests_d.log
tail -4 $O/tests_d.log
timeout 900 python tools/bench_config1.py > $O/config1.json 2> $O/config1.err; echo "config1 rc=$?"; cut -c1-1500 $O/config1.json
python tools/ab_kernels.py --config 2 --ef 64 --nq 1250,600,148,1 --set 0:0:2:0 --set 8:2:2:0 --set 7:1:2:0 --reps 30 --tag autoW > $O/ab_d.jsonl 2> $O/ab_d.err
timeout 2400 python bench.py --steps 20 --warmup 5 --extras 4 --no-cpu-baseline > $O/bench_d.json 2> $O/bench_d.err; echo "bench rc=$?"
grep -v absent $O/bench_d.err | tail -8 | cut -c1-300
