#!/bin/bash
# CTA-pair exact filter (tests under a timeout, then timing), device-assisted build at size, heavy-deletion tests
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_exact_tc.py -x -q -m gpu -k "query_stationary" > $O/tests_e_pair.log 2>&1; echo "pair tests rc=$?" >> $O/tests_e_pair.log
tail -4 $O/tests_e_pair.log
if grep -q "pair tests rc=0" $O/tests_e_pair.log; then
  : > $O/exact_e.jsonl
  MUNINN_EXACT_PAIR=1 timeout 300 python tools/bench_exact.py --n 1000000 --nq 10000 --k 10 --modes 3:0 >> $O/exact_e.jsonl 2>> $O/exact_e.err
  MUNINN_EXACT_PAIR=0 timeout 300 python tools/bench_exact.py --n 1000000 --nq 10000 --k 10 --modes 3:0 >> $O/exact_e.jsonl 2>> $O/exact_e.err
  MUNINN_EXACT_PAIR=1 timeout 300 python tools/bench_exact.py --n 6250000 --nq 10000 --k 100 --modes 3:0 --reps 2 >> $O/exact_e.jsonl 2>> $O/exact_e.err
  MUNINN_EXACT_PAIR=1 timeout 300 python tools/bench_exact.py --n 1000000 --dim 96 --metric 2 --nq 10000 --k 10 --modes 3:0 --reps 2 >> $O/exact_e.jsonl 2>> $O/exact_e.err
  cut -c1-260 $O/exact_e.jsonl
fi
timeout 600 python -m pytest tests/test_gpu_patch.py tests/test_gpu_vtab.py -x -q -m gpu > $O/tests_e.log 2>&1; echo "tests rc=$?" >> $O/tests_e.log; tail -3 $O/tests_e.log
timeout 900 python tools/check_device_build.py 300000 128 16 200 1024 > $O/devbuild_300k.json 2> $O/devbuild_300k.err; cat $O/devbuild_300k.json
MUNINN_BUILD_LOOP=c timeout 900 python tools/check_device_build.py 300000 128 16 200 4096 > $O/devbuild_300k_c4096.json 2>> $O/devbuild_300k.err; cat $O/devbuild_300k_c4096.json
