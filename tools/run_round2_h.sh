#!/bin/bash
# arrival-watermark e2e path: parity, headline e2e, then the traffic recapture for the changed kernel sources
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_vtab.py -x -q -m gpu > $O/tests_h.log 2>&1; echo "tests rc=$?" >> $O/tests_h.log; tail -3 $O/tests_h.log
for i in 1 2; do
timeout 900 python bench.py --steps 20 --warmup 5 --extras none --no-cpu-baseline > $O/h_bench$i.json 2> $O/h_bench$i.err; echo "bench rc=$?"
grep "^{" $O/h_bench$i.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('N1 value',round(d['value']),'e2e',round(d['e2e']['value']),'ratio',round(d['e2e']['value']/d['value'],4),'frac',d['roofline']['frac'],'verified',d['verified'],d['clocks'])"
done
bash tools/capture_final.sh > $O/capture_h.log 2>&1; tail -3 $O/capture_h.log
