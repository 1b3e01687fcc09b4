#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 --extras none > $O/n1b_bench.json 2> $O/n1b_bench.err; echo "n1 rc=$?"
grep "^{" $O/n1b_bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('N1 value',round(d['value']),'e2e',round(d['e2e']['value']),'frac',d['roofline']['frac'],'verified',d['verified'],d['clocks'])"
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 2 --steps 20 --warmup 5 --extras split10k > $O/n2b_bench.json 2> $O/n2b_bench.err; echo "n2 rc=$?"
grep "^{" $O/n2b_bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('N2 value',round(d['value']),'e2e',round(d['e2e']['value']),'frac',d['roofline']['frac'],'verified',d['verified'], 'split', round(d['extra']['split_10k']['value']), round(d['extra']['split_10k']['e2e']['value']), d['extra']['split_10k']['verified'])"
grep -i "error\|Traceback" -A5 $O/n2b_bench.err | head -20
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "bench_shapes or golden" > $O/tests_g.log 2>&1; tail -2 $O/tests_g.log
