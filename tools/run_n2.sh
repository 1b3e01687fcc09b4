#!/bin/bash
# 2-GPU validation: the NCCL tests and the bench at N=2 exactly as the driver launches it
cd "$(dirname "$0")/.."
mkdir -p gpurun_out; O=gpurun_out
nvidia-smi -L > $O/n2_gpus.txt
timeout 600 python -m pytest tests/test_multi.py -x -q -m gpu > $O/n2_tests.log 2>&1; echo "tests rc=$?" >> $O/n2_tests.log
tail -5 $O/n2_tests.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 --extras ${EXTRAS:-split10k,3,5} > $O/n2_bench.json 2> $O/n2_bench.err; echo "bench rc=$?"
tail -c 600 $O/n2_bench.json; grep -v "absent" $O/n2_bench.err | tail -12
