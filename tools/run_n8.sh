#!/bin/bash
# N-GPU run exactly as the driver launches it (reference arm first, then ours), reduced extras to bound the box time
cd "$(dirname "$0")/.."
N=${1:-8}; EXTRAS=${2:-split10k,5}
mkdir -p gpurun_out; O=gpurun_out
nvidia-smi -L | wc -l > $O/n${N}_gpus.txt; free -g | head -2 >> $O/n${N}_gpus.txt; nproc >> $O/n${N}_gpus.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --impl reference --gpus $N --steps 5 --warmup 1 > $O/n${N}_ref.json 2> $O/n${N}_ref.err; echo "ref rc=$?"
timeout 1800 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus $N --steps 20 --warmup 5 --extras $EXTRAS > $O/n${N}_bench.json 2> $O/n${N}_bench.err; echo "bench rc=$?"
grep "^{" $O/n${N}_bench.json | cut -c1-700; grep -v absent $O/n${N}_bench.err | grep -v "^\[bench\] rank [1-9]" | tail -6 | cut -c1-300
