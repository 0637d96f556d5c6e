#!/bin/bash
# multi-GPU bench (run under gpurun --gpus N).  usage: gpu_multi.sh <tag> <N>
set -u
TAG=$1; N=$2
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${TAG}_n$N.log 2> gpurun_out/bench_${TAG}_n$N.err
echo "bench N=$N rc=$?"; tail -1 gpurun_out/bench_${TAG}_n$N.log | cut -c1-700; tail -5 gpurun_out/bench_${TAG}_n$N.err | grep -v "^#" | cut -c1-300
timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 > gpurun_out/bench_${TAG}_n1.log 2> gpurun_out/bench_${TAG}_n1.err
echo "bench N=1 rc=$?"; tail -1 gpurun_out/bench_${TAG}_n1.log | cut -c1-400
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_${TAG}_ref.log 2>&1; tail -1 gpurun_out/bench_${TAG}_ref.log | cut -c1-400
