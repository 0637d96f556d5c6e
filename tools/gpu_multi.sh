#!/bin/bash
# multi-GPU bench only (run under gpurun --gpus N).  usage: gpu_multi.sh <tag> <N>
set -u
TAG=$1; N=$2
mkdir -p gpurun_out
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${TAG}_n$N.log 2> gpurun_out/bench_${TAG}_n$N.err
echo "bench N=$N rc=$?"; tail -1 gpurun_out/bench_${TAG}_n$N.log | cut -c1-900; grep -v "^#" gpurun_out/bench_${TAG}_n$N.err | tail -5 | cut -c1-300
