#!/bin/bash
# quick GPU visit: selected tests + bench with trace.  usage: gpu_quick.sh <tag> <pytest -k expr|none> [extra bench args]
set -u
TAG=$1; TESTS=$2; shift 2
mkdir -p gpurun_out
if [ "$TESTS" != "none" ]; then
  timeout 1500 python -m pytest tests -m gpu -q -x -k "$TESTS" > gpurun_out/pytest_gpu_$TAG.log 2>&1
  echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_$TAG.log; tail -12 gpurun_out/pytest_gpu_$TAG.log
fi
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --trace-out gpurun_out/trace_$TAG.json "$@" > gpurun_out/bench_$TAG.log 2> gpurun_out/bench_$TAG.err
echo "bench rc=$?"; tail -1 gpurun_out/bench_$TAG.log | cut -c1-400; grep "^#" gpurun_out/bench_$TAG.err | head -12; tail -5 gpurun_out/bench_$TAG.err | grep -v "^#"
