#!/bin/bash
# usage: gpu_layers.sh <tag> <pytest expr|none> [layer_bench args]
set -u
TAG=$1; TESTS=$2; shift 2
mkdir -p gpurun_out
if [ "$TESTS" != "none" ]; then
  timeout 1500 python -m pytest tests -m gpu -q -k "$TESTS" > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"
  grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu_$TAG.log | cut -c1-200 | tail
fi
timeout 600 python tools/layer_bench.py "$@" --out gpurun_out/layers_$TAG.json > gpurun_out/layers_$TAG.log 2>&1; echo "layer_bench rc=$?"; cat gpurun_out/layers_$TAG.log | tail -60
