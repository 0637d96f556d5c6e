#!/bin/bash
# GPU visit: run a pytest -k selection without -x, keep the log.  usage: gpu_tests.sh <tag> <expr>
set -u
TAG=$1; EXPR=$2
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -k "$EXPR" > gpurun_out/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_$TAG.log
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu_$TAG.log | tail -40
