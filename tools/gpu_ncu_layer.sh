#!/bin/bash
# ncu --set full on one kernel family of the layer micro-benchmark.
# usage: gpu_ncu_layer.sh <tag> <kernel regex> <layer_bench args...>
set -u
TAG=$1; KRE=$2; shift 2
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$KRE" -c 6 \
  -o gpurun_out/ncu_$TAG -f python tools/layer_bench.py "$@" --reps 1 > gpurun_out/ncu_$TAG.log 2>&1
echo "ncu rc=$?"; tail -5 gpurun_out/ncu_$TAG.log
