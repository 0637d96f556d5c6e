#!/bin/bash
set -u
TAG=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "test_gpu_models and f32 or im2col" > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu_$TAG.log | cut -c1-200 | tail
timeout 600 python tools/layer_bench.py --only conv --reps 3 --out gpurun_out/layers_$TAG.json > gpurun_out/layers_$TAG.log 2>&1; echo "layer_bench rc=$?"; cat gpurun_out/layers_$TAG.log | grep wgrad
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:conv_wgrad_tc" -c 4 -o gpurun_out/wgrad_$TAG -f python tools/layer_bench.py --only conv --convs 9,1 --reps 1 > gpurun_out/ncu_wgrad_$TAG.log 2>&1; echo "ncu rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --trace-out gpurun_out/trace_$TAG.json > gpurun_out/bench_$TAG.log 2> gpurun_out/bench_$TAG.err; grep "^#" gpurun_out/bench_$TAG.err | head -8; cut -c1-200 gpurun_out/bench_$TAG.log
