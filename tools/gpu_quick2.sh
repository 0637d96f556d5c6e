#!/bin/bash
# short GPU visit: op-level parity, the train-step parity of the models that use the touched
# kernels, two bench lines.  usage: gpu_quick2.sh <tag>
set -u
TAG=${1:-q}
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_ops.py -q > gpurun_out/pytest_ops_$TAG.log 2>&1
echo "ops rc=$?"; tail -3 gpurun_out/pytest_ops_$TAG.log
timeout 500 python -m pytest tests/test_gpu_models.py -q -k "ResNet or ShuffleNet or SqueezeNet or VGG or graph" > gpurun_out/pytest_models_$TAG.log 2>&1
echo "models rc=$?"; tail -3 gpurun_out/pytest_models_$TAG.log
timeout 300 python bench.py --steps 10 --warmup 4 --no-cpu-baseline --trace-out gpurun_out/trace_$TAG.json > gpurun_out/bench_$TAG.log 2> gpurun_out/bench_$TAG.err
echo "bench rc=$?"; tail -1 gpurun_out/bench_$TAG.log | cut -c1-260; grep "^#" gpurun_out/bench_$TAG.err | head -12
bash tools/gpu_zoo_bench.sh $TAG shufflenet_g3
