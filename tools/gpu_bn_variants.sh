#!/bin/bash
# BatchNorm loop shapes (CNB_BN_{BWD,FWD,STATS}_VARIANT; usage: [MODEL=mobilenet] gpu_bn_variants.sh <tag> BWD|FWD|STATS <variants...>): parity of each, then the ResNet-50 bench line.
set -u
TAG=${1:-v}; KIND=${2:-BWD}; shift; shift
mkdir -p gpurun_out
for V in "$@"; do
  env CNB_BN_${KIND}_VARIANT=$V timeout 300 python -m pytest tests/test_gpu_ops.py -q -k "batch_norm or bn_ or _bn" > gpurun_out/pytest_bn_${TAG}_$V.log 2>&1
  echo "variant $V parity rc=$?"; tail -1 gpurun_out/pytest_bn_${TAG}_$V.log
  env CNB_BN_${KIND}_VARIANT=$V timeout 300 python bench.py ${MODEL:+--model $MODEL} --steps 10 --warmup 4 --no-cpu-baseline --no-e2e --trace-out gpurun_out/trace_${TAG}_$V.json > gpurun_out/bench_${TAG}_$V.log 2> gpurun_out/bench_${TAG}_$V.err
  echo "variant $V bench rc=$?"; tail -1 gpurun_out/bench_${TAG}_$V.log | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('  %.0f img/s  %.2f ms/step'%(d['value'], d['ms_per_step']))
except Exception as e: print('no json', e)"
  grep "^#" gpurun_out/bench_${TAG}_$V.err | grep -E "bn_bwd|bn_fwd|im2col"
done
