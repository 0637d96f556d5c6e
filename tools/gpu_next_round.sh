#!/bin/bash
# First GPU visit of the next round: validate what was written after the previous round's GPU
# minutes ran out, then measure it.  usage: bash tools/gpu_next_round.sh <tag>
#   1. full-size property tests (tests/test_gpu_properties.py)
#   2. cnb_conv2d_dgrad_add kernel check, batched filter pack check, row-im2col stem check
#   3. train-step parity + step-graph tests with CNB_FUSED_SKIP_GRAD=1 CNB_BATCHED_PACK=1 CNB_STEM_ROWS=1
#   4. ResNet-50 bench: default, each flag, both flags
set -u
TAG=${1:-n}
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_properties.py -q > gpurun_out/pytest_props_$TAG.log 2>&1
echo "properties rc=$?"; tail -3 gpurun_out/pytest_props_$TAG.log
timeout 200 python tools/check_dgrad_add.py > gpurun_out/check_dgrad_add_$TAG.log 2>&1
echo "dgrad_add rc=$?"; tail -8 gpurun_out/check_dgrad_add_$TAG.log
CNB_BATCHED_PACK=1 timeout 200 python tools/check_batched_pack.py > gpurun_out/check_pack_$TAG.log 2>&1
echo "batched pack rc=$?"; tail -6 gpurun_out/check_pack_$TAG.log
timeout 200 python tools/check_stem_rows.py > gpurun_out/check_stem_$TAG.log 2>&1
echo "stem rows rc=$?"; tail -8 gpurun_out/check_stem_$TAG.log
CNB_FUSED_SKIP_GRAD=1 CNB_BATCHED_PACK=1 CNB_STEM_ROWS=1 timeout 600 python -m pytest tests/test_gpu_models.py -q \
  > gpurun_out/pytest_models_flags_$TAG.log 2>&1
echo "models with flags rc=$?"; tail -3 gpurun_out/pytest_models_flags_$TAG.log
for cfg in "" "CNB_FUSED_SKIP_GRAD=1" "CNB_BATCHED_PACK=1" "CNB_STEM_ROWS=1" "CNB_FUSED_SKIP_GRAD=1 CNB_BATCHED_PACK=1 CNB_STEM_ROWS=1"; do
  name=$(echo "$cfg" | tr ' =' '__'); name=${name:-default}
  env $cfg timeout 300 python bench.py --steps 10 --warmup 4 --no-cpu-baseline --no-e2e \
    > gpurun_out/bench_${TAG}_$name.log 2> gpurun_out/bench_${TAG}_$name.err
  echo "bench [$cfg] rc=$?"; tail -1 gpurun_out/bench_${TAG}_$name.log | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('  %.0f img/s  %.2f ms/step'%(d['value'], d['ms_per_step']))
except Exception as e: print('  no json', e)"
done
# 5. statistics-kernel loop shapes on the model that uses them most (BatchNorm after depthwise convs)
MODEL=mobilenet bash tools/gpu_bn_variants.sh ${TAG}s STATS 0 1 2
