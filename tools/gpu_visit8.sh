#!/bin/bash
# usage: bash tools/gpu_visit8.sh <tag> — the model-level half of the GPU suite (the conv / dense / fused-path
# halves ran green in r02g), bench with the BatchNorm-sum epilogue on and off, the depthwise-statistics cases
set -u
TAG=${1:-r02h}
O=gpurun_out
mkdir -p $O
timeout 560 python -m pytest tests/test_gpu_fused_paths.py tests/test_gpu_layers_in_network.py tests/test_gpu_models.py \
  tests/test_gpu_ops.py tests/test_gpu_properties.py tests/test_optim.py -m gpu -q -rfP --timeout=200 \
  -k "not depthwise_emits and not dgrad_add_epilogue and not rgb_stem and not batched_pack" > $O/pytest_$TAG.log 2>&1
echo "pytest rc=$?"; grep -E "^FAILED|^ERROR" $O/pytest_$TAG.log | cut -c1-400 | head -40; tail -1 $O/pytest_$TAG.log
grep -E "^PARITY|^LAYERS|^EVAL|^DENSE|BatchNorm backward calls|on-vs-off" $O/pytest_$TAG.log > $O/parity_$TAG.txt
grep -E "BatchNorm backward calls" $O/parity_$TAG.txt | cut -c1-230 | head -14
line() { python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('  %.0f img/s  %.2f ms/step  e2e %.0f (%.2f ms)  roof %s %.3f'%(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['roofline']['kernel'], d['roofline']['frac']))
except Exception as e: print('  no json', e)"; }
timeout 150 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --trace-out $O/trace_$TAG.json > $O/bench_$TAG.log 2> $O/bench_$TAG.err
echo "bench rc=$?"; tail -1 $O/bench_$TAG.log | line; grep -v "^#" $O/bench_$TAG.err | grep -v -i warn | tail -3 | cut -c1-300
grep "^#" $O/bench_$TAG.err > $O/trace_families_$TAG.txt; head -8 $O/trace_families_$TAG.txt
python tools/trace_table.py $O/trace_$TAG.json 90 > $O/trace_layers_$TAG.txt 2>&1
CNB_BN_BWD_SUMS=0 timeout 150 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_${TAG}_nosums.log 2> $O/bench_${TAG}_nosums.err
echo "bench CNB_BN_BWD_SUMS=0 rc=$?"; tail -1 $O/bench_${TAG}_nosums.log | line; grep "^#" $O/bench_${TAG}_nosums.err | head -3
timeout 100 python -m pytest tests/test_gpu_fused_paths.py -m gpu -q -x --timeout=60 -k "depthwise_emits" > $O/pytest_dw_$TAG.log 2>&1
echo "depthwise statistics cases rc=$?"; tail -2 $O/pytest_dw_$TAG.log | cut -c1-200
