#!/bin/bash
# usage: bash tools/gpu_visit6.sh <tag> — quick gate, full tests, bench A/B of this session's switches, per-config lines
set -u
TAG=${1:-r02f}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_fused_paths.py tests/test_gpu_dense.py -m gpu -q -x --timeout=150 \
  -k "test_conv_tensor_core or test_dgrad_add_epilogue or test_depthwise_emits or test_prefix or test_stats_from or test_dense_block" \
  > $O/pytest_gate_$TAG.log 2>&1
rc=$?; echo "gate rc=$rc"; tail -3 $O/pytest_gate_$TAG.log | cut -c1-300
if [ $rc -ne 0 ]; then grep -E "^FAILED|^ERROR|Error|assert|^E " $O/pytest_gate_$TAG.log | head -40 | cut -c1-300; fi
timeout 1800 python -m pytest tests -m gpu -q -rfP --timeout=600 > $O/pytest_$TAG.log 2>&1
echo "pytest rc=$?"; grep -E "^FAILED|^ERROR" $O/pytest_$TAG.log | cut -c1-400 | head -40; tail -1 $O/pytest_$TAG.log
grep -E "^PARITY|^LAYERS|^EVAL|^DENSE|Appendix A|on-vs-off" $O/pytest_$TAG.log > $O/parity_$TAG.txt
line() { python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('  %.0f img/s  %.2f ms/step  e2e %.0f (%.2f ms)  roof %s %.3f'%(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['roofline']['kernel'], d['roofline']['frac']))
except Exception as e: print('  no json', e)"; }
timeout 300 python bench.py --steps 10 --warmup 4 --no-cpu-baseline --trace-out $O/trace_$TAG.json > $O/bench_$TAG.log 2> $O/bench_$TAG.err
echo "bench rc=$?"; tail -1 $O/bench_$TAG.log | line; tail -3 $O/bench_$TAG.err | cut -c1-300
grep "^#" $O/bench_$TAG.err > $O/trace_families_$TAG.txt; head -8 $O/trace_families_$TAG.txt
python tools/trace_table.py $O/trace_$TAG.json 90 > $O/trace_layers_$TAG.txt 2>&1
for sw in CNB_BRES CNB_LAZY_SKIP_GRAD; do
  env $sw=0 timeout 300 python bench.py --steps 10 --warmup 4 --no-cpu-baseline > $O/bench_${TAG}_$sw.log 2>/dev/null
  echo "bench $sw=0 rc=$?"; tail -1 $O/bench_${TAG}_$sw.log | line
done
timeout 300 python tools/layer_bench.py --reps 5 --only conv --convs 0,1,2,3,5,7 > $O/layers_$TAG.txt 2>&1; echo "layer_bench rc=$?"; cat $O/layers_$TAG.txt
CNB_BRES=0 timeout 300 python tools/layer_bench.py --reps 5 --only conv --convs 0,1,2,3 > $O/layers_${TAG}_nobres.txt 2>&1; echo "layer_bench CNB_BRES=0 rc=$?"; cat $O/layers_${TAG}_nobres.txt
zoo() {  # name, env assignment, suffix, steps
  env $2 timeout 400 python bench.py --model $1 --steps $4 --warmup 4 --no-cpu-baseline > $O/zoo_${TAG}_$1$3.log 2> $O/zoo_${TAG}_$1$3.err
  echo "== $1 [$2] rc=$?"; tail -1 $O/zoo_${TAG}_$1$3.log | line; grep "^#" $O/zoo_${TAG}_$1$3.err | head -6; grep -v "^#" $O/zoo_${TAG}_$1$3.err | tail -2 | cut -c1-300
}
zoo mobilenet A=1 "" 10
zoo mobilenet CNB_DW_STATS=0 _nodwstats 10
zoo densenet121 A=1 "" 10
zoo densenet121 CNB_DENSE_BUFFER=0 _nobuffer 10
zoo shufflenet_g3 A=1 "" 10
zoo resnext50 A=1 "" 20
zoo se_resnet50 A=1 "" 20
zoo sknet50 A=1 "" 10
timeout 400 python tools/infer_bench.py > $O/infer_$TAG.jsonl 2> $O/infer_$TAG.err; echo "infer rc=$?"; cut -c1-150 $O/infer_$TAG.jsonl; tail -3 $O/infer_$TAG.err | cut -c1-200
du -sm $O
