#!/bin/bash
# usage: bash tools/gpu_visit5.sh <tag> — gate, full tests, bench A/B (filter-resident, depthwise statistics), e2e check
set -u
TAG=${1:-r02e}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_conv.py tests/test_gpu_fused_paths.py tests/test_optim.py -m gpu -q -x --timeout=120 > $O/pytest_gate_$TAG.log 2>&1
rc=$?; echo "gate rc=$rc"; tail -3 $O/pytest_gate_$TAG.log
if [ $rc -ne 0 ]; then grep -E "^FAILED|^ERROR|Error|assert" $O/pytest_gate_$TAG.log | head -30; exit 1; fi
timeout 1500 python -m pytest tests -m gpu -q -rfP --timeout=600 > $O/pytest_$TAG.log 2>&1
echo "pytest rc=$?"; grep -E "^FAILED|^ERROR" $O/pytest_$TAG.log | cut -c1-400 | head -40; tail -1 $O/pytest_$TAG.log
grep -E "^PARITY|^LAYERS|^EVAL|Appendix A|on-vs-off" $O/pytest_$TAG.log > $O/parity_$TAG.txt
line() { python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('  %.0f img/s  %.2f ms/step  e2e %.0f (%.2f ms)  roof %s %.3f'%(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['roofline']['kernel'], d['roofline']['frac']))
except Exception as e: print('  no json', e)"; }
timeout 300 python bench.py --steps 10 --warmup 4 --no-cpu-baseline --trace-out $O/trace_$TAG.json > $O/bench_$TAG.log 2> $O/bench_$TAG.err
echo "bench rc=$?"; tail -1 $O/bench_$TAG.log | line
grep "^#" $O/bench_$TAG.err > $O/trace_families_$TAG.txt; head -8 $O/trace_families_$TAG.txt
python tools/trace_table.py $O/trace_$TAG.json 90 > $O/trace_layers_$TAG.txt 2>&1
CNB_BRES=0 timeout 300 python bench.py --steps 10 --warmup 4 --no-cpu-baseline > $O/bench_${TAG}_nobres.log 2>/dev/null
echo "bench CNB_BRES=0 rc=$?"; tail -1 $O/bench_${TAG}_nobres.log | line
timeout 300 python tools/layer_bench.py --reps 5 --only conv --convs 0,1,2,3,5,7 > $O/layers_$TAG.txt 2>&1; echo "layer_bench rc=$?"; cat $O/layers_$TAG.txt
CNB_BRES=0 timeout 300 python tools/layer_bench.py --reps 5 --only conv --convs 0,1,2,3 > $O/layers_${TAG}_nobres.txt 2>&1; echo "layer_bench nobres rc=$?"; cat $O/layers_${TAG}_nobres.txt
for m in mobilenet shufflenet_g3 resnext50; do
  timeout 300 python bench.py --model $m --steps 10 --warmup 4 --no-cpu-baseline > $O/zoo_${TAG}_$m.log 2> $O/zoo_${TAG}_$m.err
  echo "== $m rc=$?"; tail -1 $O/zoo_${TAG}_$m.log | line; grep "^#" $O/zoo_${TAG}_$m.err | head -6
done
CNB_DW_STATS=0 timeout 300 python bench.py --model mobilenet --steps 10 --warmup 4 --no-cpu-baseline > $O/zoo_${TAG}_mobilenet_nodwstats.log 2> $O/zoo_${TAG}_mobilenet_nodwstats.err
echo "== mobilenet CNB_DW_STATS=0 rc=$?"; tail -1 $O/zoo_${TAG}_mobilenet_nodwstats.log | line; grep "^#" $O/zoo_${TAG}_mobilenet_nodwstats.err | head -6
CNB_BRES=0 timeout 300 python bench.py --model resnext50 --steps 20 --warmup 4 --no-cpu-baseline > $O/zoo_${TAG}_resnext50_nobres.log 2>/dev/null
echo "== resnext50 CNB_BRES=0, 20 steps rc=$?"; tail -1 $O/zoo_${TAG}_resnext50_nobres.log | line
timeout 400 python tools/infer_bench.py > $O/infer_$TAG.jsonl 2> $O/infer_$TAG.err; echo "infer rc=$?"; cut -c1-150 $O/infer_$TAG.jsonl
du -sm $O
