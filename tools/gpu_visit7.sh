#!/bin/bash
# usage: bash tools/gpu_visit7.sh <tag> — per-feature gates (a failing feature is switched off for the rest of
# the visit), bench + A/B, full tests, per-config lines, launch list under ncu
set -u
TAG=${1:-r02g}
O=gpurun_out
mkdir -p $O
gate() {  # name, -k expression, files
  timeout 420 python -m pytest $3 -m gpu -q -x --timeout=200 -k "$2" -rP > $O/gate_${1}_$TAG.log 2>&1
  local rc=$?; echo "gate $1 rc=$rc: $(tail -1 $O/gate_${1}_$TAG.log | cut -c1-120)"
  grep -E "sums=|DENSE block|on-vs-off" $O/gate_${1}_$TAG.log | cut -c1-220 | head -16
  if [ $rc -ne 0 ]; then grep -E "^FAILED|^ERROR|Error|^E " $O/gate_${1}_$TAG.log | head -12 | cut -c1-260; fi
  return $rc
}
gate sums "test_bn_backward_sums_and_lazy_skip" tests/test_gpu_fused_paths.py || { export CNB_BN_BWD_SUMS=0 CNB_LAZY_SKIP_GRAD=0; echo "SWITCHED OFF: BN sums, lazy skip"; }
gate dense "test_dense_block or test_prefix or test_stats_from" tests/test_gpu_dense.py || { export CNB_DENSE_BUFFER=0; echo "SWITCHED OFF: dense buffer"; }
gate epi "test_dgrad_add_epilogue or test_depthwise_emits" tests/test_gpu_fused_paths.py
line() { python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('  %.0f img/s  %.2f ms/step  e2e %.0f (%.2f ms)  roof %s %.3f'%(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['roofline']['kernel'], d['roofline']['frac']))
except Exception as e: print('  no json', e)"; }
timeout 400 python bench.py --steps 20 --warmup 5 --trace-out $O/trace_$TAG.json > $O/bench_$TAG.log 2> $O/bench_$TAG.err
echo "bench rc=$?"; tail -1 $O/bench_$TAG.log | line; grep -v "^#" $O/bench_$TAG.err | tail -3 | cut -c1-300
grep "^#" $O/bench_$TAG.err > $O/trace_families_$TAG.txt; head -12 $O/trace_families_$TAG.txt
python tools/trace_table.py $O/trace_$TAG.json 90 > $O/trace_layers_$TAG.txt 2>&1
CNB_BN_BWD_SUMS=0 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_${TAG}_nosums.log 2> $O/bench_${TAG}_nosums.err
echo "bench CNB_BN_BWD_SUMS=0 rc=$?"; tail -1 $O/bench_${TAG}_nosums.log | line; grep "^#" $O/bench_${TAG}_nosums.err | head -4
CNB_BN_BWD_SUMS=0 CNB_LAZY_SKIP_GRAD=0 CNB_BRES=0 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_${TAG}_alloff.log 2>/dev/null
echo "bench sums, lazy skip, filter-resident all off rc=$?"; tail -1 $O/bench_${TAG}_alloff.log | line
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_${TAG}_again.log 2>/dev/null
echo "bench (default, again) rc=$?"; tail -1 $O/bench_${TAG}_again.log | line
timeout 1100 python -m pytest tests -m gpu -q -rfP --timeout=300 > $O/pytest_$TAG.log 2>&1
echo "pytest rc=$?"; grep -E "^FAILED|^ERROR" $O/pytest_$TAG.log | cut -c1-400 | head -40; tail -1 $O/pytest_$TAG.log
grep -E "^PARITY|^LAYERS|^EVAL|^DENSE|sums=|Appendix A|on-vs-off" $O/pytest_$TAG.log > $O/parity_$TAG.txt
zoo() {  # name, env assignment, suffix
  env $2 timeout 300 python bench.py --model $1 --steps 10 --warmup 4 --no-cpu-baseline > $O/zoo_${TAG}_$1$3.log 2> $O/zoo_${TAG}_$1$3.err
  echo "== $1 [$2] rc=$?"; tail -1 $O/zoo_${TAG}_$1$3.log | line; grep "^#" $O/zoo_${TAG}_$1$3.err | head -6; grep -v "^#" $O/zoo_${TAG}_$1$3.err | grep -v Warning | tail -2 | cut -c1-300
}
zoo densenet121 A=1 ""
zoo densenet121 CNB_DENSE_BUFFER=0 _nobuffer
zoo shufflenet_g3 A=1 ""
zoo mobilenet A=1 ""
BC="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-graph"
timeout 200 $BC > $O/bench_plain_$TAG.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv \
  --log-file $O/launches_$TAG.csv $BC > $O/ncu_list_$TAG.log 2>&1
echo "ncu list rc=$?"
gzip -f $O/launches_$TAG.csv
du -sm $O
