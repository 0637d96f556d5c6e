#!/bin/bash
# usage: bash tools/gpu_visit3.sh <tag>  — tests, bench, layer bench, sanitizer on the g4 inference crash,
# inference bench, ncu (CSV exports only), per-config bench lines
set -u
TAG=${1:-r02c}
O=gpurun_out
mkdir -p $O
# gate: the rewritten issue loops / scatter epilogue must pass the conv cases before anything else runs
timeout 420 python -m pytest tests/test_gpu_conv.py tests/test_optim.py -m gpu -q -x --timeout=120 > $O/pytest_gate_$TAG.log 2>&1
rc=$?; echo "gate rc=$rc"; tail -3 $O/pytest_gate_$TAG.log
if [ $rc -ne 0 ]; then grep -E "^FAILED|^ERROR|Error|assert" $O/pytest_gate_$TAG.log | head -30; exit 1; fi
timeout 1500 python -m pytest tests -m gpu -q -rfP --timeout=600 > $O/pytest_$TAG.log 2>&1
echo "pytest rc=$?"; grep -E "^FAILED|^ERROR" $O/pytest_$TAG.log | cut -c1-400 | head -40; tail -1 $O/pytest_$TAG.log
grep -E "^PARITY|^LAYERS|^EVAL|Appendix A|on-vs-off" $O/pytest_$TAG.log > $O/parity_$TAG.txt
timeout 300 python bench.py --steps 10 --warmup 4 --trace-out $O/trace_$TAG.json > $O/bench_$TAG.log 2> $O/bench_$TAG.err
echo "bench rc=$?"; tail -1 $O/bench_$TAG.log | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('  %.0f img/s  %.2f ms/step  e2e %.0f  roof %s %.3f  launches %s'%(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['kernel'], d['roofline']['frac'], d.get('gpu_launches')))
except Exception as e: print('  no json', e)"
grep "^#" $O/bench_$TAG.err > $O/trace_families_$TAG.txt; head -14 $O/trace_families_$TAG.txt
python tools/trace_table.py $O/trace_$TAG.json 90 > $O/trace_layers_$TAG.txt 2>&1
CNB_FUSED_ADAM=0 timeout 300 python bench.py --steps 10 --warmup 4 --no-cpu-baseline > $O/bench_${TAG}_torchadam.log 2>/dev/null
echo "bench torch-adam rc=$?"; tail -1 $O/bench_${TAG}_torchadam.log | cut -c1-120
timeout 300 python bench.py --steps 10 --warmup 4 --no-cpu-baseline --input f32 > $O/bench_${TAG}_f32in.log 2>/dev/null
echo "bench f32-input rc=$?"; tail -1 $O/bench_${TAG}_f32in.log | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('  f32 input: %.0f img/s e2e %.0f'%(d['value'], d['e2e']['value']))"
timeout 300 python tools/layer_bench.py --reps 5 --only conv > $O/layers_$TAG.txt 2>&1; echo "layer_bench rc=$?"; head -60 $O/layers_$TAG.txt
# the inference crash of the r02b visit (ShuffleNet-v1 g4, 32x32, batch 256): name the kernel
timeout 600 compute-sanitizer --tool memcheck --print-limit 5 python tools/infer_bench.py --only ShuffleNet --batches 1 --eager > $O/sanitizer_$TAG.log 2>&1
echo "sanitizer rc=$?"; grep -E "Invalid|Error|at .*cnb|by thread|=========     in" $O/sanitizer_$TAG.log | head -20
timeout 400 python tools/infer_bench.py > $O/infer_$TAG.jsonl 2> $O/infer_$TAG.err; echo "infer rc=$?"; cut -c1-160 $O/infer_$TAG.jsonl
cap() {  # name, layer_bench args, kernel regex, count
  local LB="python tools/layer_bench.py --reps 1 $2"
  timeout 200 $LB > $O/lb_plain_$1.txt 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k "regex:$3" -c $4 -f -o /tmp/full_$1 $LB > $O/ncu_$1.log 2>&1
  echo "ncu $1 rc=$?"
  ncu -i /tmp/full_$1.ncu-rep --page raw --csv > $O/ncu_raw_${TAG}_$1.csv 2>/dev/null
  ncu -i /tmp/full_$1.ncu-rep --page source --csv > $O/ncu_src_${TAG}_$1.csv 2>/dev/null
  gzip -f $O/ncu_src_${TAG}_$1.csv
}
cap c64k3 "--only conv --convs 1" "conv_fprop|conv_wgrad" 6
cap c256k3 "--only conv --convs 9" "conv_fprop|conv_wgrad" 6
cap c128s2 "--only conv --convs 5" "conv_fprop|conv_wgrad|dgrad_zero" 10
# BASELINE.json configs as stated (row d2): one line each
zoo() {  # name, extra args
  timeout 400 python bench.py --model $1 $2 --steps 6 --warmup 4 --cpu-steps 2 > $O/zoo_${TAG}_$1.log 2> $O/zoo_${TAG}_$1.err
  echo "== $1 $2 rc=$?"; tail -1 $O/zoo_${TAG}_$1.log | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('  %.0f img/s %.2f ms/step e2e %.0f | %s %.3f | cpu %s'%(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['kernel'], d['roofline']['frac'], d.get('cpu_baseline',{}).get('value')))
except Exception as e: print('  no json', e)"
  grep "^#" $O/zoo_${TAG}_$1.err | head -5
}
zoo vgg11 "--hw 32 --dtype f32 --batch 64"
zoo vgg16 "--hw 32 --dtype f32 --batch 64"
zoo mobilenet ""
zoo shufflenet_g3 ""
zoo resnext50 ""
zoo densenet121 ""
zoo se_resnet50 ""
zoo sknet50 ""
du -sm $O
