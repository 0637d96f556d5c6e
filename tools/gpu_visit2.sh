#!/bin/bash
# usage: bash tools/gpu_visit2.sh <tag>   — tests, BatchNorm backward sweep, bench, inference bench, ncu (CSV exports only)
set -u
TAG=${1:-r02b}
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -rfP --timeout=600 > $O/pytest_$TAG.log 2>&1
echo "pytest rc=$?"; grep -E "^FAILED|^ERROR" $O/pytest_$TAG.log | cut -c1-400 | head -40; tail -1 $O/pytest_$TAG.log
grep -E "^PARITY|^LAYERS|^EVAL|Appendix A|on-vs-off" $O/pytest_$TAG.log > $O/parity_$TAG.txt
wc -l $O/parity_$TAG.txt
for mb in 0 64 128; do
  CNB_BN_BWD_FUSED_MB=$mb timeout 200 python tools/layer_bench.py --only bn --reps 5 > $O/bn_fused_${mb}_$TAG.txt 2>&1
  echo "bn sweep $mb rc=$?"; grep "bn bwd" $O/bn_fused_${mb}_$TAG.txt | grep -E "C 256 H 14|C 512 H  7|C 128 H 28|C2048" | head -8
done
for mb in 0 64 128; do
  CNB_BN_BWD_FUSED_MB=$mb timeout 300 python bench.py --steps 10 --warmup 4 --no-cpu-baseline --trace-out $O/trace_${TAG}_mb$mb.json \
    > $O/bench_${TAG}_mb$mb.log 2> $O/bench_${TAG}_mb$mb.err
  echo "bench [fused mb $mb] rc=$?"; tail -1 $O/bench_${TAG}_mb$mb.log | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('  %.0f img/s  %.2f ms/step  e2e %.0f  roof %s %.3f'%(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['kernel'], d['roofline']['frac']))
except Exception as e: print('  no json', e)"
done
python tools/trace_table.py $O/trace_${TAG}_mb64.json 80 > $O/trace_layers_$TAG.txt 2>&1
grep "^#" $O/bench_${TAG}_mb64.err > $O/trace_families_$TAG.txt; head -12 $O/trace_families_$TAG.txt
timeout 400 python tools/infer_bench.py > $O/infer_$TAG.jsonl 2> $O/infer_$TAG.err; echo "infer rc=$?"; cut -c1-200 $O/infer_$TAG.jsonl
timeout 300 python tools/layer_bench.py --reps 5 --only conv > $O/layers_$TAG.txt 2>&1; echo "layer_bench rc=$?"
# ncu --set full on the conv kernel families: one layer per capture, CSV pages only travel back
cap() {  # name, layer_bench args, kernel regex, count
  local LB="python tools/layer_bench.py --reps 1 $2"
  timeout 200 $LB > $O/lb_plain_$1.txt 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k "regex:$3" -c $4 -f -o /tmp/full_$1 $LB > $O/ncu_$1.log 2>&1
  echo "ncu $1 rc=$?"
  ncu -i /tmp/full_$1.ncu-rep --page raw --csv > $O/ncu_raw_${TAG}_$1.csv 2>/dev/null
  ncu -i /tmp/full_$1.ncu-rep --page source --csv > $O/ncu_src_${TAG}_$1.csv 2>/dev/null
  gzip -f $O/ncu_src_${TAG}_$1.csv
}
cap c64k3 "--only conv --convs 1" "conv_fprop|conv_wgrad|sum_splits" 8
cap c256k3 "--only conv --convs 9" "conv_fprop|conv_wgrad|sum_splits" 8
cap c128s2 "--only conv --convs 5" "conv_fprop|dgrad_interleave|conv_wgrad" 14
cap c512k1 "--only conv --convs 13" "conv_fprop|conv_wgrad|sum_splits" 8
cap bn "--only bn --bns 2,5" "bn_bwd|bn_apply" 24
# launch list + DRAM traffic of one eager step (same command first without ncu)
BC="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-graph"
timeout 300 $BC > $O/bench_plain_$TAG.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 6000 --csv \
  --log-file $O/launches_$TAG.csv $BC > $O/ncu_list_$TAG.log 2>&1
echo "ncu list rc=$?"
gzip -f $O/launches_$TAG.csv
du -sm $O; ls -la $O | sort -k5 -n -r | head -8
