#!/bin/bash
# Round-2 first GPU visit: full GPU test suite at HEAD (fused paths default-on, floor-bound parity
# bars, in-network per-layer checks), bench A/B of the three switches, ncu evidence for the conv
# kernel families and BatchNorm backward.   usage: bash tools/gpu_visit1.sh <tag>
set -u
TAG=${1:-r02a}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi_$TAG.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -rA --timeout=600 > $O/pytest_$TAG.log 2>&1
echo "pytest rc=$?"; grep -E "passed|failed|error" $O/pytest_$TAG.log | tail -3
grep -E "^PARITY|^LAYERS|Appendix A|on-vs-off" $O/pytest_$TAG.log > $O/parity_$TAG.txt; wc -l $O/parity_$TAG.txt
for cfg in "" "CNB_FUSED_SKIP_GRAD=0" "CNB_BATCHED_PACK=0" "CNB_STEM_ROWS=0"; do
  name=$(echo "$cfg" | tr ' =' '__'); name=${name:-default}
  env $cfg timeout 300 python bench.py --steps 10 --warmup 4 --no-cpu-baseline --trace-out $O/trace_${TAG}_$name.json \
    > $O/bench_${TAG}_$name.log 2> $O/bench_${TAG}_$name.err
  echo "bench [$cfg] rc=$?"; tail -1 $O/bench_${TAG}_$name.log | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('  %.0f img/s  %.2f ms/step  e2e %.0f  roof %s %.3f'%(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['kernel'], d['roofline']['frac']))
except Exception as e: print('  no json', e)"
done
python tools/trace_table.py $O/trace_${TAG}_default.json 70 > $O/trace_layers_$TAG.txt 2>&1
grep "^#" $O/bench_${TAG}_default.err > $O/trace_families_$TAG.txt
# per-layer micro-benchmark, then the same command under ncu --set full (conv families + BatchNorm)
LB="python tools/layer_bench.py --reps 1 --convs 1,5,9,13 --bns 2,5"
timeout 300 python tools/layer_bench.py --reps 5 > $O/layers_$TAG.txt 2>&1; echo "layer_bench rc=$?"
timeout 300 $LB > $O/layers_plain_$TAG.txt 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on \
  -k "regex:conv_fprop|conv_wgrad|dgrad_interleave|sum_splits|bn_bwd|bn_apply" -c 80 -f -o $O/full_$TAG $LB > $O/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
# launch list + DRAM traffic of one eager step (same command first without ncu)
BC="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-graph"
timeout 300 $BC > $O/bench_plain_$TAG.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 6000 --csv \
  --log-file $O/launches_$TAG.csv $BC > $O/ncu_list_$TAG.log 2>&1
echo "ncu list rc=$?"
gzip -f $O/launches_$TAG.csv
ls -la $O | tail -25
