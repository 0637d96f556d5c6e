#!/bin/bash
# usage: bash tools/gpu_visit10.sh <tag> — the whole GPU suite at HEAD, the default bench line, the launch list
set -u
TAG=${1:-r02j}
O=gpurun_out
mkdir -p $O
timeout 420 python -m pytest tests -m gpu -q -rfP --timeout=150 > $O/pytest_$TAG.log 2>&1
echo "pytest rc=$?"; grep -E "^FAILED|^ERROR" $O/pytest_$TAG.log | cut -c1-300 | head -30; tail -1 $O/pytest_$TAG.log
grep -E "^PARITY|^LAYERS|^EVAL|^DENSE|BatchNorm backward calls|on-vs-off" $O/pytest_$TAG.log > $O/parity_$TAG.txt
timeout 200 python bench.py --trace-out $O/trace_$TAG.json > $O/bench_$TAG.log 2> $O/bench_$TAG.err
echo "bench rc=$?"; tail -1 $O/bench_$TAG.log | cut -c1-900
grep "^#" $O/bench_$TAG.err > $O/trace_families_$TAG.txt; head -12 $O/trace_families_$TAG.txt
python tools/trace_table.py $O/trace_$TAG.json 90 > $O/trace_layers_$TAG.txt 2>&1
BC="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-graph"
timeout 170 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3600 --csv \
  --log-file $O/launches_$TAG.csv $BC > $O/ncu_list_$TAG.log 2>&1
echo "ncu list rc=$?"
gzip -f $O/launches_$TAG.csv
