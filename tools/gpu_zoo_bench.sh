#!/bin/bash
# one bench line per BASELINE.json config model (1 GPU).  usage: gpu_zoo_bench.sh <tag> <models...>
set -u
TAG=$1; shift
mkdir -p gpurun_out
for m in "$@"; do
  timeout 600 python bench.py --model $m --steps 5 --warmup 3 --no-cpu-baseline --trace-out gpurun_out/trace_${TAG}_$m.json \
    > gpurun_out/bench_${TAG}_$m.log 2> gpurun_out/bench_${TAG}_$m.err
  echo "== $m rc=$?"; tail -1 gpurun_out/bench_${TAG}_$m.log | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print(d['metric'], '%.0f img/s'%d['value'], '%.2f ms/step'%d['ms_per_step'], 'e2e %.0f'%d['e2e']['value'], d['config']['workload'])
except Exception as e: print('no json', e)"
  grep "^#" gpurun_out/bench_${TAG}_$m.err | head -6; grep -v "^#" gpurun_out/bench_${TAG}_$m.err | tail -3 | cut -c1-300
done
