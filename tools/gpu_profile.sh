#!/bin/bash
# GPU-box visit: (optional) tests, bench, then the ncu launch list and one --set full capture.
# usage: bash tools/gpu_profile.sh <tag> [kernel-regex for the full capture] [pytest -k expr|none]
set -u
TAG=${1:-r01}; KRE=${2:-bn_bwd}; TESTS=${3:-all}
mkdir -p gpurun_out
if [ "$TESTS" != "none" ]; then
  if [ "$TESTS" = "all" ]; then
    timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_$TAG.log 2>&1
  else
    timeout 2400 python -m pytest tests -m gpu -q -k "$TESTS" > gpurun_out/pytest_gpu_$TAG.log 2>&1
  fi
  echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_$TAG.log; tail -15 gpurun_out/pytest_gpu_$TAG.log
fi
timeout 900 python bench.py --steps 10 --warmup 3 --trace-out gpurun_out/trace_$TAG.json > gpurun_out/bench_$TAG.log 2> gpurun_out/bench_$TAG.err
echo "bench rc=$?"; tail -2 gpurun_out/bench_$TAG.log; grep "^#" gpurun_out/bench_$TAG.err | head -14
# launch list (same command as the bench, fewer steps; numbers printed under ncu are not bench values)
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv \
  --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e \
  > gpurun_out/ncu_list_$TAG.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:$KRE" -s 10 -c 6 \
  -o gpurun_out/full_$TAG -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e \
  > gpurun_out/ncu_full_$TAG.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | tail -12
