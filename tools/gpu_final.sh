#!/bin/bash
# Round-end GPU visit: full parity suite, the bench line (both arms), ncu launch list, DRAM traffic
# of the BatchNorm-backward family and one --set full capture.  usage: gpu_final.sh <tag> [skip-tests]
set -u
TAG=${1:-r01z}; SKIP=${2:-no}
mkdir -p gpurun_out
if [ "$SKIP" = "no" ]; then
  timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_$TAG.log 2>&1
  echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_$TAG.log; tail -4 gpurun_out/pytest_gpu_$TAG.log
fi
timeout 600 python bench.py --steps 20 --warmup 5 --trace-out gpurun_out/trace_$TAG.json > gpurun_out/bench_$TAG.log 2> gpurun_out/bench_$TAG.err
echo "bench rc=$?"; tail -1 gpurun_out/bench_$TAG.log | cut -c1-300; grep "^#" gpurun_out/bench_$TAG.err | head -12
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$TAG.log 2> gpurun_out/bench_ref_$TAG.err
echo "reference rc=$?"; tail -1 gpurun_out/bench_ref_$TAG.log | cut -c1-300
# the ncu passes enqueue eagerly (--no-graph) so that every kernel is a separate launch record
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4500 --csv \
  --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-graph \
  > gpurun_out/ncu_list_$TAG.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
  -k regex:bn_bwd -c 700 --csv --log-file gpurun_out/dram_bn_bwd_$TAG.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/ncu_dram_$TAG.log 2>&1; echo "ncu dram rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bn_bwd -s 330 -c 4 \
  -o gpurun_out/full_$TAG -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-graph \
  > gpurun_out/ncu_full_$TAG.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | tail -8
