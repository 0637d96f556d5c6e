#!/bin/bash
# Short round-end visit (no conv sweep): op + model parity, bench line, smoke, ncu launch list,
# DRAM traffic of the BatchNorm-backward family.  usage: gpu_final2.sh <tag>
set -u
TAG=${1:-z}
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_ops.py tests/test_gpu_models.py -q > gpurun_out/pytest_$TAG.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/pytest_$TAG.log
timeout 200 python bench.py --steps 20 --warmup 5 --trace-out gpurun_out/trace_$TAG.json > gpurun_out/bench_$TAG.log 2> gpurun_out/bench_$TAG.err
echo "bench rc=$?"; tail -1 gpurun_out/bench_$TAG.log | cut -c1-330; grep "^#" gpurun_out/bench_$TAG.err | head -12
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_$TAG.log
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 4500 --csv \
  --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-graph \
  > gpurun_out/ncu_list_$TAG.log 2>&1; echo "ncu list rc=$?"
timeout 200 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
  -k regex:bn_bwd -c 700 --csv --log-file gpurun_out/dram_bn_bwd_$TAG.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-graph > gpurun_out/ncu_dram_$TAG.log 2>&1; echo "ncu dram rc=$?"
