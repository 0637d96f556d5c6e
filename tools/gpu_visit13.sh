#!/bin/bash
# Visit 13: CNB_SKIP_ZERO_GRADS (conv functions take None for the gradients of their partial-sum
# outputs instead of materialised zeros) — A/B of the ResNet-50 step, then the tests that exercise
# the conv functions with the switch on.
O=gpurun_out; mkdir -p $O
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e"
CNB_SKIP_ZERO_GRADS=1 timeout 70 $B > $O/r02l_bench_on.log 2> $O/r02l_bench_on.err; echo "on rc=$?"; tail -1 $O/r02l_bench_on.log | cut -c1-260
CNB_SKIP_ZERO_GRADS=0 timeout 70 $B > $O/r02l_bench_off.log 2> $O/r02l_bench_off.err; echo "off rc=$?"; tail -1 $O/r02l_bench_off.log | cut -c1-260
CNB_SKIP_ZERO_GRADS=1 timeout 75 python -m pytest tests/test_gpu_fused_paths.py tests/test_gpu_dense.py tests/test_gpu_models.py -m gpu -x -q --timeout=60 > $O/r02l_pytest_on.log 2>&1
echo "pytest rc=$?"; tail -2 $O/r02l_pytest_on.log | cut -c1-300
