#!/bin/bash
# Visit 14: tensor-map encode from a thread without a current context (new test, default switches),
# the stem tests in default mode, then the conv-function tests with CNB_SKIP_ZERO_GRADS=1.
O=gpurun_out; mkdir -p $O
timeout 60 python -m pytest tests/test_gpu_fused_paths.py -m gpu -x -q -k "thread or rgb_stem" > $O/r02m_thread_default.log 2>&1
echo "default rc=$?"; tail -2 $O/r02m_thread_default.log | cut -c1-300
CNB_SKIP_ZERO_GRADS=1 timeout 100 python -m pytest tests/test_gpu_fused_paths.py tests/test_gpu_dense.py tests/test_gpu_models.py -m gpu -x -q --timeout=60 > $O/r02m_pytest_on.log 2>&1
echo "pytest-on rc=$?"; tail -2 $O/r02m_pytest_on.log | cut -c1-300
