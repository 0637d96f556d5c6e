#!/bin/bash
# Visit 12: the allocation-free BatchNorm forward (tests/test_zz_no_alloc.py) alone in a fresh
# process, then after the other BatchNorm / elementwise C-ABI tests in one process.
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02k_noalloc_alone.log 2>&1
echo "alone exit $?" | tee -a gpurun_out/r02k_noalloc_alone.log
tail -3 gpurun_out/r02k_noalloc_alone.log
timeout 150 python -m pytest tests/test_gpu_ops.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02k_noalloc_after.log 2>&1
echo "after exit $?" | tee -a gpurun_out/r02k_noalloc_after.log
tail -3 gpurun_out/r02k_noalloc_after.log
