#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
line() { python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); print('  %.0f img/s  %.2f ms/step  e2e %.0f (%.2f ms)  roof %s %.3f'%(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['roofline']['kernel'], d['roofline']['frac']))
except Exception as e: print('  no json', e)"; }
timeout 120 python -m pytest tests/test_gpu_fused_paths.py -m gpu -q --timeout=50 -k "test_bn_backward_sums and (ResNet-18 or ResNet-50-50-64 or ResNeXt)" > $O/pytest_sums_quick.log 2>&1; echo "sums test rc=$?"; tail -1 $O/pytest_sums_quick.log
CNB_BN_BWD_SUMS=1 timeout 150 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > $O/bench_sums_pf.log 2> $O/bench_sums_pf.err
echo "bench sums ON (with prefetch) rc=$?"; tail -1 $O/bench_sums_pf.log | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('  %.0f img/s %.2f ms'%(d['value'], d['ms_per_step']))"; grep "^#" $O/bench_sums_pf.err | head -4
CNB_BN_BWD_SUMS=0 timeout 150 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > $O/bench_sums_off2.log 2> $O/bench_sums_off2.err
echo "bench sums OFF rc=$?"; tail -1 $O/bench_sums_off2.log | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('  %.0f img/s %.2f ms'%(d['value'], d['ms_per_step']))"; grep "^#" $O/bench_sums_off2.err | head -4
