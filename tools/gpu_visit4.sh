#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
CUDA_LAUNCH_BLOCKING=1 timeout 200 python tools/find_crash.py ShuffleNetV1 g4 --hw 32 --batch 256 > $O/crash_g4_eval.txt 2>&1; echo "g4 eval rc=$?"; tail -4 $O/crash_g4_eval.txt | cut -c1-200
CUDA_LAUNCH_BLOCKING=1 timeout 200 python tools/find_crash.py ShuffleNetV1 g4 --hw 32 --batch 256 --train > $O/crash_g4_train.txt 2>&1; echo "g4 train rc=$?"; tail -4 $O/crash_g4_train.txt | cut -c1-200
CUDA_LAUNCH_BLOCKING=1 timeout 200 python tools/find_crash.py ShuffleNetV1 g4 --hw 32 --batch 8 > $O/crash_g4_eval_b8.txt 2>&1; echo "g4 eval b8 rc=$?"; tail -2 $O/crash_g4_eval_b8.txt | cut -c1-200
timeout 300 python tools/e2e_probe.py resnext50 > $O/e2e_probe_resnext50.txt 2>&1; echo "probe rc=$?"; tail -22 $O/e2e_probe_resnext50.txt
timeout 300 python tools/e2e_probe.py resnet50 > $O/e2e_probe_resnet50.txt 2>&1; echo "probe rc=$?"; tail -12 $O/e2e_probe_resnet50.txt
