#!/bin/bash
# 1-GPU session for the implicit-GEMM convolutions: bring-up check, unit tests, regression tests of the bf16 path,
# ResNet-50 step with the implicit GEMM on and off.
mkdir -p gpurun_out
timeout 150 python tools/debug_igemm.py > gpurun_out/igemm_debug.txt 2>&1; rc=$?
tail -45 gpurun_out/igemm_debug.txt
if [ $rc -ne 0 ] && [ $rc -ne 1 ]; then echo "debug_igemm rc=$rc (hang or crash): stopping"; exit 0; fi
timeout 400 python -m pytest tests/test_conv_igemm_gpu.py tests/test_act16_gpu.py tests/test_mixed_gpu.py -q -x 2>&1 | tail -8
if [ $rc -eq 0 ]; then
  for ig in 1 0; do
    TFDE_CONV_IGEMM=$ig timeout 300 python tools/bench_resnet.py --batch 64 --dtype bf16 --steps 10 --warmup 4 2>gpurun_out/resnet_ig$ig.err | tee gpurun_out/resnet_ig$ig.json
    tail -2 gpurun_out/resnet_ig$ig.err
  done
fi
