#!/bin/bash
# 1-GPU session for the implicit-GEMM convolutions / TMA-store epilogue: bring-up check, the GPU test-suite,
# ResNet-50 step with each switch off, loss parity, per-kernel launch list of one step, ncu of one conv kernel.
mkdir -p gpurun_out
timeout 150 python tools/debug_igemm.py > gpurun_out/igemm_debug.txt 2>&1; rc=$?
tail -4 gpurun_out/igemm_debug.txt
if [ $rc -ne 0 ]; then echo "debug_igemm rc=$rc: stopping"; exit 0; fi
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
timeout 200 python tools/resnet_parity.py 2>&1 | tail -4
R="python tools/bench_resnet.py --batch 64 --dtype bf16 --steps 20 --warmup 4"
for cfg in "1 1" "1 0" "0 1"; do
  set -- $cfg
  echo "igemm=$1 tma_store=$2"
  TFDE_CONV_IGEMM=$1 TFDE_TMA_STORE=$2 timeout 300 $R 2>gpurun_out/resnet_$1$2.err | tee gpurun_out/resnet_ig$1_st$2.json
done
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
  --log-file gpurun_out/resnet_launches.csv python tools/bench_resnet.py --batch 64 --dtype bf16 --steps 1 --warmup 3 > gpurun_out/resnet_ncu_run.log 2>&1
tail -1 gpurun_out/resnet_ncu_run.log
python tools/agg_step.py gpurun_out/resnet_launches.csv > gpurun_out/resnet_breakdown.md 2>&1; tail -45 gpurun_out/resnet_breakdown.md
