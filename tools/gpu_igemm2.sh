#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 300 python tools/bench_resnet.py --batch 64 --dtype bf16 --steps 20 --warmup 4 2>gpurun_out/resnet_v2.err | tee gpurun_out/resnet_v2.json
for s in "64 56 56 64 64" "64 28 28 128 128" "64 14 14 256 256" "64 7 7 512 512"; do timeout 60 python tools/one_conv.py $s; done 2>&1 | tee gpurun_out/one_conv.txt
timeout 60 python tools/one_gemm.py 200704 256 64 persistent | tee gpurun_out/one_gemm_k64.txt && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:persistent -s 4 -c 1 -f -o gpurun_out/gemm_k64_n256 python tools/one_gemm.py 200704 256 64 persistent > gpurun_out/ncu_gemm.log 2>&1
tail -2 gpurun_out/ncu_gemm.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:persistent -s 6 -c 1 -f -o gpurun_out/conv_igemm_fwd_c64 python tools/one_conv.py 64 56 56 64 64 > gpurun_out/ncu_conv.log 2>&1
tail -2 gpurun_out/ncu_conv.log
