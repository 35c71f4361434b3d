#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
R="python tools/bench_resnet.py --batch 64 --dtype bf16 --steps 20 --warmup 4"
timeout 300 $R 2>gpurun_out/resnet_v3.err | tee gpurun_out/resnet_v3.json
TFDE_FUSE_BN_STATS=1 timeout 300 $R 2>gpurun_out/resnet_v3_bnstats.err | tee gpurun_out/resnet_v3_bnstats.json
timeout 60 python tools/one_gemm.py 200704 256 64 persistent | tee gpurun_out/one_gemm_k64.txt
timeout 60 python tools/one_gemm.py 200704 64 256 persistent | tee -a gpurun_out/one_gemm_k64.txt
timeout 60 python tools/one_conv.py 64 56 56 64 64 | tee gpurun_out/one_conv_v3.txt
timeout 300 ncu --set full --clock-control none --import-source on -k regex:persistent -s 4 -c 1 -f -o gpurun_out/gemm_k64_n256_v3 python tools/one_gemm.py 200704 256 64 persistent > gpurun_out/ncu_gemm.log 2>&1
tail -1 gpurun_out/ncu_gemm.log
