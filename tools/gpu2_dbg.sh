#!/bin/bash
# 2-GPU check of the streamed-input execution under rank skew (soak) + driver-protocol bench
export TFDE_P2P_TIMEOUT_S=5
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29541 tests/mp_worker_fused.py --soak 1000 > gpurun_out/mpw_stream.log 2>&1; echo "stream rc=$?"
grep -v '^\*\|OMP\|^$' gpurun_out/mpw_stream.log | grep -i "error\|fused\|assert\|raise\|timed" | head -12
timeout 200 $TR --master-port 29536 bench.py --gpus 2 --steps 20 --warmup 5 2>gpurun_out/bench_n2_stream.err | tee gpurun_out/bench_n2_stream.json | python tools/bench_summary.py
