#!/bin/bash
# 2-GPU check of the streamed-input execution: multi-GPU tests, driver-protocol bench (streamed and staged inputs)
export TFDE_P2P_TIMEOUT_S=5
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29535 tests/mp_worker_fused.py --soak 500 2>&1 | grep -v '^\*\|OMP\|^$' | tail -4
timeout 200 $TR --master-port 29536 bench.py --gpus $N --steps 20 --warmup 5 2>gpurun_out/bench_n${N}_stream.err | tee gpurun_out/bench_n${N}_stream.json | python tools/bench_summary.py
tail -2 gpurun_out/bench_n${N}_stream.err
TFDE_FUSED_STREAM=0 timeout 200 $TR --master-port 29537 bench.py --gpus $N --steps 20 --warmup 5 2>/dev/null | tee gpurun_out/bench_n${N}_staged.json | python tools/bench_summary.py
