#!/bin/bash
# Final 8-GPU records: headline bench (driver protocol) and BASELINE.json config 3 (global batch 8192, bf16 policy, Adam)
export TFDE_P2P_TIMEOUT_S=5
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
mkdir -p gpurun_out
timeout 200 $TR --master-port 29536 bench.py --gpus $N --steps 20 --warmup 5 2>gpurun_out/final_n$N.err | tee gpurun_out/final_bench_n$N.json | python tools/bench_summary.py
timeout 300 $TR --master-port 29541 bench.py --gpus $N --batch-per-gpu 1024 --dtype bf16 --optimizer adam --steps 20 --warmup 5 2>gpurun_out/final_c3_n$N.err | tee gpurun_out/final_config3_n$N.json | python tools/bench_summary.py
tail -3 gpurun_out/final_c3_n$N.err
timeout 100 $TR --master-port 29537 tools/fused_step_phases.py --steps 8 2>&1 | grep -v '^\*\|OMP\|^$' | tail -20 | tee gpurun_out/final_phases_n$N.txt
