#!/bin/bash
# 1-GPU session: full GPU test-suite, headline bench (ours + graphed torch stand-in), Model B bench, ncu of the
# whole-step kernel (only after the same program exited 0 without ncu).
export TFDE_P2P_TIMEOUT_S=5
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -15
timeout 200 python bench.py --steps 20 --warmup 5 2>gpurun_out/bench1.err | tee gpurun_out/bench_n1.json | python tools/bench_summary.py
timeout 200 python bench.py --impl torch --steps 20 --warmup 5 2>gpurun_out/torch1.err | tee gpurun_out/torch_graph_n1.json | python tools/bench_summary.py
timeout 200 python bench.py --model b --batch-per-gpu 128 --steps 20 --warmup 5 2>gpurun_out/benchb.err | tee gpurun_out/bench_b_n1.json | python tools/bench_summary.py
tail -3 gpurun_out/benchb.err
timeout 100 python tools/fused_step_phases.py --steps 8 > gpurun_out/phases_n1.txt 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fused_step -s 2 -c 1 -f -o gpurun_out/fused_step python tools/fused_step_phases.py --steps 8 > gpurun_out/ncu_run.log 2>&1
tail -3 gpurun_out/ncu_run.log; cat gpurun_out/phases_n1.txt | tail -15
