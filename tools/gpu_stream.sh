#!/bin/bash
export TFDE_P2P_TIMEOUT_S=5
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_fused_step_gpu.py -q -x 2>&1 | tail -5
timeout 200 python bench.py --steps 20 --warmup 5 2>gpurun_out/bench_stream.err | tee gpurun_out/bench_stream_n1.json | python tools/bench_summary.py
tail -3 gpurun_out/bench_stream.err
TFDE_FUSED_STREAM=0 timeout 200 python bench.py --steps 20 --warmup 5 2>gpurun_out/bench_staged.err | tee gpurun_out/bench_staged_n1.json | python tools/bench_summary.py
timeout 200 python bench.py --steps 200 --warmup 5 2>>gpurun_out/bench_stream.err | tee gpurun_out/bench_stream_k200.json | python tools/bench_summary.py
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
