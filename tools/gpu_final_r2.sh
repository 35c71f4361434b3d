#!/bin/bash
# Round-2 closing 1-GPU session: GPU test-suite, headline / Model B / ResNet-50 benches, ResNet per-kernel launch list.
export TFDE_P2P_TIMEOUT_S=5
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -4
timeout 200 python bench.py --steps 20 --warmup 5 2>gpurun_out/bench1.err | tee gpurun_out/bench_n1.json | python tools/bench_summary.py
timeout 200 python bench.py --model b --batch-per-gpu 128 --steps 20 --warmup 5 2>gpurun_out/benchb.err | tee gpurun_out/bench_b_n1.json | python tools/bench_summary.py
timeout 300 python tools/bench_resnet.py --batch 64 --dtype bf16 --steps 20 --warmup 4 2>gpurun_out/resnet_v4.err | tee gpurun_out/resnet_v4.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
  --log-file gpurun_out/resnet_launches.csv python tools/bench_resnet.py --batch 64 --dtype bf16 --steps 1 --warmup 3 > gpurun_out/resnet_ncu_run.log 2>&1
python tools/agg_step.py gpurun_out/resnet_launches.csv > gpurun_out/resnet_breakdown.md 2>&1; tail -42 gpurun_out/resnet_breakdown.md
