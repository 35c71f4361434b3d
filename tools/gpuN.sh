#!/bin/bash
# N-GPU session: whole-step kernel validation, headline bench (ours + graphed torch stand-in), phase profile,
# all-reduce sweep 1 KB - 1 GB. Usage: bash tools/gpuN.sh N [sweep]
export TFDE_P2P_TIMEOUT_S=5
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
mkdir -p gpurun_out
timeout 300 $TR --master-port 29535 tests/mp_worker_fused.py --soak 1000 2>&1 | grep -v '^\*\|OMP\|^$' | tail -8
timeout 200 $TR --master-port 29536 bench.py --gpus $N --steps 20 --warmup 5 2>gpurun_out/bench_n$N.err | tee gpurun_out/bench_n$N.json | python tools/bench_summary.py
timeout 200 $TR --master-port 29538 bench.py --gpus $N --steps 200 --warmup 5 --steps-per-execution 50 2>>gpurun_out/bench_n$N.err | tee gpurun_out/bench_n${N}_k200.json | python tools/bench_summary.py
timeout 200 $TR --master-port 29539 bench.py --impl torch --gpus $N --steps 20 --warmup 5 2>gpurun_out/torch_n$N.err | tee gpurun_out/torch_graph_n$N.json | python tools/bench_summary.py
timeout 100 $TR --master-port 29537 tools/fused_step_phases.py --steps 8 2>&1 | grep -v '^\*\|OMP\|^$' | tail -16 | tee gpurun_out/phases_n$N.txt
if [ "$2" == "sweep" ]; then
  timeout 900 $TR --master-port 29540 tests/mp_worker_p2p.py --sweep-only 2>&1 | grep -v '^\*\|OMP\|^$' | tail -45
fi
