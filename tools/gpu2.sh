#!/bin/bash
# N-GPU validation + bench + phase profile of the whole-step kernel (run under gpurun --gpus N)
export TFDE_P2P_TIMEOUT_S=5
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29535 tests/mp_worker_fused.py --soak 2000 2>&1 | grep -v '^\*\|OMP\|^$' | tail
timeout 200 $TR --master-port 29536 bench.py --gpus $N --steps 20 --warmup 5 2>&1 | python tools/bench_summary.py
timeout 100 $TR --master-port 29537 tools/fused_step_phases.py --steps 8 2>&1 | grep -v '^\*\|OMP\|^$' | tail -16
