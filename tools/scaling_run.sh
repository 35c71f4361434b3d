#!/bin/bash
# 1/2/4/8-GPU headline runs (ours and the torch stand-in) + 8-rank validation of the P2P kernels.
# Usage (on a box with 8 GPUs):  bash tools/scaling_run.sh [steps]
STEPS=${1:-300}
OUT=gpurun_out
mkdir -p $OUT
NG=$(nvidia-smi -L | wc -l)
python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29544 tests/mp_worker_p2p.py > $OUT/p2p_all.log 2>&1
echo "rc=$?" >> $OUT/p2p_all.log
for impl in ours torch; do
  for n in 1 2 4 8; do
    [ $n -gt $NG ] && continue
    if [ $n = 1 ]; then
      python bench.py --impl $impl --gpus 1 --steps $STEPS --warmup 20 > $OUT/scale_${impl}_$n.log 2>&1
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2956$n bench.py --impl $impl --gpus $n --steps $STEPS --warmup 20 > $OUT/scale_${impl}_$n.log 2>&1
    fi
  done
done
grep -E "P2P|FAIL" $OUT/p2p_all.log | head -5
