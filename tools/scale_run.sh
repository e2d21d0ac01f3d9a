#!/bin/bash
# bench.py at N GPUs of one box the way the driver launches it (run on the GPU box): tools/scale_run.sh <ngpus>
N=${1:-8}
if [ "$N" = 1 ]; then
  python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline 2>&1 | tail -1
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline 2>&1 | tail -1
fi
