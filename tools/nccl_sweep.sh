#!/bin/bash
# N-GPU bench under different NCCL CTA limits (run on the GPU box): tools/nccl_sweep.sh <ngpus>
N=${1:-2}
run() {
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | tail -1 | \
    python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$*', d['ms_per_step'], d['value'], d['e2e']['value'])"
}
run BQA_TAG=default
for c in 1 2 4 8 16 32; do run NCCL_MAX_CTAS=$c; done
run NCCL_NVLS_ENABLE=0
run NCCL_ALGO=Ring
