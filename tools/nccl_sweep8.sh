#!/bin/bash
# 8-GPU bench under different NCCL CTA limits (run on the GPU box)
N=${1:-8}
run() {
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 15 --warmup 4 --no-cpu-baseline 2>/dev/null | tail -1 | \
    python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$*', d['ms_per_step'], d['value'], d['e2e']['value'], round(d['roofline']['by_kind']['gemm_wgrad']['ms_per_step'],3))"
}
run BQA_TAG=default
run NCCL_MAX_CTAS=4
run NCCL_MAX_CTAS=8
run NCCL_MAX_CTAS=16
run NCCL_MIN_CTAS=32
