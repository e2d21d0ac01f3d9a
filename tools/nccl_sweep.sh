#!/bin/bash
# N-GPU bench under different NCCL CTA limits and SM budgets (run on the GPU box): tools/nccl_sweep.sh <ngpus>
N=${1:-2}
run() {
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | tail -1 | \
    python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$*', d['ms_per_step'], d['value'], d['e2e']['value'])"
}
python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('single', d['ms_per_step'], d['value'])"
run BQA_TAG=default
for b in 140 132 124 116; do run BQA_SM_BUDGET=$b; done
for c in 8 16; do for b in 148 140 132; do run NCCL_MAX_CTAS=$c BQA_SM_BUDGET=$b; done; done
