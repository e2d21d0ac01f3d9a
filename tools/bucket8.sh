#!/bin/bash
# N-GPU sweep of the gradient bucket size (usage under gpurun --gpus 8: [NGPU=8] bash tools/bucket8.sh 50 100)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
port=29720
for mb in "$@"; do
  port=$((port + 1))
  BQA_BUCKET_MB=$mb python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NGPU:-8} --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus ${NGPU:-8} --steps 20 --warmup 3 > gpurun_out/bucket_$mb.json 2> gpurun_out/bucket_$mb.err
  echo "bucket_mb=$mb rc=$? $(python -c "import json; d=json.load(open('gpurun_out/bucket_$mb.json')); print(d['value'], d['ms_per_step'], {k: v['ms_per_step'] for k, v in d['roofline']['by_kind'].items()})" 2>&1 | tail -1)"
done
