#!/bin/bash
# One 8-GPU box: N-rank NCCL parity, then the three workloads at N = 8, the bf16 gradient exchange option, N = 1 on the
# same box (SCALE8_SWEEP=1 adds a small sweep of NCCL's CTA budget).
# usage (under gpurun --gpus 8): bash tools/scale8.sh
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() {  # name, port, extra env..., -- bench args
  name=$1; port=$2; shift 2
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus 8 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  echo "$name rc=$? $(python -c "import json,sys; d=json.load(open('gpurun_out/$name.json')); print(d['value'], d['ms_per_step'], d['roofline']['by_kind'].get('gemm_wgrad'))" 2>&1 | tail -1)"
}
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29700 \
    tests/dist_parity_worker.py --shape base --out gpurun_out/r02_ddp_parity_8gpu.json > gpurun_out/ddp8.log 2>&1
echo "ddp parity rc=$? $(grep -c DDP_PARITY gpurun_out/ddp8.log)"
run r02_n8_base 29701 -- --steps 20 --warmup 3
if [ "$SCALE8_SWEEP" = "1" ]; then  # NCCL CTA budget sweep (recorded once in profiles/r02_scale8.json: limiting it lengthens the step)
run r02_n8_base_cta8 29702 NCCL_MAX_CTAS=8 -- --steps 20 --warmup 3
run r02_n8_base_cta4 29703 NCCL_MAX_CTAS=4 -- --steps 20 --warmup 3
run r02_n8_base_cta16 29704 NCCL_MAX_CTAS=16 -- --steps 20 --warmup 3
fi
run r02_n8_base_bf16x 29707 BQA_GRAD_EXCHANGE=bf16 -- --steps 20 --warmup 3
run r02_n8_large 29705 -- --steps 10 --warmup 3 --workload large512
run r02_n8_decode 29706 -- --steps 16 --warmup 3 --workload decode512
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r02_n1_samebox.json 2> gpurun_out/r02_n1_samebox.err
echo "n1 rc=$?"
