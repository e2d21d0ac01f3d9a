"""N-rank data-parallel optimizer step on NCCL vs the single-rank step on the concatenated batch.

Launched by tests/test_gpu_entrypoints.py (and by hand) as

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tests/dist_parity_worker.py [--shape base|tiny] [--per-rank B] [--steps K] [--out file.json]

What is checked (reference semantics: DistributedDataParallel averages the ranks' gradients,
modules/model/trainer/trainer.py:136-142, so N ranks with b samples each must produce the weights
one process produces on the N*b samples, because every loss term is a mean over the batch):
  * after the bucketed all-reduce every rank holds the same gradient, and it equals the
    single-rank gradient of the concatenated batch within the gradient tolerance of
    tests/test_gpu_model.py (2e-2 norm-wise, 5e-2 per tensor + floor);
  * after K clip + AdamW steps the weight UPDATE equals the single-rank update within 10 %
    (Adam normalises every element to ~lr, so elements whose gradient is rounding noise differ).
Dropout is off so that the two runs are comparable.  No oracle is involved: both sides run the
library; what is tested is the exchange (bucket plan, side stream ordering, mean).
"""
import argparse
import importlib
import json
import os
import sys
import types

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SHAPES = {
    "tiny": dict(num_layers=2, hidden=128, heads=2, intermediate=512, vocab=1000, max_pos=512, type_vocab=2),
    "base": dict(num_layers=12, hidden=768, heads=12, intermediate=3072, vocab=30522, max_pos=512, type_vocab=2),
}


def rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-20)).item()


def make_batch(B, S, vocab, seed, device):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(1000 if vocab > 2000 else 103, vocab, (B, S), generator=g)
    ids[:, 0] = 101
    ids[:, 65] = 102
    ids[:, -1] = 102
    tt = (torch.arange(S)[None, :] > 65).long().expand(B, S).contiguous()
    inputs = {"input_ids": ids.to(device), "token_type_ids": tt.to(device), "attention_mask": None}
    labels = {"start_class": torch.randint(0, S, (B,), generator=g).to(device),
              "end_class": torch.randint(0, S, (B,), generator=g).to(device),
              "start_reg": torch.rand(B, generator=g).to(device),
              "end_reg": torch.rand(B, generator=g).to(device),
              "cls": torch.randint(0, 5, (B,), generator=g).to(device)}
    return inputs, labels


def take(d, lo, hi):
    return {k: (v[lo:hi].contiguous() if v is not None else None) for k, v in d.items()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="base")
    ap.add_argument("--per-rank", type=int, default=4)
    ap.add_argument("--seq", type=int, default=384)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group(backend="nccl")

    from __graft_entry__ import load_package
    load_package()
    M = importlib.import_module("bqa_b200.modules.model.model")
    I = importlib.import_module("bqa_b200.modules.init")
    T = importlib.import_module("bqa_b200.modules.model.trainer.trainer")
    shape = SHAPES[args.shape]

    def build(local_rank):
        mp = types.SimpleNamespace(model="bert-base-uncased", model_shape=shape, hidden_dropout_prob=0.0,
                                   attention_probs_dropout_prob=0.0, layer_norm_eps=1e-12)
        torch.manual_seed(11)
        model = M.BertForQuestionAnswering(mp, 5).to(dev)
        hp = types.SimpleNamespace(lr=1e-4, weight_decay=1e-2, optimizer="adam", finetune=False, apex_level=None)
        opt = I.init_optimizer(hp, model)
        lp = types.SimpleNamespace(loss="smooth", smooth_alpha=0.01, focal_alpha=1, focal_gamma=2, w_start=1, w_end=1,
                                   w_start_reg=1, w_end_reg=1, w_cls=1)
        tr = T.Trainer(model=model, loss=I.init_loss(lp, {"label_weights": None}), collate_fun=None, optimizer=opt,
                       device=dev, local_rank=local_rank, n_epochs=1, train_batch_size=args.per_rank, batch_split=1,
                       n_jobs=0, warmup_coef=0.0, max_grad_norm=1.0)
        tr.set_train()
        opt.zero_grad()
        return model, opt, tr

    model, opt, tr = build(rank)
    assert tr.reducer is not None and len(tr.reducer.plan) >= (2 if args.shape == "base" else 1)
    w0 = model.engine.params.clone()
    b = args.per_rank
    report = {"world": world, "shape": args.shape, "per_rank": b, "seq": args.seq, "steps": args.steps,
              "buckets": len(tr.reducer.plan), "grad": [], "ok": True}
    grads_n = []
    for s in range(args.steps):
        inputs, labels = make_batch(world * b, args.seq, shape["vocab"], 100 + s, dev)
        tr.train_micro_batch(take(inputs, rank * b, (rank + 1) * b), take(labels, rank * b, (rank + 1) * b),
                             last_of_step=True)
        torch.cuda.synchronize()
        g = model.engine.grads.clone()
        # every rank must hold the same averaged gradient
        g0 = g.clone()
        dist.broadcast(g0, src=0)
        same = bool(torch.equal(g, g0))
        flags = torch.tensor([int(same)], device=dev)
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        report["grad"].append({"identical_on_all_ranks": bool(flags.item())})
        grads_n.append(g)
        tr.optimizer_step()
    torch.cuda.synchronize()
    wN = model.engine.params.clone()
    dist.barrier()

    if rank == 0:
        model1, opt1, tr1 = build(-1)
        assert tr1.reducer.world == 1  # local_rank -1: no exchange, although this process has a process group
        assert torch.equal(model1.engine.params, w0), "replicas must start from rank 0's weights"
        for s in range(args.steps):
            inputs, labels = make_batch(world * b, args.seq, shape["vocab"], 100 + s, dev)
            tr1.train_micro_batch(inputs, labels, last_of_step=True)
            torch.cuda.synchronize()
            g1 = model1.engine.grads
            eng = model1.engine
            entry = report["grad"][s]
            entry["global_rel"] = rel(grads_n[s], g1)
            floor = 1e-3 * max(eng.view(g1, k).norm().item() for k in eng.layout)
            worst, worst_k = 0.0, None
            for k in eng.layout:
                a, r = eng.view(grads_n[s], k), eng.view(g1, k)
                excess = max(0.0, (a - r).norm().item() - floor) / (r.norm().item() + 1e-30)
                if excess > worst:
                    worst, worst_k = excess, k
            entry["worst_tensor"] = {"name": worst_k, "rel": worst}
            if not (entry["identical_on_all_ranks"] and entry["global_rel"] < 2e-2 and worst < 5e-2):
                report["ok"] = False
            tr1.optimizer_step()
        torch.cuda.synchronize()
        w1 = model1.engine.params
        report["update_rel"] = rel(wN - w0, w1 - w0)
        report["weights_max_abs_diff"] = (wN - w1).abs().max().item()
        if not report["update_rel"] < 0.1:
            report["ok"] = False
        print("DDP_PARITY " + json.dumps(report), flush=True)
        if args.out:
            os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
            with open(args.out, "w") as f:
                json.dump(report, f, indent=1)
    ok = torch.tensor([int(report["ok"])], device=dev)
    dist.broadcast(ok, src=0)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok.item() else 1)


if __name__ == "__main__":
    main()
