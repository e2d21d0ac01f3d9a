"""Runs the REFERENCE's own training loop on the host cores and times it.  TEST INFRASTRUCTURE ONLY
(bench.py's `cpu_baseline` / `--impl reference` legs; the product never imports this).

What runs is the unmodified reference: `Trainer.train()` -> `Trainer._train`
(modules/model/trainer/trainer.py:266-300) over `DummyDataset` (dummy_dataset.py:36-51) with
`collate_fun`, `BertForQuestionAnswering` (model.py:13-73), `WeightedLoss` (loss.py:74-106),
`clip_grad_norm_` and the optimizer built by `init_optimizer` (init.py:125-138), imported from
/root/reference or from the staged copy oracle/_ref (oracle/stage_ref.py) through the stand-ins of
oracle/reference_shims.py for what this image lacks (nltk, configargparse, transformers.AdamW,
offline BertModel.from_pretrained).
"""
import os
import statistics
import time
import types
from collections import defaultdict


def time_reference_train(steps, warmup, batch, seq, threads=None):
    """`warmup + steps` optimizer steps of BASELINE configs[0] (BERT-base random init, fp32, dropout 0.1,
    loss=smooth, lr 1e-5, weight decay 1e-4, clip 1, batch_split 1) -> dict(value, unit, cores, kind, sample)."""
    import numpy as np
    import torch

    from oracle.reference_shims import StubTokenizer, import_reference
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    ref = import_reference()
    torch.manual_seed(1234)
    np.random.seed(1234)
    mp = types.SimpleNamespace(model="bert-base-uncased", model_name="bert", hidden_dropout_prob=0.1,
                               attention_probs_dropout_prob=0.1, layer_norm_eps=1e-12)
    tp = types.SimpleNamespace(loss="smooth", smooth_alpha=0.01, focal_alpha=1, focal_gamma=2, w_start=1, w_end=1,
                               w_start_reg=1, w_end_reg=1, w_cls=1, lr=1e-5, weight_decay=1e-4, optimizer="adam",
                               finetune=False, apex_level=None)
    model = ref.model.BertForQuestionAnswering(mp, num_labels=5)
    opt = ref.init.init_optimizer(tp, model)
    tok = StubTokenizer()
    n_steps = warmup + steps
    ds = ref.dataset.DummyDataset(tok, max_seq_len=seq, max_question_len=64, dataset_len=batch * n_steps)
    tr = ref.trainer.Trainer(model=model, loss=ref.init.init_loss(tp, defaultdict(lambda: None)),
                             collate_fun=lambda items: ref.dataset.collate_fun(items, tok), optimizer=opt,
                             train_dataset=ds, test_dataset=None, device=torch.device("cpu"), n_epochs=1,
                             train_batch_size=batch, batch_split=1, n_jobs=0, warmup_coef=0.6, max_grad_norm=1)
    stamps = []
    orig_step = opt.step

    def step_and_stamp(*a, **k):
        r = orig_step(*a, **k)
        stamps.append(time.perf_counter())
        return r

    opt.step = step_and_stamp
    t0 = time.perf_counter()
    tr.train()
    assert len(stamps) == n_steps, (len(stamps), n_steps)
    # a step = everything between two optimizer steps (data loading, forward, loss, backward, clip, AdamW, zero_grad)
    times = [b - a for a, b in zip([t0] + stamps[:-1], stamps)][warmup:]
    sec = statistics.median(times)
    return {"value": round(batch / sec, 3), "unit": "samples/s", "cores": cores, "kind": "reference",
            "sample": f"{steps} optimizer steps of batch {batch}, seq {seq}, BERT-base fp32, dropout 0.1: the reference's "
                      f"own Trainer._train on DummyDataset (modules/model/trainer/trainer.py:266-300, staged copy), "
                      f"median step {sec:.2f} s after {warmup} warm-up"}
