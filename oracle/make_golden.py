"""Generates tests/golden/*.npz by running the REFERENCE's own classes (imported from
/root/reference through oracle/reference_shims.py) on fixed seeds.  TEST INFRASTRUCTURE ONLY.

Run here (the authoring container; /root/reference does not exist on the GPU box):
    python -m oracle.make_golden
The fixtures pin oracle/bertqa_oracle.py (tests/test_oracle_golden.py) and, through it, the
CUDA path.  Shapes are small so that the files stay small, but legal for the kernels
(hidden = 64 * heads).
"""
import os
import sys
import types
from collections import defaultdict

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.reference_shims import StubTokenizer, import_reference  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
SHAPE = dict(num_hidden_layers=2, hidden_size=64, num_attention_heads=1, intermediate_size=128, vocab_size=200,
             max_position_embeddings=128, type_vocab_size=2)
NAME = "bert-golden-tiny"


def model_params(drop=0.0):
    return types.SimpleNamespace(model=NAME, model_name="bert", hidden_dropout_prob=drop,
                                 attention_probs_dropout_prob=drop, layer_norm_eps=1e-12)


def trainer_params(**kw):
    p = types.SimpleNamespace(loss="smooth", smooth_alpha=0.01, focal_alpha=1, focal_gamma=2, w_start=1, w_end=1,
                              w_start_reg=1, w_end_reg=1, w_cls=1, lr=1e-3, weight_decay=1e-2, optimizer="adam",
                              finetune=False, apex_level=None)
    p.__dict__.update(kw)
    return p


def np_dict(d, prefix):
    return {f"{prefix}/{k}": v.detach().cpu().numpy() for k, v in d.items()}


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = import_reference(extra_shapes={NAME: SHAPE})
    torch.manual_seed(1234)
    np.random.seed(1234)

    # ---- (1) forward / loss / gradients of the reference model on a padded batch ---------------
    model = ref.model.BertForQuestionAnswering(model_params(0.0), num_labels=5)
    # non-trivial LayerNorm / bias values so that every gradient path is exercised
    g = torch.Generator().manual_seed(7)
    with torch.no_grad():
        for n, p in model.named_parameters():
            if n.endswith("LayerNorm.weight"):
                p.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g))
            elif n.endswith(".bias") and n.startswith("transformer."):
                p.copy_(0.05 * torch.randn(p.shape, generator=g))
    model.train()  # dropout probabilities are 0: train == eval arithmetic
    tok = StubTokenizer(SHAPE["vocab_size"])
    tok.pad_token_id, tok.unk_token_id, tok.cls_token_id, tok.sep_token_id = 0, 1, 2, 3
    rng = np.random.RandomState(5)
    items = []
    for i, n_doc in enumerate((70, 41, 55)):
        q = rng.randint(4, 200, 12).tolist()
        d = rng.randint(4, 200, n_doc).tolist()
        ids = [tok.cls_token_id] + q + [tok.sep_token_id] + d + [tok.sep_token_id]
        items.append(ref.dataset.DatasetItem(example_id=str(i), input_ids=ids, start_id=[20, -1, 30][i],
                                             end_id=[25, -1, 33][i], label_id=[2, 4, 0][i],
                                             start_position=[0.25, 0.0, 0.6][i], end_position=[0.3, 0.0, 0.7][i]))
    inputs, labels = ref.dataset.collate_fun(items, tok)
    loss = ref.init.init_loss(trainer_params(), defaultdict(lambda: None))
    out = model(**inputs)
    meters = {}
    total = loss(out, labels, avg_meters=meters)
    total.backward()
    fx = {}
    fx.update(np_dict(model.state_dict(), "param"))
    fx.update(np_dict(inputs, "input"))
    fx.update(np_dict(labels, "label"))
    fx.update(np_dict(out, "out"))
    fx.update({f"loss/{k}": np.float32(v) for k, v in meters.items()})
    fx.update({f"grad/{n}": p.grad.detach().numpy() for n, p in model.named_parameters()})
    # 'ce' class loss with label weights on the same logits
    w = torch.tensor([0.1, 0.3, 0.2, 0.25, 0.15])
    loss_ce = ref.init.init_loss(trainer_params(loss="ce"), {"label_weights": w})
    meters_ce = {}
    with torch.no_grad():
        loss_ce({k: v.detach() for k, v in out.items()}, labels, avg_meters=meters_ce)
    fx["ce/label_weights"] = w.numpy()
    fx.update({f"ce/{k}": np.float32(v) for k, v in meters_ce.items()})
    np.savez_compressed(os.path.join(OUT, "ref_forward_backward.npz"), **fx)

    # ---- (2) reference Trainer: 3 optimizer steps on DummyDataset (clip + HF-AdamW shim + schedule)
    torch.manual_seed(99)
    np.random.seed(99)
    model2 = ref.model.BertForQuestionAnswering(model_params(0.0), num_labels=5)
    tp = trainer_params()
    opt = ref.init.init_optimizer(tp, model2)
    ds = ref.dataset.DummyDataset(tok, max_seq_len=48, max_question_len=8, dataset_len=8)
    init_state = {k: v.detach().clone() for k, v in model2.state_dict().items()}
    batches = []

    def recording_collate(items_):
        b = ref.dataset.collate_fun(items_, tok)
        batches.append(b)
        return b

    tr = ref.trainer.Trainer(model=model2, loss=ref.init.init_loss(tp, defaultdict(lambda: None)),
                             collate_fun=recording_collate, optimizer=opt, train_dataset=ds, test_dataset=None,
                             device=torch.device("cpu"), n_epochs=2, train_batch_size=4, batch_split=2, n_jobs=0,
                             warmup_coef=0.5, max_grad_norm=0.5)
    lrs = []
    orig_step = opt.step

    def step_and_record(*a, **k):
        lrs.append(opt.param_groups[0]["lr"])
        return orig_step(*a, **k)

    opt.step = step_and_record
    tr.train()
    fx2 = {}
    fx2.update(np_dict(init_state, "init"))
    fx2.update(np_dict(model2.state_dict(), "final"))
    for i, (inp, lab) in enumerate(batches):
        fx2.update(np_dict(inp, f"batch{i}/input"))
        fx2.update(np_dict(lab, f"batch{i}/label"))
    fx2["n_batches"] = np.int64(len(batches))
    fx2["lrs"] = np.asarray(lrs, dtype=np.float64)
    fx2["hparams"] = np.asarray([tp.lr, tp.weight_decay, 0.5, 2, 0.5], dtype=np.float64)  # lr, wd, clip, split, warmup
    fx2["global_step"] = np.int64(tr.global_step)
    np.savez_compressed(os.path.join(OUT, "ref_trainer_steps.npz"), **fx2)

    # ---- (3) decode: the reference's predictor arithmetic (predictor.py:114-120) on crafted logits
    torch.manual_seed(3)
    B, S = 6, 37
    start = torch.randn(B, S)
    end = torch.randn(B, S)
    start[1, 5] = start[1, 30] = start[1].max() + 1.0  # exact tie: first index must win
    end[2, 0] = end[2].max() + 2.0
    cls = torch.randn(B, 5)
    s_logits, s_ids = torch.max(start, dim=-1)
    e_logits, e_ids = torch.max(end, dim=-1)
    c_prob, c_ids = torch.max(torch.softmax(cls, dim=-1), dim=-1)
    scores = s_logits + e_logits - (start[:, 0] + end[:, 0])
    np.savez_compressed(os.path.join(OUT, "ref_decode.npz"), start=start.numpy(), end=end.numpy(), cls=cls.numpy(),
                        start_ids=s_ids.numpy(), end_ids=e_ids.numpy(), cls_ids=c_ids.numpy(), scores=scores.numpy(),
                        cls_probas=c_prob.numpy())

    # ---- (4) LR schedule of the installed transformers (trainer.py:116-126)
    from transformers import get_linear_schedule_with_warmup
    lin = torch.nn.Linear(1, 1)
    o = torch.optim.SGD(lin.parameters(), lr=1.0)
    sch = get_linear_schedule_with_warmup(o, num_warmup_steps=6, num_training_steps=20)
    vals = []
    for _ in range(24):
        vals.append(o.param_groups[0]["lr"])
        o.step()
        sch.step()
    np.savez_compressed(os.path.join(OUT, "ref_schedule.npz"), lr=np.asarray(vals), warmup=np.int64(6),
                        total=np.int64(20))
    # ---- (5) AdaMod (reference optim.py) and the focal class loss (reference loss.py), run as they are
    import model.trainer.optim as ref_optim
    torch.manual_seed(11)
    names = ["transformer.w.weight", "transformer.w.bias", "classifier.weight", "x.LayerNorm.weight"]
    shapes = [(7, 5), (7,), (3, 7), (7,)]
    params = [torch.nn.Parameter(torch.randn(*sh)) for sh in shapes]
    p0 = [p.detach().clone() for p in params]
    no_decay = ['bias', 'LayerNorm.bias', 'LayerNorm.weight']
    groups = [{'params': [p for n, p in zip(names, params) if not any(nd in n for nd in no_decay)], 'weight_decay': 0.01},
              {'params': [p for n, p in zip(names, params) if any(nd in n for nd in no_decay)], 'weight_decay': 0.0}]
    opt = ref_optim.AdaMod(groups, lr=1e-2)
    grads, traj = [], []
    for step in range(6):
        gs = [torch.randn(*sh) * (0.1 + step) for sh in shapes]
        for p, g in zip(params, gs):
            p.grad = g.clone()
        opt.step()
        grads.append(gs)
        traj.append([p.detach().clone() for p in params])
    fx5 = {"names": np.array(names), "lr": np.float64(1e-2), "weight_decay": np.float64(0.01)}
    for i, n in enumerate(names):
        fx5[f"p0/{n}"] = p0[i].numpy()
        for t in range(6):
            fx5[f"g{t}/{n}"] = grads[t][i].numpy()
            fx5[f"p{t + 1}/{n}"] = traj[t][i].numpy()
        fx5[f"exp_avg_lr/{n}"] = opt.state[params[i]]['exp_avg_lr'].numpy()
    np.savez_compressed(os.path.join(OUT, "ref_adamod.npz"), **fx5)

    torch.manual_seed(12)
    logits = (torch.randn(9, 5) * 2).requires_grad_(True)
    target = torch.tensor([0, 4, 2, -1, 1, 3, -1, 0, 2])
    fx6 = {"logits": logits.detach().numpy(), "target": target.numpy()}
    for alpha, gamma in ((1, 2), (0.25, 3), (2, 0), (1, 1)):
        fl = ref.loss.FocalLossWithLogits(alpha=alpha, gamma=gamma)
        val = fl(logits, target)
        (g,) = torch.autograd.grad(val, logits)
        fx6[f"loss/{alpha}/{gamma}"] = val.detach().numpy()
        fx6[f"grad/{alpha}/{gamma}"] = g.numpy()
    np.savez_compressed(os.path.join(OUT, "ref_focal.npz"), **fx6)
    # ---- (6) the reference's second encoder family: BertForQuestionAnswering over RobertaModel (model.py:9-10),
    # one token type, <pad> = 1, position ids counted over the non-padding tokens.  Rows are padded with <pad>; the
    # reference's collate_fun builds attention_mask = tokens > 0, i.e. RoBERTa padding is attended (kept as is).
    torch.manual_seed(4321)
    rshape = dict(num_hidden_layers=2, hidden_size=64, num_attention_heads=1, intermediate_size=128, vocab_size=200,
                  max_position_embeddings=130, type_vocab_size=1, pad_token_id=1, bos_token_id=0, eos_token_id=2)
    ref_r = import_reference(extra_shapes={"roberta-golden-tiny": rshape})
    mp_r = types.SimpleNamespace(model="roberta-golden-tiny", model_name="roberta", hidden_dropout_prob=0.0,
                                 attention_probs_dropout_prob=0.0, layer_norm_eps=1e-5)
    model_r = ref_r.model.BertForQuestionAnswering(mp_r, num_labels=5)
    g = torch.Generator().manual_seed(8)
    with torch.no_grad():
        for n, p in model_r.named_parameters():
            if n.endswith("LayerNorm.weight"):
                p.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g))
            elif n.endswith(".bias") and n.startswith("transformer."):
                p.copy_(0.05 * torch.randn(p.shape, generator=g))
            elif "position_embeddings" in n or "word_embeddings" in n:
                p.copy_(0.05 * torch.randn(p.shape, generator=g))  # incl. the padding rows: they must not matter twice
    model_r.train()
    tok_r = StubTokenizer(200)
    tok_r.model_name = "roberta"
    tok_r.pad_token_id, tok_r.unk_token_id, tok_r.cls_token_id, tok_r.sep_token_id = 1, 3, 0, 2
    rng = np.random.RandomState(6)
    items_r = []
    for i, n_doc in enumerate((66, 30, 49)):
        qq = rng.randint(4, 200, 10).tolist()
        dd = rng.randint(4, 200, n_doc).tolist()
        ids = [0] + qq + [2] + dd + [2]
        if i == 1:
            ids[5] = 1  # a <pad> id in the middle of a row: it does not advance the position count
        items_r.append(ref_r.dataset.DatasetItem(example_id=str(i), input_ids=ids, start_id=[14, -1, 21][i],
                                                 end_id=[19, -1, 30][i], label_id=[1, 3, 0][i],
                                                 start_position=[0.2, 0.0, 0.5][i], end_position=[0.3, 0.0, 0.8][i]))
    inputs_r, labels_r = ref_r.dataset.collate_fun(items_r, tok_r)
    loss_r = ref_r.init.init_loss(trainer_params(), defaultdict(lambda: None))
    out_r = model_r(**inputs_r)
    meters_r = {}
    total_r = loss_r(out_r, labels_r, avg_meters=meters_r)
    total_r.backward()
    fx7 = {}
    fx7.update(np_dict({k: v for k, v in model_r.state_dict().items() if "position_ids" not in k and "token_type_ids" not in k},
                       "param"))
    fx7.update(np_dict(inputs_r, "input"))
    fx7.update(np_dict(labels_r, "label"))
    fx7.update(np_dict(out_r, "out"))
    fx7.update({f"loss/{k}": np.float32(v) for k, v in meters_r.items()})
    fx7.update({f"grad/{n}": p.grad.detach().numpy() for n, p in model_r.named_parameters()})
    np.savez_compressed(os.path.join(OUT, "ref_roberta_forward_backward.npz"), **fx7)

    # ---- (7) a checkpoint written by the reference's own Trainer.save_state_dict (trainer.py:355-379) after two
    # optimizer steps on DummyDataset: model / optimizer / scheduler / global_step, the wire format the drop-in loads
    torch.manual_seed(77)
    np.random.seed(77)
    model_c = ref.model.BertForQuestionAnswering(model_params(0.0), num_labels=5)
    tp_c = trainer_params()
    opt_c = ref.init.init_optimizer(tp_c, model_c)
    ds_c = ref.dataset.DummyDataset(tok, max_seq_len=48, max_question_len=8, dataset_len=8)
    tr_c = ref.trainer.Trainer(model=model_c, loss=ref.init.init_loss(tp_c, defaultdict(lambda: None)),
                               collate_fun=lambda items_: ref.dataset.collate_fun(items_, tok), optimizer=opt_c,
                               train_dataset=ds_c, test_dataset=None, device=torch.device("cpu"), n_epochs=1,
                               train_batch_size=4, batch_split=1, n_jobs=0, warmup_coef=0.5, max_grad_norm=0.5)
    tr_c.train()
    tr_c.save_state_dict(os.path.join(OUT, "ref_checkpoint.ch"))
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
