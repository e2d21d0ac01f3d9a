"""Pins oracle/bertqa_oracle.py to fixtures produced by the reference's own classes
(oracle/make_golden.py): model forward, all five losses, every gradient, a 4-step Trainer run
(grad accumulation, clip, HF-AdamW shim, warm-up schedule), decode, LR schedule.  CPU only."""
import numpy as np
import torch

from conftest import load_npz
from oracle import bertqa_oracle as O

TINY = O.OracleConfig(num_layers=2, hidden=64, heads=1, intermediate=128, vocab=200, max_pos=128, type_vocab=2,
                      hidden_dropout=0.0, attn_dropout=0.0)


def _group(fx, prefix):
    return {k[len(prefix) + 1:]: torch.from_numpy(np.asarray(v)) for k, v in fx.items() if k.startswith(prefix + "/")}


def test_forward_loss_backward_match_reference():
    fx = load_npz("ref_forward_backward.npz")
    P, inp, lab, out_ref, grad_ref = (_group(fx, p) for p in ("param", "input", "label", "out", "grad"))
    assert list(P.keys()) == O.param_names(TINY)
    total, parts, grads, out = O.loss_and_grads(P, TINY, inp, lab, O.LossConfig())
    for k in out_ref:
        torch.testing.assert_close(out[k], out_ref[k], rtol=2e-5, atol=2e-6)
    for k in ("start_class", "end_class", "start_reg", "end_reg", "cls"):
        assert abs(parts[k].item() - float(fx[f"loss/{k}"])) < 2e-6 * max(1.0, abs(float(fx[f"loss/{k}"])))
    assert abs(total.item() - float(fx["loss/loss"])) < 1e-5
    for k, g in grad_ref.items():
        torch.testing.assert_close(grads[k], g, rtol=2e-4, atol=2e-7, msg=lambda m, k=k: f"{k}: {m}")


ROBERTA_TINY = O.OracleConfig(num_layers=2, hidden=64, heads=1, intermediate=128, vocab=200, max_pos=130, type_vocab=1,
                              hidden_dropout=0.0, attn_dropout=0.0, ln_eps=1e-5, pad_token_id=1, position_mode=1)


def test_roberta_forward_loss_backward_match_reference():
    """The reference's class over RobertaModel (model.py:9-10): one token type, <pad> = 1, position ids =
    padding_idx + running count of non-padding tokens (a <pad> in the middle of a row does not advance it), padding
    rows of the word AND position tables without gradient, attention_mask = tokens > 0 from the reference collate
    (which masks <s> = 0 and attends <pad> = 1: kept as the reference does it)."""
    fx = load_npz("ref_roberta_forward_backward.npz")
    P, inp, lab, out_ref, grad_ref = (_group(fx, p) for p in ("param", "input", "label", "out", "grad"))
    # (the installed transformers registers RoBERTa's token-type table before its position table; the reference's
    # pinned 3.5.1 uses BERT's order, which is also the engine's: compare as sets)
    assert sorted(P.keys()) == sorted(O.param_names(ROBERTA_TINY))
    # the reference's collate masks ids == 0: with RoBERTa that is <s>, while the <pad> = 1 positions are attended
    assert bool((inp["input_ids"] == 1).any()) and torch.equal(inp["attention_mask"], inp["input_ids"] > 0)
    assert int(inp["token_type_ids"].max()) == 0
    total, parts, grads, out = O.loss_and_grads(P, ROBERTA_TINY, inp, lab, O.LossConfig())
    for k in out_ref:
        torch.testing.assert_close(out[k], out_ref[k], rtol=2e-5, atol=2e-6)
    assert abs(total.item() - float(fx["loss/loss"])) < 1e-5
    for k, g in grad_ref.items():
        torch.testing.assert_close(grads[k], g, rtol=2e-4, atol=2e-7, msg=lambda m, k=k: f"{k}: {m}")
    assert float(grads["transformer.embeddings.position_embeddings.weight"][1].abs().max()) == 0.0
    assert float(grads["transformer.embeddings.word_embeddings.weight"][1].abs().max()) == 0.0


def test_reference_checkpoint_wire_format(pkg):
    """A checkpoint written by the reference's own Trainer.save_state_dict (trainer.py:355-379): the drop-in's
    state_dict has exactly its model keys and shapes, and its optimizer section has the per-parameter
    {step, exp_avg, exp_avg_sq} layout that FusedAdamW.load_state_dict consumes (GPU side: tests/test_gpu_model.py)."""
    import importlib
    import os
    import types
    from conftest import GOLDEN
    ck = torch.load(os.path.join(GOLDEN, "ref_checkpoint.ch"), map_location="cpu", weights_only=False)
    assert set(ck) == {"model", "optimizer", "scheduler", "global_step"} and ck["global_step"] == 2
    M = importlib.import_module("bqa_b200.modules.model.model")
    shape = dict(num_layers=2, hidden=64, heads=1, intermediate=128, vocab=200, max_pos=128, type_vocab=2)
    mp = types.SimpleNamespace(model="bert-base-uncased", model_shape=shape, hidden_dropout_prob=0.0,
                               attention_probs_dropout_prob=0.0, layer_norm_eps=1e-12)
    sd = M.BertForQuestionAnswering(mp, 5).state_dict()
    ref_keys = [k for k in ck["model"] if not k.endswith("position_ids") and not k.endswith("embeddings.token_type_ids")]
    assert list(sd.keys()) == ref_keys
    assert all(tuple(sd[k].shape) == tuple(ck["model"][k].shape) for k in sd)
    st = ck["optimizer"]["state"]
    assert len(ck["optimizer"]["param_groups"]) == 2 and len(st) == len(sd)
    assert all(set(v) == {"step", "exp_avg", "exp_avg_sq"} for v in st.values())


def test_ce_class_loss_with_label_weights():
    fx = load_npz("ref_forward_backward.npz")
    out, lab = _group(fx, "out"), _group(fx, "label")
    lc = O.LossConfig(loss="ce", label_weights=torch.from_numpy(fx["ce/label_weights"]))
    total, parts = O.weighted_loss(out, lab, lc)
    assert abs(parts["cls"].item() - float(fx["ce/cls"])) < 1e-6
    assert abs(total.item() - float(fx["ce/loss"])) < 1e-5


def test_trainer_trajectory_matches_reference():
    """Reference Trainer.train(): 2 epochs x 4 micro-batches, batch_split 2 -> 4 optimizer steps."""
    fx = load_npz("ref_trainer_steps.npz")
    P = {k: v.clone() for k, v in _group(fx, "init").items()}
    final = _group(fx, "final")
    lr, wd, clip, split, warmup_coef = fx["hparams"].tolist()
    split = int(split)
    n = int(fx["n_batches"])
    steps = n // split
    warmup = int(steps * warmup_coef)
    state, lrs = {}, []
    acc = None
    for i in range(n):
        inp, lab = _group(fx, f"batch{i}/input"), _group(fx, f"batch{i}/label")
        _, _, grads, _ = O.loss_and_grads(P, TINY, inp, lab, O.LossConfig(), loss_scale=1.0 / split)
        acc = grads if acc is None else {k: acc[k] + grads[k] for k in acc}
        if (i + 1) % split == 0:
            step = (i + 1) // split - 1
            cur = lr * O.linear_schedule_lambda(step, warmup, steps)
            lrs.append(cur)
            O.clip_grad_norm_(acc, clip)
            O.hf_adamw_step(P, acc, state, cur, wd)
            acc = None
    np.testing.assert_allclose(lrs, fx["lrs"], rtol=1e-12)
    assert int(fx["global_step"]) == steps
    for k in final:
        torch.testing.assert_close(P[k], final[k], rtol=1e-4, atol=2e-6, msg=lambda m, k=k: f"{k}: {m}")


def test_decode_matches_reference():
    fx = load_npz("ref_decode.npz")
    d = O.decode({k: torch.from_numpy(fx[k]) for k in ("start", "end", "cls")} |
                 {"start_class": torch.from_numpy(fx["start"]), "end_class": torch.from_numpy(fx["end"])})
    assert (d["start_ids"].numpy() == fx["start_ids"]).all()
    assert (d["end_ids"].numpy() == fx["end_ids"]).all()
    assert (d["cls_ids"].numpy() == fx["cls_ids"]).all()
    assert (d["scores"].numpy() == fx["scores"]).all()  # same float32 association order -> bit-exact


def test_schedule_matches_transformers():
    fx = load_npz("ref_schedule.npz")
    got = [O.linear_schedule_lambda(t, int(fx["warmup"]), int(fx["total"])) for t in range(len(fx["lr"]))]
    np.testing.assert_allclose(got, fx["lr"], rtol=1e-12, atol=0)


def test_dummy_batch_shape_matches_dummy_dataset():
    inputs, labels = O.dummy_batch(3, 96, vocab=500, max_question_len=16)
    ids = inputs["input_ids"]
    assert ids.shape == (3, 96) and (ids[:, 0] == 101).all() and (ids[:, 17] == 102).all() and (ids[:, -1] == 102).all()
    assert inputs["attention_mask"].all()
    assert (inputs["token_type_ids"][:, :18] == 0).all() and (inputs["token_type_ids"][:, 18:] == 1).all()
    assert labels["end_class"].tolist() == [95] * 3 and labels["end_reg"].tolist() == [1.0] * 3


def test_bf16_mode_stays_close_to_fp32():
    fx = load_npz("ref_forward_backward.npz")
    P, inp = _group(fx, "param"), _group(fx, "input")
    a = O.forward(P, TINY, inp["input_ids"], inp["attention_mask"], inp["token_type_ids"])
    b = O.forward(P, TINY, inp["input_ids"], inp["attention_mask"], inp["token_type_ids"], bf16=True)
    for k in a:
        assert (a[k] - b[k]).abs().max() < 0.05


def test_adamod_matches_reference_class():
    """oracle.adamod_step against six steps of the reference's own AdaMod (optim.py:42-100), two
    weight-decay groups as built by init.py:121-129."""
    fx = load_npz("ref_adamod.npz")
    names = [str(n) for n in fx["names"]]
    P = {n: torch.from_numpy(np.asarray(fx[f"p0/{n}"])).clone() for n in names}
    state = {}
    for t in range(6):
        G = {n: torch.from_numpy(np.asarray(fx[f"g{t}/{n}"])) for n in names}
        O.adamod_step(P, G, state, lr=float(fx["lr"]), weight_decay=float(fx["weight_decay"]))
        for n in names:
            torch.testing.assert_close(P[n], torch.from_numpy(np.asarray(fx[f"p{t + 1}/{n}"])), rtol=1e-6, atol=1e-7,
                                       msg=lambda m, n=n, t=t: f"step {t} {n}: {m}")
    for n in names:
        torch.testing.assert_close(state[n]["exp_avg_lr"], torch.from_numpy(np.asarray(fx[f"exp_avg_lr/{n}"])),
                                   rtol=1e-6, atol=1e-9)


def test_focal_class_loss_matches_reference_class():
    """oracle.class_loss(loss='focal') and its autograd gradient against the reference's
    FocalLossWithLogits (loss.py:57-71), including ignored rows (target -1) and gamma = 0."""
    fx = load_npz("ref_focal.npz")
    target = torch.from_numpy(np.asarray(fx["target"]))
    for alpha, gamma in ((1, 2), (0.25, 3), (2, 0), (1, 1)):
        logits = torch.from_numpy(np.asarray(fx["logits"])).clone().requires_grad_(True)
        val = O.class_loss(logits, target, O.LossConfig(loss="focal", focal_alpha=alpha, focal_gamma=gamma))
        (g,) = torch.autograd.grad(val, logits)
        assert abs(val.item() - float(fx[f"loss/{alpha}/{gamma}"])) < 1e-6
        torch.testing.assert_close(g, torch.from_numpy(np.asarray(fx[f"grad/{alpha}/{gamma}"])), rtol=1e-5, atol=1e-7)
