"""Parity of the whole step on the GPU: engine forward / loss / backward vs the oracle (which is
pinned to the reference by tests/test_oracle_golden.py), on the reference-generated golden batch,
on DummyDataset batches at BERT-base size, through the drop-in model class, the Trainer and the
Predictor.

Tolerances.  The CUDA path computes in bf16 with fp32 accumulation; compared with the oracle in
its bf16-faithful mode (same rounding points) the remaining differences are accumulation order
and bf16 rounding of back-propagated activations: outputs within 1e-2 relative (norm-wise),
losses within 3e-3, the concatenated gradient within 2e-2, each tensor within 5e-2 of its norm
plus a floor of 1e-3 of the largest tensor-gradient norm (gradients that are exactly zero in
exact arithmetic — span bias, key biases — are pure rounding noise in both implementations).
Against the fp32 reference fixture the bounds are 3x looser.  Decoded indices are bit-exact."""
import importlib
import types

import numpy as np
import pytest
import torch

from conftest import load_npz
from oracle import bertqa_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"

TINY = dict(num_layers=2, hidden=128, heads=2, intermediate=512, vocab=1000, max_pos=512, type_vocab=2)
GOLD = dict(num_layers=2, hidden=64, heads=1, intermediate=128, vocab=200, max_pos=128, type_vocab=2)
BASE = dict(num_layers=12, hidden=768, heads=12, intermediate=3072, vocab=30522, max_pos=512, type_vocab=2)
# two layers at the BERT-large widths (BASELINE configs[3]: S = 512, 16 heads, H = 1024, I = 4096)
LARGE2 = dict(num_layers=2, hidden=1024, heads=16, intermediate=4096, vocab=2000, max_pos=512, type_vocab=2)
# BASELINE configs[3] at full depth: bert-large-uncased
LARGE24 = dict(num_layers=24, hidden=1024, heads=16, intermediate=4096, vocab=30522, max_pos=512, type_vocab=2)
OUT_KEYS = ("start_class", "end_class", "cls", "start_reg", "end_reg")


def rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-20)).item()


def make_engine(pkg, shape, p_drop, P):
    E = pkg.engine
    eng = E.Engine(E.make_config(shape, hidden_dropout=p_drop, attn_dropout=p_drop), DEV)
    for k, v in eng.param_views().items():
        v.copy_(P[k])
    eng.sync_weights()
    return eng


def perturbed_params(cfgo, seed=1234):
    P = O.init_params(cfgo, seed, device=DEV)
    g = torch.Generator().manual_seed(5)
    for k in P:
        if k.endswith("LayerNorm.weight"):
            P[k] = (1.0 + 0.1 * torch.randn(P[k].shape, generator=g)).to(DEV)
        elif k.endswith(".bias") and k.startswith("transformer."):
            P[k] = (0.02 * torch.randn(P[k].shape, generator=g)).to(DEV)
    return P


def check_grads(eng, grads, glob_tol=2e-2, tensor_tol=5e-2):
    gv = eng.grad_views()
    ref_flat = torch.cat([grads[k].flatten() for k in grads])
    got_flat = torch.cat([gv[k].flatten() for k in grads])
    assert rel(got_flat, ref_flat) < glob_tol
    floor = 1e-3 * max(g.norm().item() for g in grads.values())
    for k, g in grads.items():
        err = (gv[k].float() - g).norm().item()
        assert err <= tensor_tol * g.norm().item() + floor, f"{k}: err {err:.3e} vs norm {g.norm().item():.3e}"


@pytest.mark.parametrize("shape,B,S,p_drop,use_mask", [(TINY, 4, 128, 0.0, False), (TINY, 4, 128, 0.1, False),
                                                       (TINY, 3, 200, 0.1, True), (BASE, 4, 384, 0.0, False),
                                                       (LARGE2, 12, 512, 0.1, True),
                                                       # BASELINE configs[1] exactly (the shape bench.py times): every
                                                       # persistent kernel walks several tiles per CTA, the GEMMs run 48
                                                       # m-tiles through the cluster-launch-control scheduler, and the
                                                       # fused side outputs (row-dot D, GELU-backward column sums,
                                                       # attention bias gradients) run at full size
                                                       (BASE, 32, 384, 0.1, False),
                                                       # BASELINE configs[3] exactly: 24 layers, B = 16, S = 512
                                                       (LARGE24, 16, 512, 0.1, True)])
def test_step_matches_oracle(pkg, shape, B, S, p_drop, use_mask):
    E = pkg.engine
    cfgo = O.OracleConfig(hidden_dropout=p_drop, attn_dropout=p_drop, **shape)
    P = perturbed_params(cfgo)
    eng = make_engine(pkg, shape, p_drop, P)
    inputs, labels = O.dummy_batch(B, S, vocab=shape["vocab"], device=DEV)
    if use_mask:
        lens = torch.randint(S // 2, S + 1, (B,), device=DEV)
        m = torch.arange(S, device=DEV)[None, :] < lens[:, None]
        inputs["input_ids"] = inputs["input_ids"] * m
        inputs["attention_mask"] = m
        labels["start_class"][0] = -1
    else:
        inputs["attention_mask"] = None
    seed, step = 4242, 3
    outs = eng.forward(inputs["input_ids"], inputs["token_type_ids"], inputs["attention_mask"], training=True,
                       seed=seed, step=step)
    losses = eng.loss(outs, labels, E.make_loss_config(), with_grad=True).clone()
    eng.backward()
    torch.cuda.synchronize()
    drop = O.DropSpec(seed, step) if p_drop > 0 else None
    total, parts, grads, oout = O.loss_and_grads(P, cfgo, inputs, labels, O.LossConfig(), drop=drop, bf16=True)
    for k, t in zip(OUT_KEYS, outs):
        # rounding accumulates with depth: 1e-2 at 2 layers, 2.5e-2 at 12, 4e-2 at 24
        assert rel(t, oout[k]) < (4e-2 if shape is LARGE24 else 2.5e-2 if shape is BASE else 1e-2), k
    ref_l = [parts[k].item() for k in ("start_class", "end_class", "start_reg", "end_reg", "cls")] + [total.item()]
    for a, b in zip(losses.tolist(), ref_l):
        assert abs(a - b) <= 3e-3 * max(1.0, abs(b))
    check_grads(eng, grads)
    # exact-arithmetic zeros: softmax is invariant to a key bias, CE gradients sum to zero over positions
    gv = eng.grad_views()
    gmax = max(g.norm().item() for g in grads.values())
    assert gv["position_outputs.bias"].abs().max().item() < 1e-4 * max(1.0, gmax)


# Measured on B200 (round 2, dropout off, B = 4, S = 384, BERT-base, perturbed LayerNorm weights and biases), CUDA
# path (bf16 operands, fp32 accumulation) against the oracle in plain fp32 — i.e. against the arithmetic of the
# reference's own PyTorch path (profiles/r02_parity_fp32_base.json): outputs 0.9e-3 .. 1.6e-2 norm-wise (the span
# logits are the worst), the six losses within 6.8e-4 relative, concatenated gradient 9.8e-3, worst single tensor
# 8.5e-3 (token-type table).  The stated tolerance is ~2x that; the test writes the measured values to
# gpurun_out/parity_fp32_base.json when that directory exists.
FP32_BOUNDS = dict(outputs=3e-2, losses=2e-3, grad_global=2e-2, grad_tensor=2.5e-2)


def test_step_within_stated_fp32_tolerance_at_bert_base(pkg):
    """The north star's "within a stated fp32/bf16 tolerance of the reference's own PyTorch path" at BERT-base
    size: no bf16 emulation on the oracle side."""
    import json
    import os
    E = pkg.engine
    cfgo = O.OracleConfig(hidden_dropout=0.0, attn_dropout=0.0, **BASE)
    P = perturbed_params(cfgo)
    eng = make_engine(pkg, BASE, 0.0, P)
    inputs, labels = O.dummy_batch(4, 384, vocab=BASE["vocab"], device=DEV)
    inputs["attention_mask"] = None
    outs = eng.forward(inputs["input_ids"], inputs["token_type_ids"], None, training=True, seed=1, step=1)
    losses = eng.loss(outs, labels, E.make_loss_config(), with_grad=True).clone()
    eng.backward()
    torch.cuda.synchronize()
    total, parts, grads, oout = O.loss_and_grads(P, cfgo, inputs, labels, O.LossConfig(), drop=None, bf16=False)
    gv = eng.grad_views()
    measured = {"outputs": {k: rel(t, oout[k]) for k, t in zip(OUT_KEYS, outs)}}
    ref_l = [parts[k].item() for k in ("start_class", "end_class", "start_reg", "end_reg", "cls")] + [total.item()]
    measured["losses"] = [abs(a - b) / max(1.0, abs(b)) for a, b in zip(losses.tolist(), ref_l)]
    ref_flat = torch.cat([grads[k].flatten() for k in grads])
    got_flat = torch.cat([gv[k].flatten() for k in grads])
    measured["grad_global"] = rel(got_flat, ref_flat)
    floor = 1e-3 * max(g.norm().item() for g in grads.values())
    per = {k: max(0.0, (gv[k].float() - g).norm().item() - floor) / (g.norm().item() + 1e-30) for k, g in grads.items()}
    worst = max(per, key=per.get)
    measured["grad_tensor_worst"] = {"name": worst, "rel": per[worst]}
    if os.path.isdir("gpurun_out"):
        with open("gpurun_out/parity_fp32_base.json", "w") as f:
            json.dump(measured, f, indent=1)
    assert max(measured["outputs"].values()) < FP32_BOUNDS["outputs"], measured
    assert max(measured["losses"]) < FP32_BOUNDS["losses"], measured
    assert measured["grad_global"] < FP32_BOUNDS["grad_global"], measured
    assert per[worst] < FP32_BOUNDS["grad_tensor"], measured


def test_step_matches_reference_golden(pkg):
    """Inputs, weights, outputs, losses and gradients produced by the reference's own classes (fp32)."""
    E = pkg.engine
    fx = load_npz("ref_forward_backward.npz")
    grp = lambda p: {k[len(p) + 1:]: torch.from_numpy(np.asarray(v)).to(DEV) for k, v in fx.items()  # noqa: E731
                     if k.startswith(p + "/")}
    P, inp, lab, out_ref, grad_ref = (grp(p) for p in ("param", "input", "label", "out", "grad"))
    eng = make_engine(pkg, GOLD, 0.0, P)
    outs = eng.forward(inp["input_ids"], inp["token_type_ids"], inp["attention_mask"], training=True, seed=1, step=1)
    losses = eng.loss(outs, lab, E.make_loss_config(), with_grad=True).clone()
    eng.backward()
    torch.cuda.synchronize()
    valid = inp["attention_mask"]
    for k, t in zip(OUT_KEYS, outs):
        a, b = (t[valid], out_ref[k][valid]) if t.dim() == 2 and t.shape == valid.shape else (t, out_ref[k])
        assert rel(a, b) < 3e-2, k
    ref_l = [float(fx[f"loss/{k}"]) for k in ("start_class", "end_class", "start_reg", "end_reg", "cls", "loss")]
    for a, b in zip(losses.tolist(), ref_l):
        assert abs(a - b) <= 1e-2 * max(1.0, abs(b))
    check_grads(eng, grad_ref, glob_tol=6e-2, tensor_tol=0.15)
    # padding row of the word table: nn.Embedding(padding_idx=0) -> exactly zero gradient
    assert eng.grad_views()["transformer.embeddings.word_embeddings.weight"][0].abs().max().item() == 0.0


ROBERTA_GOLD = dict(num_layers=2, hidden=64, heads=1, intermediate=128, vocab=200, max_pos=130, type_vocab=1,
                    pad_token_id=1, position_mode=1)


def test_roberta_step_matches_reference_golden(pkg):
    """The reference's second `--model` choice (model.py:9-10, RobertaModel): one token type, <pad> = 1, position ids
    counted over the non-padding tokens on the device, padding rows of the word and position tables without
    gradient.  Inputs, weights, outputs, losses and gradients come from the reference's own class (fp32)."""
    E = pkg.engine
    fx = load_npz("ref_roberta_forward_backward.npz")
    grp = lambda p: {k[len(p) + 1:]: torch.from_numpy(np.asarray(v)).to(DEV) for k, v in fx.items()  # noqa: E731
                     if k.startswith(p + "/")}
    P, inp, lab, out_ref, grad_ref = (grp(p) for p in ("param", "input", "label", "out", "grad"))
    eng = E.Engine(E.make_config(ROBERTA_GOLD, ln_eps=1e-5, hidden_dropout=0.0, attn_dropout=0.0), DEV)
    for k, v in eng.param_views().items():
        v.copy_(P[k])
    eng.sync_weights()
    outs = eng.forward(inp["input_ids"], inp["token_type_ids"], inp["attention_mask"], training=True, seed=1, step=1)
    losses = eng.loss(outs, lab, E.make_loss_config(), with_grad=True).clone()
    eng.backward()
    torch.cuda.synchronize()
    eng.check_inputs()
    valid = inp["attention_mask"]
    for k, t in zip(OUT_KEYS, outs):
        a, b = (t[valid], out_ref[k][valid]) if t.dim() == 2 and t.shape == valid.shape else (t, out_ref[k])
        assert rel(a, b) < 3e-2, k
    ref_l = [float(fx[f"loss/{k}"]) for k in ("start_class", "end_class", "start_reg", "end_reg", "cls", "loss")]
    for a, b in zip(losses.tolist(), ref_l):
        assert abs(a - b) <= 1e-2 * max(1.0, abs(b))
    check_grads(eng, grad_ref, glob_tol=6e-2, tensor_tol=0.15)
    gv = eng.grad_views()
    assert gv["transformer.embeddings.word_embeddings.weight"][1].abs().max().item() == 0.0
    assert gv["transformer.embeddings.position_embeddings.weight"][1].abs().max().item() == 0.0
    assert gv["transformer.embeddings.position_embeddings.weight"][0].abs().max().item() == 0.0  # never indexed


def test_checkpoint_written_by_the_reference_loads(pkg, tmp_path):
    """tests/golden/ref_checkpoint.ch was written by the reference's own Trainer.save_state_dict after two
    optimizer steps (oracle/make_golden.py).  Trainer.load_state_dict restores weights, AdamW moments, scheduler
    and global_step from it; init_model(checkpoint=...) restores the weights alone (init.py:43-48, strict=False)."""
    import os
    from conftest import GOLDEN
    I = importlib.import_module("bqa_b200.modules.init")
    T = importlib.import_module("bqa_b200.modules.model.trainer.trainer")
    D = importlib.import_module("bqa_b200.modules.model.dataset")
    path = os.path.join(GOLDEN, "ref_checkpoint.ch")
    ck = torch.load(path, map_location="cpu", weights_only=False)
    model = _model(pkg, GOLD)
    hp = types.SimpleNamespace(lr=1e-3, weight_decay=1e-2, optimizer="adam", finetune=False, apex_level=None)
    opt = I.init_optimizer(hp, model)
    tok = I.DummyTokenizer(vocab_size=200)
    ds = D.DummyDataset(tok, max_seq_len=48, max_question_len=8, dataset_len=8)
    tr = T.Trainer(model=model, loss=_loss_obj(), collate_fun=I.init_collate_fun(tok), optimizer=opt, train_dataset=ds,
                   device=torch.device(DEV), n_epochs=1, train_batch_size=4, batch_split=1, n_jobs=0, warmup_coef=0.5,
                   max_grad_norm=0.5)
    tr.load_state_dict(path)
    assert tr.global_step == 2
    sd = model.state_dict()
    for k in sd:
        assert torch.equal(sd[k].cpu(), ck["model"][k]), k
    eng = model.engine
    for i, (p, key) in enumerate(opt._keys):
        saved = ck["optimizer"]["state"][i]
        assert torch.equal(eng.view(opt._m, key).cpu(), saved["exp_avg"]), key
        assert torch.equal(eng.view(opt._v, key).cpu(), saved["exp_avg_sq"]), key
        assert opt.state[p]["step"] == saved["step"] == 2
    assert tr.scheduler.state_dict()["last_epoch"] == ck["scheduler"]["last_epoch"]
    # training goes on from the restored state (the restored schedule has reached its end, lr = 0: set one)
    assert opt.param_groups[0]["lr"] == 0.0
    for g_ in opt.param_groups:
        g_["lr"] = 1e-3
    opt.zero_grad()
    tr.set_train()
    inputs, labels = O.dummy_batch(4, 48, vocab=200, max_question_len=8, device=DEV)
    tr.train_micro_batch(inputs, labels, last_of_step=True)
    tr.optimizer_step()
    torch.cuda.synchronize()
    assert torch.isfinite(eng.params).all() and not torch.equal(model.state_dict()["classifier.1.weight"].cpu(),
                                                               ck["model"]["classifier.1.weight"])
    # weights-only restore of init_model
    mp = types.SimpleNamespace(model="bert-base-uncased", model_shape=GOLD, hidden_dropout_prob=0.0,
                               attention_probs_dropout_prob=0.0, layer_norm_eps=1e-12, vocab_file=None)
    model2, _ = I.init_model(mp, checkpoint=path, device=torch.device(DEV))
    for k, v in model2.state_dict().items():
        assert torch.equal(v.cpu(), ck["model"][k]), k


def test_ce_loss_with_label_weights_matches_reference_golden(pkg):
    E = pkg.engine
    L = pkg._lib
    fx = load_npz("ref_forward_backward.npz")
    M = importlib.import_module("bqa_b200.modules.model.model.loss")
    out = {k[4:]: torch.from_numpy(v).to(DEV) for k, v in fx.items() if k.startswith("out/")}
    lab = {k[6:]: torch.from_numpy(v).to(DEV) for k, v in fx.items() if k.startswith("label/")}
    w = torch.from_numpy(fx["ce/label_weights"]).to(DEV)
    lc = E.make_loss_config(loss="ce", cls_weight=w)
    losses, _ = M._run_loss_kernel([out[k] for k in ("start_class", "end_class", "start_reg", "end_reg", "cls")], lab, lc,
                                   False)
    got = losses.tolist()
    for i, k in enumerate(("start_class", "end_class", "start_reg", "end_reg", "cls", "loss")):
        assert abs(got[i] - float(fx[f"ce/{k}"])) < 2e-5 * max(1.0, abs(float(fx[f"ce/{k}"]))), k


def _model(pkg, shape, p_drop=0.0, seed=11):
    M = importlib.import_module("bqa_b200.modules.model.model")
    mp = types.SimpleNamespace(model="bert-base-uncased", model_shape=shape, hidden_dropout_prob=p_drop,
                               attention_probs_dropout_prob=p_drop, layer_norm_eps=1e-12)
    torch.manual_seed(seed)
    return M.BertForQuestionAnswering(mp, 5).to(DEV)


def _loss_obj(**kw):
    I = importlib.import_module("bqa_b200.modules.init")
    p = types.SimpleNamespace(loss="smooth", smooth_alpha=0.01, focal_alpha=1, focal_gamma=2, w_start=1, w_end=1,
                              w_start_reg=1, w_end_reg=1, w_cls=1)
    p.__dict__.update(kw)
    return I.init_loss(p, {"label_weights": None})


def test_dropin_autograd_path_equals_fused_path(pkg):
    """model(**inputs) -> WeightedLoss -> (loss / split).backward() gives the same gradients as the
    fused forward_loss_backward, and they land in p.grad of the reference-named parameters."""
    model = _model(pkg, TINY)
    loss = _loss_obj()
    inputs, labels = O.dummy_batch(4, 128, vocab=1000, device=DEV)
    model.train()
    meters = {}
    out = model(**inputs)
    assert set(out.keys()) == {"start_class", "end_class", "start_reg", "end_reg", "cls"}
    full = loss(out, labels, avg_meters=meters)
    (full / 2).backward()
    torch.cuda.synchronize()
    g_auto = {k: p.grad.clone() for k, p in model.named_parameters()}
    assert set(meters) == {"start_class", "end_class", "start_reg", "end_reg", "cls", "loss"}
    assert abs(meters["loss"] - full.item()) < 1e-6
    for p in model.parameters():
        p.grad.zero_()
    lc = loss.loss_config(loss_scale=0.5, device=torch.device(DEV))
    _, losses = model.forward_loss_backward(inputs, labels, lc)
    torch.cuda.synchronize()
    assert abs(losses[5].item() - full.item()) < 1e-6
    floor = 1e-5 * max(g.abs().max().item() for g in g_auto.values())  # exact-zero gradients are rounding noise
    for k, p in model.named_parameters():
        torch.testing.assert_close(p.grad, g_auto[k], rtol=2e-3, atol=floor, msg=lambda m, k=k: f"{k}: {m}")
    # eval path: no grad, no dropout, deterministic
    model.eval()
    with torch.no_grad():
        a = model(**inputs)
        b = model(**inputs)
    assert all(torch.equal(a[k], b[k]) for k in a)


@pytest.mark.parametrize("optimizer", ["adam", "adamod"])
def test_training_trajectory_matches_oracle(pkg, optimizer):
    """4 optimizer steps through the package's Trainer (grad accumulation over 2 micro-batches, clip,
    fused AdamW / AdaMod with the two decay groups, warm-up schedule) vs the oracle loop on the same batches."""
    I = importlib.import_module("bqa_b200.modules.init")
    T = importlib.import_module("bqa_b200.modules.model.trainer.trainer")
    shape = TINY
    model = _model(pkg, shape)
    P = {k: v.detach().clone() for k, v in model.state_dict().items()}
    hp = types.SimpleNamespace(lr=1e-3, weight_decay=1e-2, optimizer=optimizer, finetune=False, apex_level=None)
    opt = I.init_optimizer(hp, model)
    assert type(opt).__name__ == ("FusedAdamW" if optimizer == "adam" else "AdaMod")
    tr = T.Trainer(model=model, loss=_loss_obj(), collate_fun=None, optimizer=opt, device=torch.device(DEV), n_epochs=1,
                   train_batch_size=4, batch_split=2, n_jobs=0, warmup_coef=0.0, max_grad_norm=0.5)
    steps, warmup = 4, 2
    tr.scheduler = torch.optim.lr_scheduler.LambdaLR(opt, T.linear_warmup_lambda(warmup, steps))
    tr.set_train()
    opt.zero_grad()
    cfgo = O.OracleConfig(hidden_dropout=0.0, attn_dropout=0.0, **shape)
    state, norms = {}, []
    for s in range(steps):
        acc = None
        for mb in range(2):
            inputs, labels = O.dummy_batch(2, 128, vocab=1000, seed=100 + 2 * s + mb, device=DEV)
            losses = tr.train_micro_batch(inputs, labels, last_of_step=(mb == 1))
            total, _, grads, _ = O.loss_and_grads(P, cfgo, inputs, labels, O.LossConfig(), bf16=True, loss_scale=0.5)
            assert abs(losses[5].item() - total.item()) < 3e-3 * max(1.0, total.item())
            acc = grads if acc is None else {k: acc[k] + grads[k] for k in acc}
        lr = 1e-3 * O.linear_schedule_lambda(s, warmup, steps)
        assert abs(opt.param_groups[0]["lr"] - lr) < 1e-12
        norms.append(O.clip_grad_norm_(acc, 0.5).item())
        if optimizer == "adam":
            O.hf_adamw_step(P, acc, state, lr, 1e-2)
        else:
            O.adamod_step(P, acc, state, lr, 1e-2)
        tr.optimizer_step()
        assert abs(opt.grad_norm() - norms[-1]) < 2e-2 * norms[-1]
    torch.cuda.synchronize()
    sd = model.state_dict()
    upd_ref = torch.cat([(P[k] - v0).flatten() for k, v0 in ((k, _model(pkg, shape).state_dict()[k]) for k in P)])
    upd_got = torch.cat([(sd[k] - _model(pkg, shape).state_dict()[k]).flatten() for k in P])
    assert rel(upd_got, upd_ref) < 0.1          # the UPDATE (not the weights) within 10 %
    for k in P:
        assert (sd[k] - P[k]).abs().max().item() < 2e-3, k
    assert (model.engine.grads == 0).all()      # optimizer.zero_grad() semantics
    assert all(opt.state[p]["step"] == steps for p in model.parameters())
    if optimizer == "adamod":
        k0 = "transformer.encoder.layer.0.attention.self.query.weight"
        got = opt.state[dict(model.named_parameters())[k0]]["exp_avg_lr"]
        assert rel(got, state[k0]["exp_avg_lr"]) < 5e-2


def test_checkpoint_roundtrip_and_state_dict_layout(pkg, tmp_path):
    I = importlib.import_module("bqa_b200.modules.init")
    T = importlib.import_module("bqa_b200.modules.model.trainer.trainer")
    hp = types.SimpleNamespace(lr=1e-3, weight_decay=1e-2, optimizer="adam", finetune=False, apex_level=None)

    def make(touch_optimizer=True):
        model = _model(pkg, TINY)
        opt = I.init_optimizer(hp, model)
        tr = T.Trainer(model=model, loss=_loss_obj(), collate_fun=None, optimizer=opt, device=torch.device(DEV),
                       n_epochs=1, train_batch_size=2, batch_split=1, n_jobs=0, warmup_coef=0.0, max_grad_norm=1.0)
        tr.set_train()
        if touch_optimizer:
            opt.zero_grad()
        return model, opt, tr

    model, opt, tr = make()
    inputs, labels = O.dummy_batch(2, 128, vocab=1000, device=DEV)
    for _ in range(2):
        tr.train_micro_batch(inputs, labels, last_of_step=True)
        tr.optimizer_step()
        tr.global_step += 1
    path = tmp_path / "last.ch"
    tr.save_state_dict(path)
    ck = torch.load(path, map_location="cpu")
    assert set(ck) == {"model", "optimizer", "scheduler", "global_step"} and ck["global_step"] == 2
    assert list(ck["model"].keys()) == O.param_names(O.OracleConfig(**TINY))
    st0 = ck["optimizer"]["state"][0]
    assert set(st0) == {"step", "exp_avg", "exp_avg_sq"} and len(ck["optimizer"]["param_groups"]) == 2
    # the resume path of train.py --last: the optimizer has never been stepped or zeroed (its flat moment
    # buffers do not exist yet) when the checkpoint is loaded
    model2, opt2, tr2 = make(touch_optimizer=False)
    tr2.load_state_dict(path)
    eng2 = model2.engine
    for i, (p, key) in enumerate(opt2._keys):
        saved = ck["optimizer"]["state"][i]
        assert torch.equal(eng2.view(opt2._m, key).cpu(), saved["exp_avg"]), key
        assert torch.equal(eng2.view(opt2._v, key).cpu(), saved["exp_avg_sq"]), key
        assert opt2.state[p]["step"] == 2
    assert float(opt2._m.abs().sum()) > 0.0
    opt2.zero_grad()
    for (k, a), (_, b) in zip(model.state_dict().items(), model2.state_dict().items()):
        assert torch.equal(a, b), k
    # same next step from the restored state (moments restored into the flat buffers)
    for t in (tr, tr2):
        t.train_micro_batch(inputs, labels, last_of_step=True)
        t.optimizer_step()
    torch.cuda.synchronize()
    # Split-K reduce-adds and atomics make gradients reproducible only to rounding; Adam (eps 1e-6) turns
    # rounding noise on exact-zero gradients into updates of up to ~lr * 1e-2.  Lost moments would show up
    # as differences of the order of lr = 1e-3 on most elements.
    for (k, a), (_, b) in zip(model.state_dict().items(), model2.state_dict().items()):
        if k.endswith("key.bias") or k == "position_outputs.bias":
            continue  # exact-zero gradients: their Adam updates are driven by rounding noise alone
        assert (a - b).abs().max().item() < 1e-4, k
        assert (a - b).abs().mean().item() < 2e-6, k


def test_out_of_range_ids_are_reported_and_harmless(pkg):
    """nn.Embedding raises IndexError for ids outside the table (the reference's behaviour).  The kernels clamp
    such ids, flag them, the loss turns NaN, Engine.check_inputs() raises IndexError, and the backward writes
    nothing outside the gradient of row 0 (no out-of-bounds scatter)."""
    E = pkg.engine
    cfgo = O.OracleConfig(hidden_dropout=0.0, attn_dropout=0.0, **TINY)
    P = O.init_params(cfgo, 3, device=DEV)
    eng = make_engine(pkg, TINY, 0.0, P)
    inputs, labels = O.dummy_batch(2, 128, vocab=1000, device=DEV)
    ids = inputs["input_ids"].clone()
    ids[1, 7] = 1000 + 12345            # vocab is 1000
    tt = inputs["token_type_ids"].clone()
    tt[0, 3] = 5                         # type_vocab is 2
    outs = eng.forward(ids, tt, None, training=True, seed=1, step=1)
    losses = eng.loss(outs, labels, E.make_loss_config(), with_grad=True).clone()
    eng.backward()
    torch.cuda.synchronize()
    assert torch.isnan(losses).all()
    with pytest.raises(IndexError):
        eng.check_inputs()
    assert torch.isfinite(eng.grads).all()
    # a clean batch afterwards clears the flag
    outs = eng.forward(inputs["input_ids"], inputs["token_type_ids"], None, training=True, seed=1, step=2)
    losses = eng.loss(outs, labels, E.make_loss_config(), with_grad=True).clone()
    torch.cuda.synchronize()
    assert torch.isfinite(losses).all()
    eng.check_inputs()


def test_predictor_decodes_like_reference_arithmetic(pkg):
    """Predictor over synthetic padded chunks: device decode == predictor.py:114-120 on the same logits,
    and the per-document candidate rule is applied."""
    Pm = importlib.import_module("bqa_b200.modules.model.inference.predictor")
    V = importlib.import_module("bqa_b200.modules.validate")
    I = importlib.import_module("bqa_b200.modules.init")
    model = _model(pkg, dict(TINY, vocab=30522))
    tok = I.DummyTokenizer()
    ds = V.SyntheticChunkDataset(5, max_seq_len=256, max_question_len=32, doc_stride=64, max_doc_len=600)
    pred = Pm.Predictor(model, DEV, collate_fun=I.init_collate_fun(tok, return_items=True), batch_size=8, n_jobs=0)
    pred(ds, save_dump=True)
    assert pred.dump
    model.eval()
    for scores, s_ids, e_ids, c_ids, items in pred.dump:
        inputs, _, _ = I.init_collate_fun(tok, return_items=True)(items)
        with torch.no_grad():
            out = model(**{k: v.to(DEV) for k, v in inputs.items()})
        d = O.decode({k: v.cpu() for k, v in out.items()})
        assert (d["start_ids"].numpy() == s_ids).all() and (d["end_ids"].numpy() == e_ids).all()
        assert (d["cls_ids"].numpy() == c_ids).all() and (d["scores"].numpy() == scores).all()
    for item_id, cand in pred.candidates.items():
        item = pred.items[item_id]
        assert cand.start_id <= cand.end_id and cand.start_id >= item.question_len + 2


def test_bad_inputs_fail_loudly(pkg):
    model = _model(pkg, TINY)
    with pytest.raises(RuntimeError, match="out of range"):
        model.eval()
        with torch.no_grad():
            model(torch.zeros(1, 600, dtype=torch.int64, device=DEV))
    with pytest.raises(NotImplementedError):
        model(torch.zeros(1, 8, dtype=torch.int64, device=DEV), head_mask=torch.ones(2))


def test_inputs_built_on_device_equal_collate_fun(pkg):
    """bqa_prepare_inputs (token types up to the first [SEP], mask = ids > 0, padding rules) against the
    host collate_fun on ragged rows, through Trainer._to_device's compact path."""
    ds = importlib.import_module("bqa_b200.modules.model.dataset")
    rng = np.random.default_rng(3)
    tok = types.SimpleNamespace(pad_token_id=0, sep_token_id=102, cls_token_id=101, model_name='bert')
    items = []
    for i in range(37):
        q = rng.integers(1, 60)
        d = rng.integers(1, 300)
        row = [101] + rng.integers(103, 30000, q).tolist() + [102] + rng.integers(103, 30000, d).tolist() + [102]
        items.append(ds.DatasetItem(example_id=str(i), input_ids=row, start_id=0, end_id=len(row) - 1, label_id=0,
                                    start_position=0.0, end_position=1.0))
    items.append(ds.DatasetItem(example_id='sep-first', input_ids=[102, 7, 8], start_id=0, end_id=1, label_id=0,
                                start_position=0.0, end_position=1.0))
    full, _ = ds.collate_fun(items, tok)
    compact, _ = ds.collate_fun(items, tok, compact=True, pin_memory=True)
    assert compact["no_padding"] is False
    got = ds.expand_compact_inputs(compact, torch.device(DEV))
    torch.cuda.synchronize()
    # a batch without padding elides the (all-True) mask
    same_len = [it for it in items if len(it.input_ids) == len(items[0].input_ids)] or items[:1]
    dense = ds.expand_compact_inputs(ds.collate_fun(same_len[:1] * 3, tok, compact=True)[0], torch.device(DEV))
    assert dense["attention_mask"] is None
    for k in ("input_ids", "token_type_ids", "attention_mask"):
        assert got[k].dtype == full[k].dtype, k
        assert torch.equal(got[k].cpu(), full[k]), k
    # RoBERTa rule: type_coef = 0 -> all token types 0
    tok_r = types.SimpleNamespace(pad_token_id=0, sep_token_id=102, cls_token_id=101, model_name='roberta')
    full_r, _ = ds.collate_fun(items, tok_r)
    got_r = ds.expand_compact_inputs(ds.collate_fun(items, tok_r, compact=True)[0], torch.device(DEV))
    assert torch.equal(got_r["token_type_ids"].cpu(), full_r["token_type_ids"])


def test_finetune_mode_trains_only_selected_modules(pkg):
    """params.finetune with finetune_class only (init.py:86-119, trainer.py:226-233): the model stays in eval
    mode except the classifier, so the encoder runs without dropout, the classifier's dropout fires, and only
    the classifier's parameters move."""
    I = importlib.import_module("bqa_b200.modules.init")
    T = importlib.import_module("bqa_b200.modules.model.trainer.trainer")
    model = _model(pkg, TINY, p_drop=0.5)
    hp = types.SimpleNamespace(lr=1e-2, weight_decay=0.0, optimizer="adam", finetune=True, finetune_transformer=False,
                               finetune_position=False, finetune_position_reg=False, finetune_class=True,
                               apex_level=None)
    opt = I.init_optimizer(hp, model)
    tr = T.Trainer(model=model, loss=_loss_obj(), collate_fun=None, optimizer=opt, device=torch.device(DEV), n_epochs=1,
                   train_batch_size=4, batch_split=1, n_jobs=0, warmup_coef=0.0, max_grad_norm=1.0)
    tr.set_train()
    assert not model.training and not model.transformer.training and model.classifier.training
    sd0 = {k: v.detach().clone() for k, v in model.state_dict().items()}
    inputs, labels = O.dummy_batch(4, 128, vocab=1000, seed=5, device=DEV)
    spans, clss = [], []
    for _ in range(2):
        outs, _ = model.forward_loss_backward(inputs, labels, tr.loss.loss_config(device=torch.device(DEV)))
        spans.append(outs["start_class"].clone())
        clss.append(outs["cls"].clone())
        model.engine.grads.zero_()
    assert torch.equal(spans[0], spans[1])        # encoder: no dropout although hidden_dropout_prob = 0.5
    assert not torch.equal(clss[0], clss[1])      # classifier dropout: a new mask every step
    for s in range(2):
        tr.train_micro_batch(inputs, labels, last_of_step=True)
        tr.optimizer_step()
    torch.cuda.synchronize()
    for k, v in model.state_dict().items():
        if k.startswith("classifier."):
            assert not torch.equal(v, sd0[k]), k
        else:
            assert torch.equal(v, sd0[k]), k
