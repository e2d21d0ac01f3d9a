"""CPU oracle: plain-PyTorch fp32 restatement of the reference's BERT-QA step.  TEST INFRASTRUCTURE ONLY.

Nothing under ml-recipe-distributed-pytorch_b200/ imports this file.  It restates, as
straight-line functional code over a ``{state_dict key: tensor}`` dictionary:

  forward      reference modules/model/model/model.py:43-73 on top of transformers' BertModel
               (site-packages/transformers/models/bert/modeling_bert.py:72-112 embeddings,
               :115-140,168-207 attention, :294-298 self-output, :339-342 intermediate,
               :352-356 output, :462-468 pooler)
  loss         reference modules/model/model/loss.py:5-38,74-100 with the loss set of
               modules/init.py:18-40
  backward     torch autograd over the above (the reference does the same, trainer.py:197-204)
  clip         torch.nn.utils.clip_grad_norm_ as called in trainer.py:221-225
  AdamW        transformers 3.5.1 optimization.AdamW(correct_bias=False) as constructed in
               init.py:125-138 — that class no longer exists in the installed transformers,
               so this part is a restatement of the published algorithm
  schedule     get_linear_schedule_with_warmup (trainer.py:116-126)
  decode       reference modules/model/inference/predictor.py:114-120

PARITY PINS: tests/golden/*.npz hold inputs/outputs produced by the reference's own classes
(oracle/make_golden.py imports /root/reference and the installed transformers BertModel);
tests/test_oracle_golden.py checks this file against them.  The AdamW restatement has no
reference-side pin (upstream has no tests and the class is gone): "parity unpinned" for that
one function.

Two extras that the reference does not have, both needed to test the CUDA path exactly:
  * dropout masks come from oracle/dropout_hash.py (the library's stateless rule) instead of
    torch's Philox stream;
  * ``bf16=True`` rounds weights and every stored activation to bfloat16 at the same points
    as the kernels do, which turns a loose fp32-vs-bf16 tolerance into a tight one.
"""
import math
from dataclasses import dataclass
from typing import Dict, Optional

import torch
import torch.nn.functional as F

from .dropout_hash import _M64, keep_mask, splitmix64


@dataclass
class OracleConfig:
    num_layers: int = 12
    hidden: int = 768
    heads: int = 12
    intermediate: int = 3072
    vocab: int = 30522
    max_pos: int = 512
    type_vocab: int = 2
    num_labels: int = 5
    ln_eps: float = 1e-12
    hidden_dropout: float = 0.1
    attn_dropout: float = 0.1
    # embedding conventions of the encoder family: BERT (0, 0) / RoBERTa (1, 1), see forward()
    pad_token_id: int = 0
    position_mode: int = 0


@dataclass
class DropSpec:
    """Selects the dropout streams of one training step (seed, step) as the library does."""
    seed: int
    step: int

    def stream_seed(self):
        return (self.seed ^ splitmix64((self.step + 0x51ED270B9F3A6C1D) & _M64)) & _M64


def param_names(cfg: OracleConfig):
    """state_dict keys of the reference model, in named_parameters() order."""
    n = ["transformer.embeddings.word_embeddings.weight",
         "transformer.embeddings.position_embeddings.weight",
         "transformer.embeddings.token_type_embeddings.weight",
         "transformer.embeddings.LayerNorm.weight",
         "transformer.embeddings.LayerNorm.bias"]
    for l in range(cfg.num_layers):
        p = f"transformer.encoder.layer.{l}."
        for m in ("attention.self.query", "attention.self.key", "attention.self.value", "attention.output.dense",
                  "attention.output.LayerNorm", "intermediate.dense", "output.dense", "output.LayerNorm"):
            n += [p + m + ".weight", p + m + ".bias"]
    n += ["transformer.pooler.dense.weight", "transformer.pooler.dense.bias",
          "position_outputs.weight", "position_outputs.bias",
          "classifier.1.weight", "classifier.1.bias",
          "reg_start.0.weight", "reg_start.0.bias",
          "reg_end.0.weight", "reg_end.0.bias"]
    return n


def param_shapes(cfg: OracleConfig):
    H, I = cfg.hidden, cfg.intermediate
    s = {"transformer.embeddings.word_embeddings.weight": (cfg.vocab, H),
         "transformer.embeddings.position_embeddings.weight": (cfg.max_pos, H),
         "transformer.embeddings.token_type_embeddings.weight": (cfg.type_vocab, H),
         "transformer.embeddings.LayerNorm.weight": (H,),
         "transformer.embeddings.LayerNorm.bias": (H,)}
    for l in range(cfg.num_layers):
        p = f"transformer.encoder.layer.{l}."
        for m in ("attention.self.query", "attention.self.key", "attention.self.value", "attention.output.dense"):
            s[p + m + ".weight"], s[p + m + ".bias"] = (H, H), (H,)
        for m in ("attention.output.LayerNorm", "output.LayerNorm"):
            s[p + m + ".weight"], s[p + m + ".bias"] = (H,), (H,)
        s[p + "intermediate.dense.weight"], s[p + "intermediate.dense.bias"] = (I, H), (I,)
        s[p + "output.dense.weight"], s[p + "output.dense.bias"] = (H, I), (H,)
    s["transformer.pooler.dense.weight"], s["transformer.pooler.dense.bias"] = (H, H), (H,)
    s["position_outputs.weight"], s["position_outputs.bias"] = (2, H), (2,)
    s["classifier.1.weight"], s["classifier.1.bias"] = (cfg.num_labels, H), (cfg.num_labels,)
    s["reg_start.0.weight"], s["reg_start.0.bias"] = (1, H), (1,)
    s["reg_end.0.weight"], s["reg_end.0.bias"] = (1, H), (1,)
    return s


def init_params(cfg: OracleConfig, seed: int, device="cpu") -> Dict[str, torch.Tensor]:
    """Random init with the reference's distributions: HF BERT init for the encoder
    (N(0, 0.02) weights, zero biases, LayerNorm 1/0; transformers modeling_utils._init_weights),
    torch's default nn.Linear init for the four heads (model.py:30-41).  Values differ from what
    torch.manual_seed(seed) + the reference constructor would draw; only the law matters."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for name, shape in param_shapes(cfg).items():
        head = not name.startswith("transformer.")
        if name.endswith("LayerNorm.weight"):
            t = torch.ones(shape)
        elif name.endswith("LayerNorm.bias"):
            t = torch.zeros(shape)
        elif head:
            fan_in = cfg.hidden
            bound = 1.0 / math.sqrt(fan_in)  # kaiming_uniform(a=sqrt(5)) on weight, same bound on bias
            t = (torch.rand(shape, generator=g) * 2 - 1) * bound
        elif name.endswith(".bias"):
            t = torch.zeros(shape)
        else:
            t = torch.randn(shape, generator=g) * 0.02
        out[name] = t.to(device)
    return out


def _q(x, bf16):
    return x.to(torch.bfloat16).to(torch.float32) if bf16 else x


def _drop(x, p, spec, site, flat_index_shape=None):
    """x * keep / (1 - p) with the library's decision rule; element index = C-order position.
    spec == "torch" uses torch's own dropout stream instead (what the reference does; used only
    when the oracle is TIMED as the CPU baseline, never for parity)."""
    if spec is None or p <= 0:
        return x
    if isinstance(spec, str):
        return F.dropout(x, p, training=True)
    keep = keep_mask(spec.stream_seed(), site, p, x.numel(), device=x.device).view(x.shape)
    return torch.where(keep, x * (1.0 / (1.0 - p)), torch.zeros_like(x))


def _layer_norm(x, w, b, eps):
    return F.layer_norm(x, (x.shape[-1],), w, b, eps)


def forward(P: Dict[str, torch.Tensor], cfg: OracleConfig, input_ids, attention_mask=None, token_type_ids=None,
            drop: Optional[DropSpec] = None, bf16: bool = False, return_hidden: bool = False):
    """Returns the reference's output dict {start_class, end_class, start_reg, end_reg, cls}.

    drop=None is eval mode (model.eval()).  bf16=True mirrors the kernels' rounding points."""
    B, S = input_ids.shape
    H, A, L = cfg.hidden, cfg.heads, cfg.num_layers
    d = H // A
    q = lambda t: _q(t, bf16)                       # noqa: E731  activation rounding
    wq = lambda name: _q(P[name], bf16)             # noqa: E731  GEMM weights live in bf16
    if token_type_ids is None:
        token_type_ids = torch.zeros_like(input_ids)
    e = "transformer.embeddings."
    pad = cfg.pad_token_id
    if cfg.position_mode == 0:
        # BertEmbeddings (modeling_bert.py:72-112): position id = index in the sequence
        pos_emb = P[e + "position_embeddings.weight"][torch.arange(S, device=input_ids.device)][None]
    else:
        # RobertaEmbeddings (modeling_roberta.py: create_position_ids_from_input_ids): padding_idx + number of
        # non-padding tokens up to and including this one, padding_idx for padding tokens; the position table has
        # padding_idx = pad_token_id as well (that row receives no gradient)
        nonpad = (input_ids != pad).long()
        pos_ids = torch.cumsum(nonpad, dim=1) * nonpad + pad
        pos_emb = F.embedding(pos_ids, P[e + "position_embeddings.weight"], padding_idx=pad)
    # nn.Embedding(vocab, H, padding_idx=pad_token_id) in Bert/RobertaEmbeddings: that row receives no gradient
    z0 = (F.embedding(input_ids, P[e + "word_embeddings.weight"], padding_idx=pad)
          + P[e + "token_type_embeddings.weight"][token_type_ids]) \
        + pos_emb
    z0 = q(z0)
    x = _layer_norm(z0, P[e + "LayerNorm.weight"], P[e + "LayerNorm.bias"], cfg.ln_eps)
    x = q(_drop(x, cfg.hidden_dropout, drop, 0))
    if attention_mask is not None:
        kbias = torch.zeros(B, 1, 1, S, device=x.device).masked_fill(~attention_mask.bool()[:, None, None, :],
                                                                      float("-inf"))
    else:
        kbias = None
    hidden = [x]
    for l in range(L):
        p = f"transformer.encoder.layer.{l}."
        wqkv = torch.cat([wq(p + "attention.self.query.weight"), wq(p + "attention.self.key.weight"),
                          wq(p + "attention.self.value.weight")], 0)
        bqkv = torch.cat([P[p + "attention.self.query.bias"], P[p + "attention.self.key.bias"],
                          P[p + "attention.self.value.bias"]], 0)
        qkv = q(x @ wqkv.t() + bqkv)                                   # [B,S,3H]
        qh, kh, vh = [t.view(B, S, A, d).transpose(1, 2) for t in qkv.split(H, dim=-1)]
        scores = (qh @ kh.transpose(-1, -2)) * (1.0 / math.sqrt(d))
        if kbias is not None:
            scores = scores + kbias
        probs = torch.softmax(scores, dim=-1)
        probs = _drop(probs, cfg.attn_dropout, drop, 1 + 3 * l)
        probs = q(probs)                                               # P is fed to the PV MMA as bf16
        ctx = q((probs @ vh).transpose(1, 2).reshape(B, S, H))
        ao = ctx @ wq(p + "attention.output.dense.weight").t() + P[p + "attention.output.dense.bias"]
        z1 = q(_drop(ao, cfg.hidden_dropout, drop, 2 + 3 * l) + x)
        a = q(_layer_norm(z1, P[p + "attention.output.LayerNorm.weight"], P[p + "attention.output.LayerNorm.bias"],
                          cfg.ln_eps))
        u_f = a @ wq(p + "intermediate.dense.weight").t() + P[p + "intermediate.dense.bias"]
        g = q(F.gelu(u_f))                                             # erf GELU on the fp32 pre-activation
        o = g @ wq(p + "output.dense.weight").t() + P[p + "output.dense.bias"]
        z2 = q(_drop(o, cfg.hidden_dropout, drop, 3 + 3 * l) + a)
        x = q(_layer_norm(z2, P[p + "output.LayerNorm.weight"], P[p + "output.LayerNorm.bias"], cfg.ln_eps))
        hidden.append(x)
    pooled = torch.tanh(x[:, 0] @ P["transformer.pooler.dense.weight"].t() + P["transformer.pooler.dense.bias"])
    pos_logits = x @ P["position_outputs.weight"].t() + P["position_outputs.bias"]
    start, end = pos_logits[..., 0], pos_logits[..., 1]
    cls = _drop(pooled, cfg.hidden_dropout, drop, 1 + 3 * L) @ P["classifier.1.weight"].t() + P["classifier.1.bias"]
    reg_s = torch.sigmoid(pooled @ P["reg_start.0.weight"].t() + P["reg_start.0.bias"]).squeeze(-1)
    reg_e = torch.sigmoid(pooled @ P["reg_end.0.weight"].t() + P["reg_end.0.bias"]).squeeze(-1)
    out = {"start_class": start, "end_class": end, "start_reg": reg_s, "end_reg": reg_e, "cls": cls}
    if return_hidden:
        out["_hidden"] = hidden
    return out


@dataclass
class LossConfig:
    w_start: float = 1.0
    w_end: float = 1.0
    w_start_reg: float = 1.0
    w_end_reg: float = 1.0
    w_cls: float = 1.0
    loss: str = "smooth"          # 'smooth' | 'ce' | 'focal'
    smooth_alpha: float = 0.01
    label_weights: Optional[torch.Tensor] = None
    focal_alpha: float = 1.0
    focal_gamma: float = 2.0


def class_loss(logits, target, lc: LossConfig):
    """init.py:22-31.  'smooth' = LabelSmoothingLossWithLogits (loss.py:5-38)."""
    C = logits.shape[-1]
    logp = torch.log_softmax(logits, dim=-1)
    if lc.loss == "smooth":
        if lc.smooth_alpha > 0:
            fill = lc.smooth_alpha / (C - 1)           # ignore_index=-100 is outside [0, C): num_ignore_ixs = 1
            t = torch.full_like(logp, fill)
            t.scatter_(-1, target.unsqueeze(-1), 1.0 - lc.smooth_alpha)
            return F.kl_div(logp, t, reduction="batchmean")
        return F.nll_loss(logp, target, ignore_index=-100)
    if lc.loss == "ce":
        return F.cross_entropy(logits, target, weight=lc.label_weights)
    if lc.loss == "focal":
        # FocalLossWithLogits (loss.py:57-71): NLLLoss(ignore_index=-1) of alpha (1 - p)^gamma log p
        probs = torch.exp(logp)
        return F.nll_loss(lc.focal_alpha * (1 - probs) ** lc.focal_gamma * logp, target, ignore_index=-1)
    raise NotImplementedError(lc.loss)


def weighted_loss(out, labels, lc: LossConfig):
    """WeightedLoss.__call__ (loss.py:78-100): returns (total, {key: unweighted loss})."""
    parts = {
        "start_class": F.cross_entropy(out["start_class"], labels["start_class"], ignore_index=-1),
        "end_class": F.cross_entropy(out["end_class"], labels["end_class"], ignore_index=-1),
        "start_reg": F.mse_loss(out["start_reg"], labels["start_reg"]),
        "end_reg": F.mse_loss(out["end_reg"], labels["end_reg"]),
        "cls": class_loss(out["cls"], labels["cls"], lc),
    }
    w = {"start_class": lc.w_start, "end_class": lc.w_end, "start_reg": lc.w_start_reg, "end_reg": lc.w_end_reg,
         "cls": lc.w_cls}
    total = 0
    for k in ("start_class", "end_class", "start_reg", "end_reg", "cls"):
        total = total + w[k] * parts[k]
    return total, parts


def loss_and_grads(P, cfg, inputs, labels, lc: LossConfig, drop=None, bf16=False, loss_scale=1.0):
    """One forward + backward; returns (total, parts, grads dict) with grads of total * loss_scale."""
    Pg = {k: v.detach().clone().requires_grad_(True) for k, v in P.items()}
    out = forward(Pg, cfg, inputs["input_ids"], inputs.get("attention_mask"), inputs.get("token_type_ids"),
                  drop=drop, bf16=bf16)
    total, parts = weighted_loss(out, labels, lc)
    (total * loss_scale).backward()
    grads = {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in Pg.items()}
    return total.detach(), {k: v.detach() for k, v in parts.items()}, grads, {k: v.detach() for k, v in out.items()}


def clip_grad_norm_(grads: Dict[str, torch.Tensor], max_norm: float):
    """torch.nn.utils.clip_grad_norm_ (trainer.py:221-225): returns the total norm, scales in place."""
    total = torch.sqrt(sum((g.double() ** 2).sum() for g in grads.values())).float()
    coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
    for g in grads.values():
        g.mul_(coef)
    return total


NO_DECAY = ("bias", "LayerNorm.bias", "LayerNorm.weight")   # init.py:121


def hf_adamw_step(P, grads, state, lr, weight_decay, betas=(0.9, 0.999), eps=1e-6):
    """transformers 3.5.1 AdamW.step with correct_bias=False and the two groups of init.py:125-129.
    PARITY UNPINNED (see module docstring)."""
    b1, b2 = betas
    for name, p in P.items():
        g = grads[name]
        st = state.setdefault(name, {"step": 0, "exp_avg": torch.zeros_like(p), "exp_avg_sq": torch.zeros_like(p)})
        st["step"] += 1
        st["exp_avg"].mul_(b1).add_(g, alpha=1.0 - b1)
        st["exp_avg_sq"].mul_(b2).addcmul_(g, g, value=1.0 - b2)
        denom = st["exp_avg_sq"].sqrt().add_(eps)
        p.addcdiv_(st["exp_avg"], denom, value=-lr)
        wd = 0.0 if any(nd in name for nd in NO_DECAY) else weight_decay
        if wd > 0.0:
            p.add_(p, alpha=-lr * wd)


def adamod_step(P, grads, state, lr, weight_decay, betas=(0.9, 0.999), beta3=0.999, eps=1e-8):
    """AdaMod.step (modules/model/trainer/optim.py:42-100) with the two groups of init.py:125-129.
    Pinned by tests/golden/ref_adamod.npz (the reference class itself run for several steps)."""
    b1, b2 = betas
    for name, p in P.items():
        g = grads[name]
        st = state.setdefault(name, {"step": 0, "exp_avg": torch.zeros_like(p), "exp_avg_sq": torch.zeros_like(p),
                                     "exp_avg_lr": torch.zeros_like(p)})
        st["step"] += 1
        st["exp_avg"].mul_(b1).add_(g, alpha=1.0 - b1)
        st["exp_avg_sq"].mul_(b2).addcmul_(g, g, value=1.0 - b2)
        denom = st["exp_avg_sq"].sqrt().add_(eps)
        step_size = lr * math.sqrt(1.0 - b2 ** st["step"]) / (1.0 - b1 ** st["step"])
        wd = 0.0 if any(nd in name for nd in NO_DECAY) else weight_decay
        if wd != 0.0:
            p.add_(p, alpha=-wd * lr)
        eta = torch.full_like(denom, step_size).div_(denom)
        st["exp_avg_lr"].mul_(beta3).add_(eta, alpha=1.0 - beta3)
        eta = torch.min(eta, st["exp_avg_lr"])
        p.add_(-eta * st["exp_avg"])


def linear_schedule_lambda(step, num_warmup_steps, num_training_steps):
    """get_linear_schedule_with_warmup's lr_lambda (transformers/optimization.py)."""
    if step < num_warmup_steps:
        return float(step) / float(max(1, num_warmup_steps))
    return max(0.0, float(num_training_steps - step) / float(max(1, num_training_steps - num_warmup_steps)))


def decode(out):
    """predictor.py:114-120."""
    start_logits, start_ids = torch.max(out["start_class"], dim=-1)
    end_logits, end_ids = torch.max(out["end_class"], dim=-1)
    cls_probas, cls_ids = torch.max(torch.softmax(out["cls"], dim=-1), dim=-1)
    scores = start_logits + end_logits - (out["start_class"][:, 0] + out["end_class"][:, 0])
    return {"start_ids": start_ids, "end_ids": end_ids, "cls_ids": cls_ids, "scores": scores,
            "start_logits": start_logits, "end_logits": end_logits, "cls_probas": cls_probas}


def dummy_batch(B, S, *, vocab=30522, max_question_len=64, seed=1234, device="cpu"):
    """DummyDataset.__getitem__ + collate_fun for B items (dummy_dataset.py:36-51,
    split_dataset.py:480-520): [CLS] q(64) [SEP] doc [SEP], never padded, fixed labels."""
    import numpy as np
    rng = np.random.RandomState(seed)
    pad, unk, cls, sep = 0, 100, 101, 102
    doc_len = S - max_question_len - 3
    rows = []
    for _ in range(B):
        qd = rng.randint(1, vocab, max_question_len)
        dd = rng.randint(1, vocab, doc_len)
        for a in (qd, dd):
            for w in (pad, sep, cls):
                a[a == w] = unk
        rows.append([cls] + qd.tolist() + [sep] + dd.tolist() + [sep])
    ids = torch.tensor(rows, dtype=torch.int64)
    first_sep = max_question_len + 1
    tt = (torch.arange(S)[None, :] > first_sep).to(torch.int64).expand(B, S).contiguous()
    inputs = {"input_ids": ids.to(device), "attention_mask": (ids > 0).to(device), "token_type_ids": tt.to(device)}
    labels = {"start_class": torch.zeros(B, dtype=torch.int64, device=device),
              "end_class": torch.full((B,), S - 1, dtype=torch.int64, device=device),
              "start_reg": torch.zeros(B, device=device),
              "end_reg": torch.ones(B, device=device),
              "cls": torch.zeros(B, dtype=torch.int64, device=device)}
    return inputs, labels
