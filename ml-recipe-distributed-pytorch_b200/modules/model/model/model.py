"""Drop-in for the reference's model class (modules/model/model/model.py:13-73).

Same constructor arguments, sub-module attribute names (`transformer`, `position_outputs`,
`classifier`, `reg_start`, `reg_end` — consumed by init.py:97-113), state_dict keys, forward
signature and output dictionary.  The sub-modules are parameter containers only: every FLOP of
`forward` and of its backward runs in libbqa (sm_100a kernels) through `engine.Engine`.

All parameters are views into one flat fp32 buffer owned by the engine (so that the gradient
all-reduce and the fused optimizer run over contiguous memory); `p.grad` of every parameter is a
view into the engine's flat gradient buffer, which the backward kernels accumulate into.  Plain
PyTorch optimizers and `clip_grad_norm_` therefore keep working on `model.parameters()`.
"""
import logging
import types

import torch
import torch.nn as nn

from .... import engine as _engine

logger = logging.getLogger(__name__)


class _Container(nn.Module):
    """A named bag of sub-modules / parameters; never called."""

    def forward(self, *a, **k):  # pragma: no cover
        raise RuntimeError("this module only holds parameters; the computation runs in libbqa")


def _linear(out_f, in_f):
    return nn.Linear(in_f, out_f)


def _bert_skeleton(shape):
    """Module tree with the parameter names of transformers' BertModel."""
    H, I = shape["hidden"], shape["intermediate"]
    t = _Container()
    t.embeddings = _Container()
    t.embeddings.word_embeddings = nn.Embedding(shape["vocab"], H)
    t.embeddings.position_embeddings = nn.Embedding(shape["max_pos"], H)
    t.embeddings.token_type_embeddings = nn.Embedding(shape["type_vocab"], H)
    t.embeddings.LayerNorm = nn.LayerNorm(H)
    t.encoder = _Container()
    layers = []
    for _ in range(shape["num_layers"]):
        lyr = _Container()
        lyr.attention = _Container()
        lyr.attention.self = _Container()
        lyr.attention.self.query = _linear(H, H)
        lyr.attention.self.key = _linear(H, H)
        lyr.attention.self.value = _linear(H, H)
        lyr.attention.output = _Container()
        lyr.attention.output.dense = _linear(H, H)
        lyr.attention.output.LayerNorm = nn.LayerNorm(H)
        lyr.intermediate = _Container()
        lyr.intermediate.dense = _linear(I, H)
        lyr.output = _Container()
        lyr.output.dense = _linear(H, I)
        lyr.output.LayerNorm = nn.LayerNorm(H)
        layers.append(lyr)
    t.encoder.layer = nn.ModuleList(layers)
    t.pooler = _Container()
    t.pooler.dense = _linear(H, H)
    # transformers' BertPreTrainedModel._init_weights: N(0, 0.02) weights, zero biases, LN 1/0
    for m in t.modules():
        if isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)
            nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Embedding):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)
        elif isinstance(m, nn.LayerNorm):
            nn.init.ones_(m.weight)
            nn.init.zeros_(m.bias)
    return t


class _StepFn(torch.autograd.Function):
    """Connects the engine's forward/backward to autograd so that `loss.backward()` of any
    torch loss built on the output dictionary drives the CUDA backward."""

    @staticmethod
    def forward(ctx, anchor, model, input_ids, token_type_ids, attention_mask):
        ctx.model = model
        outs = model._engine_forward(input_ids, token_type_ids, attention_mask, training=True)
        ctx.token = model._step_token
        return outs

    @staticmethod
    def backward(ctx, d_start, d_end, d_cls, d_reg_s, d_reg_e):
        model = ctx.model
        if ctx.token != model._step_token:
            raise RuntimeError("backward through a BertForQuestionAnswering forward that is no longer the most "
                               "recent one: the engine keeps the activations of one micro-batch only")
        model._engine.set_output_grads(d_start, d_end, d_cls, d_reg_s, d_reg_e)
        model._engine.backward()
        return None, None, None, None, None


class BertForQuestionAnswering(nn.Module):
    """BERT model for QA and classification tasks (reference model.py:13)."""

    def __init__(self, model_params, num_labels):
        super().__init__()
        self.model_params = model_params
        name = getattr(model_params, "model", "bert-base-uncased")
        shape = getattr(model_params, "model_shape", None) or _engine.MODEL_SHAPES.get(name)
        if shape is None:
            raise NotImplementedError(f"model {name!r}: only {sorted(_engine.MODEL_SHAPES)} have a B200 path")
        self._shape = dict(shape)
        self._num_labels = num_labels
        self._hidden_dropout = float(getattr(model_params, "hidden_dropout_prob", 0.1))
        self._attn_dropout = float(getattr(model_params, "attention_probs_dropout_prob", 0.1))
        self._ln_eps = float(getattr(model_params, "layer_norm_eps", 1e-12))

        self.transformer = _bert_skeleton(self._shape)
        self.transformer.config = types.SimpleNamespace(
            hidden_size=shape["hidden"], num_hidden_layers=shape["num_layers"], num_attention_heads=shape["heads"],
            intermediate_size=shape["intermediate"], vocab_size=shape["vocab"],
            max_position_embeddings=shape["max_pos"], type_vocab_size=shape["type_vocab"],
            pad_token_id=shape.get("pad_token_id", 0),
            hidden_dropout_prob=self._hidden_dropout, attention_probs_dropout_prob=self._attn_dropout,
            layer_norm_eps=self._ln_eps, num_labels=num_labels)
        H = shape["hidden"]
        self.position_outputs = nn.Linear(H, 2)  # start/end
        self.classifier = nn.Sequential(nn.Dropout(self._hidden_dropout), nn.Linear(H, num_labels))
        self.reg_start = nn.Sequential(nn.Linear(H, 1), nn.Sigmoid())
        self.reg_end = nn.Sequential(nn.Linear(H, 1), nn.Sigmoid())

        self._engine = None
        self._versions = None
        self._step_token = 0
        self._step = 0
        self._seed = int(torch.initial_seed() & 0x7FFFFFFFFFFFFFFF)
        self._anchor = None
        self.grads_dirty = True  # the flat gradient buffer may hold non-zero values

    # ---- engine plumbing -------------------------------------------------------------------
    def _param_device(self):
        return next(self.parameters()).device

    def _attach(self):
        """(Re)creates the engine on the parameters' CUDA device and moves every parameter into
        its flat buffer (parameter objects keep their identity, only `.data` is re-pointed)."""
        dev = self._param_device()
        if dev.type != "cuda":
            raise RuntimeError("BertForQuestionAnswering runs on sm_100a only: move the model to a CUDA device "
                               "(there is no CPU implementation of this path)")
        cfg = _engine.make_config(self._shape, num_labels=self._num_labels, ln_eps=self._ln_eps,
                                  hidden_dropout=self._hidden_dropout, attn_dropout=self._attn_dropout)
        with torch.cuda.device(dev):
            eng = _engine.Engine(cfg, dev)
        views, gviews = eng.param_views(), eng.grad_views()
        named = dict(self.named_parameters())
        if list(named.keys()) != list(views.keys()):
            raise RuntimeError("parameter names diverged from the engine layout")
        with torch.no_grad():
            for k, p in named.items():
                views[k].copy_(p.data)
                old_grad = p.grad
                p.data = views[k]
                if old_grad is not None:
                    gviews[k].copy_(old_grad)
                p.grad = gviews[k]
        self._engine = eng
        self._anchor = torch.zeros((), device=dev, requires_grad=True)
        self._versions = None
        self._sync_if_needed()

    def _engine_ready(self):
        eng = self._engine
        if eng is None:
            return False
        p = self.transformer.embeddings.word_embeddings.weight
        return p.device == eng.device and p.data_ptr() == eng.params.data_ptr()

    def _sync_if_needed(self):
        """Refresh the bf16 shadow weights when any parameter was modified outside libbqa
        (optimizer.step of a torch optimizer, load_state_dict, manual edits bump `_version`)."""
        v = sum(p._version for p in self.parameters())
        if v != self._versions:
            self._engine.sync_weights()
            self._versions = v

    def mark_weights_synced(self):
        """Called by the fused optimizer, which writes master and shadow together."""
        self._versions = sum(p._version for p in self.parameters())

    def restore_grad_views(self):
        """`optimizer.zero_grad(set_to_none=True)` detaches p.grad; re-point it at the flat buffer."""
        gviews = self._engine.grad_views()
        for k, p in self.named_parameters():
            if p.grad is None or p.grad.data_ptr() != gviews[k].data_ptr():
                p.grad = gviews[k]

    @property
    def engine(self):
        if not self._engine_ready():
            self._attach()
        return self._engine

    def _engine_forward(self, input_ids, token_type_ids, attention_mask, training):
        if not self._engine_ready():
            self._attach()
        self._sync_if_needed()
        if training:
            self._step += 1
            self._step_token += 1
            self.grads_dirty = True
        if training:
            # bqa_forward's training bit set: the reference's dropouts belong to self.transformer and to
            # self.classifier[0]; they fire only if their owner is in train mode
            training = 1 | (0 if self.transformer.training else 2) | (0 if self.classifier.training else 4)
        return self._engine.forward(input_ids, token_type_ids, attention_mask, training=training,
                                    seed=self._seed, step=self._step)

    # ---- reference surface -----------------------------------------------------------------
    def forward(self, input_ids, attention_mask=None, token_type_ids=None, position_ids=None, head_mask=None):
        if position_ids is not None or head_mask is not None:
            raise NotImplementedError("position_ids / head_mask are always None in the reference's call sites "
                                      "(trainer.py:278, predictor.py:109) and have no kernel path")
        # The reference's graph exists whenever autograd is on; which dropouts fire depends on the train flag
        # of the sub-module that owns them (fine-tuning keeps the top-level module in eval mode, init.py:92).
        needs_grad = torch.is_grad_enabled() and (self.training or self.transformer.training or
                                                  self.classifier.training or self.position_outputs.training or
                                                  self.reg_start.training or self.reg_end.training)
        if needs_grad:
            if not self._engine_ready():
                self._attach()
            self.restore_grad_views()
            outs = _StepFn.apply(self._anchor, self, input_ids, token_type_ids, attention_mask)
        else:
            outs = self._engine_forward(input_ids, token_type_ids, attention_mask, training=False)
        start, end, cls, reg_s, reg_e = outs
        return {'start_class': start, 'end_class': end, 'start_reg': reg_s, 'end_reg': reg_e, 'cls': cls}

    # ---- fused fast path (used by the package's Trainer) --------------------------------------
    def forward_loss_backward(self, inputs, labels, loss_config, *, last_phase=None):
        """One micro-batch without autograd: forward, fused loss (6 scalars on the device) and the
        backward up to `last_phase` (None = all).  Returns (outputs dict, losses tensor[6])."""
        outs = self._engine_forward(inputs['input_ids'], inputs.get('token_type_ids'), inputs.get('attention_mask'),
                                    training=True)
        losses = self._engine.loss(outs, labels, loss_config, with_grad=True)
        if last_phase is None or last_phase >= 0:
            self._engine.backward(0, last_phase)
        start, end, cls, reg_s, reg_e = outs
        return ({'start_class': start, 'end_class': end, 'start_reg': reg_s, 'end_reg': reg_e, 'cls': cls}, losses)
