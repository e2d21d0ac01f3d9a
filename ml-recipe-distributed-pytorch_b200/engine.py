"""Host-side owner of one libbqa handle: flat parameter / gradient / bf16-shadow buffers, the
activation workspace, and thin methods over the C ABI (include/bqa.h).

PyTorch is used here for device memory and streams only; every kernel is in libbqa.so.
"""
import ctypes
from collections import OrderedDict

import torch

from . import _lib as L

# model name (reference parser.py:66-68 `--model`) -> encoder shape.  bert-large-uncased is the
# documented extension needed by BASELINE config 4 (the reference's choice list stops at base).
MODEL_SHAPES = {
    "bert-base-uncased": dict(num_layers=12, hidden=768, heads=12, intermediate=3072, vocab=30522, max_pos=512,
                              type_vocab=2),
    "bert-large-uncased": dict(num_layers=24, hidden=1024, heads=16, intermediate=4096, vocab=30522, max_pos=512,
                               type_vocab=2),
    # the reference's second choice (model.py:9-10 -> RobertaModel): one token type, <pad> = 1, position ids counted
    # over the non-padding tokens and offset by the padding index (514 = 512 + 2 rows)
    "roberta-base": dict(num_layers=12, hidden=768, heads=12, intermediate=3072, vocab=50265, max_pos=514,
                         type_vocab=1, pad_token_id=1, position_mode=1),
}


def make_config(shape, *, num_labels=5, ln_eps=1e-12, hidden_dropout=0.1, attn_dropout=0.1):
    c = L.Config()
    c.num_layers, c.hidden, c.heads, c.intermediate = (shape["num_layers"], shape["hidden"], shape["heads"],
                                                       shape["intermediate"])
    c.vocab, c.max_pos, c.type_vocab, c.num_labels = shape["vocab"], shape["max_pos"], shape["type_vocab"], num_labels
    c.ln_eps, c.hidden_dropout, c.attn_dropout = ln_eps, hidden_dropout, attn_dropout
    c.pad_token_id, c.position_mode = int(shape.get("pad_token_id", 0)), int(shape.get("position_mode", 0))
    return c


def param_layout(cfg):
    """Flat offsets of every tensor under the reference's state_dict key:
    OrderedDict key -> (offset, shape), in the reference's named_parameters() order."""
    lay = L.Layout()
    L.check(L.lib.bqa_param_layout(ctypes.byref(cfg), ctypes.byref(lay)), "bqa_param_layout")
    H, I, C = cfg.hidden, cfg.intermediate, cfg.num_labels
    out = OrderedDict()
    e = "transformer.embeddings."
    out[e + "word_embeddings.weight"] = (lay.word_emb, (cfg.vocab, H))
    out[e + "position_embeddings.weight"] = (lay.pos_emb, (cfg.max_pos, H))
    out[e + "token_type_embeddings.weight"] = (lay.type_emb, (cfg.type_vocab, H))
    out[e + "LayerNorm.weight"] = (lay.emb_ln_g, (H,))
    out[e + "LayerNorm.bias"] = (lay.emb_ln_b, (H,))
    for l in range(cfg.num_layers):
        b = lay.layer0 + l * lay.layer_stride
        p = f"transformer.encoder.layer.{l}."
        for i, n in enumerate(("query", "key", "value")):
            out[p + f"attention.self.{n}.weight"] = (b + lay.qkv_w + i * H * H, (H, H))
            out[p + f"attention.self.{n}.bias"] = (b + lay.qkv_b + i * H, (H,))
        out[p + "attention.output.dense.weight"] = (b + lay.ao_w, (H, H))
        out[p + "attention.output.dense.bias"] = (b + lay.ao_b, (H,))
        out[p + "attention.output.LayerNorm.weight"] = (b + lay.ln1_g, (H,))
        out[p + "attention.output.LayerNorm.bias"] = (b + lay.ln1_b, (H,))
        out[p + "intermediate.dense.weight"] = (b + lay.i_w, (I, H))
        out[p + "intermediate.dense.bias"] = (b + lay.i_b, (I,))
        out[p + "output.dense.weight"] = (b + lay.o_w, (H, I))
        out[p + "output.dense.bias"] = (b + lay.o_b, (H,))
        out[p + "output.LayerNorm.weight"] = (b + lay.ln2_g, (H,))
        out[p + "output.LayerNorm.bias"] = (b + lay.ln2_b, (H,))
    out["transformer.pooler.dense.weight"] = (lay.pool_w, (H, H))
    out["transformer.pooler.dense.bias"] = (lay.pool_b, (H,))
    out["position_outputs.weight"] = (lay.span_w, (2, H))
    out["position_outputs.bias"] = (lay.span_b, (2,))
    out["classifier.1.weight"] = (lay.cls_w, (C, H))
    out["classifier.1.bias"] = (lay.cls_b, (C,))
    out["reg_start.0.weight"] = (lay.regs_w, (1, H))
    out["reg_start.0.bias"] = (lay.regs_b, (1,))
    out["reg_end.0.weight"] = (lay.rege_w, (1, H))
    out["reg_end.0.bias"] = (lay.rege_b, (1,))
    return out, int(lay.total)


def _numel(shape):
    n = 1
    for s in shape:
        n *= s
    return n


class Engine:
    """One model replica on one GPU."""

    def __init__(self, cfg, device, *, with_grads=True):
        if not torch.cuda.is_available():
            raise RuntimeError("bqa Engine needs a CUDA device (sm_100a); there is no CPU implementation")
        self.cfg = cfg
        self.device = torch.device(device)
        self.layout, self.total = param_layout(cfg)
        with torch.cuda.device(self.device):
            self.params = torch.zeros(self.total, dtype=torch.float32, device=self.device)
            self.grads = torch.zeros(self.total, dtype=torch.float32, device=self.device) if with_grads else None
            self.shadow = torch.zeros(self.total, dtype=torch.bfloat16, device=self.device)
            h = ctypes.c_void_p()
            L.check(L.lib.bqa_create(ctypes.byref(cfg), ctypes.byref(h)), "bqa_create")
        self.handle = h
        self.workspace = None
        self._ws_key = None
        self.num_phases = L.lib.bqa_num_backward_phases(self.handle)
        self._losses = torch.zeros(6, dtype=torch.float32, device=self.device)
        self._bind(1 << 20)

    def __del__(self):
        h = getattr(self, "handle", None)
        if h is not None and L is not None:
            L.lib.bqa_destroy(h)
            self.handle = None

    # ---- buffers -------------------------------------------------------------------------
    def view(self, flat, key):
        off, shape = self.layout[key]
        return flat[off:off + _numel(shape)].view(shape)

    def param_views(self):
        return OrderedDict((k, self.view(self.params, k)) for k in self.layout)

    def grad_views(self):
        return OrderedDict((k, self.view(self.grads, k)) for k in self.layout)

    def _bind(self, ws_bytes):
        self.workspace = torch.empty(ws_bytes, dtype=torch.uint8, device=self.device)
        L.check(L.lib.bqa_bind(self.handle, self.params.data_ptr(),
                               self.grads.data_ptr() if self.grads is not None else None, self.shadow.data_ptr(),
                               self.workspace.data_ptr(), ws_bytes), "bqa_bind")

    def _ensure_workspace(self, B, S, training):
        need = ctypes.c_uint64()
        L.check(L.lib.bqa_workspace_bytes(ctypes.byref(self.cfg), B, S, int(training), ctypes.byref(need)),
                "bqa_workspace_bytes")
        if self.workspace is None or self.workspace.numel() < need.value:
            torch.cuda.synchronize(self.device)
            self.workspace = None
            self._bind(int(need.value))

    def sync_weights(self):
        """fp32 master -> bf16 shadow (after load_state_dict or any out-of-library update)."""
        L.check(L.lib.bqa_f32_to_bf16(self.params.data_ptr(), self.shadow.data_ptr(), self.total,
                                      L.current_stream_ptr()), "bqa_f32_to_bf16")

    # ---- step ----------------------------------------------------------------------------
    def forward(self, input_ids, token_type_ids=None, attention_mask=None, *, training=False, seed=0, step=0):
        B, S = input_ids.shape
        self._ensure_workspace(B, S, training)
        dev = self.device
        ids = input_ids.to(device=dev, dtype=torch.int64).contiguous()
        tt = None if token_type_ids is None else token_type_ids.to(device=dev, dtype=torch.int64).contiguous()
        mask = None
        if attention_mask is not None:
            mask = attention_mask.to(device=dev)
            mask = (mask if mask.dtype == torch.bool else mask != 0).to(torch.uint8).contiguous()
        C = self.cfg.num_labels
        start = torch.empty(B, S, dtype=torch.float32, device=dev)
        end = torch.empty(B, S, dtype=torch.float32, device=dev)
        cls = torch.empty(B, C, dtype=torch.float32, device=dev)
        reg_s = torch.empty(B, dtype=torch.float32, device=dev)
        reg_e = torch.empty(B, dtype=torch.float32, device=dev)
        L.check(L.lib.bqa_forward(self.handle, ids.data_ptr(), L.ptr(tt), L.ptr(mask), B, S, int(training),
                                  seed & 0xFFFFFFFFFFFFFFFF, step, start.data_ptr(), end.data_ptr(), cls.data_ptr(),
                                  reg_s.data_ptr(), reg_e.data_ptr(), L.current_stream_ptr()), "bqa_forward")
        return start, end, cls, reg_s, reg_e

    def loss(self, outs, labels, lc, *, with_grad, out=None):
        """outs: (start, end, cls, reg_s, reg_e) from forward(); labels: dict as built by
        collate_fun; lc: _lib.LossConfig.  Returns the device tensor of the 6 loss scalars."""
        start, end, cls, reg_s, reg_e = outs
        dev = self.device
        ys = labels["start_class"].to(device=dev, dtype=torch.int64).contiguous()
        ye = labels["end_class"].to(device=dev, dtype=torch.int64).contiguous()
        yc = labels["cls"].to(device=dev, dtype=torch.int64).contiguous()
        yrs = labels["start_reg"].to(device=dev, dtype=torch.float32).contiguous()
        yre = labels["end_reg"].to(device=dev, dtype=torch.float32).contiguous()
        losses = self._losses if out is None else out
        L.check(L.lib.bqa_loss(self.handle, start.data_ptr(), end.data_ptr(), cls.data_ptr(), reg_s.data_ptr(),
                               reg_e.data_ptr(), ys.data_ptr(), ye.data_ptr(), yc.data_ptr(), yrs.data_ptr(),
                               yre.data_ptr(), ctypes.byref(lc), int(with_grad), losses.data_ptr(),
                               L.current_stream_ptr()), "bqa_loss")
        return losses

    def check_inputs(self):
        """Raises IndexError if the last forward saw a token / token-type id outside the embedding tables (what
        torch's nn.Embedding does in the reference).  Synchronises the current stream."""
        flag = L.c_i32(0)
        L.check(L.lib.bqa_input_status(self.handle, ctypes.byref(flag), L.current_stream_ptr()), "bqa_input_status")
        if flag.value:
            raise IndexError("index out of range in self (input_ids / token_type_ids outside the embedding tables)")

    def set_output_grads(self, d_start, d_end, d_cls, d_reg_s, d_reg_e):
        def c(t):
            return None if t is None else t.to(device=self.device, dtype=torch.float32).contiguous()
        ts = [c(t) for t in (d_start, d_end, d_cls, d_reg_s, d_reg_e)]
        L.check(L.lib.bqa_set_output_grads(self.handle, *[L.ptr(t) for t in ts], L.current_stream_ptr()),
                "bqa_set_output_grads")

    def backward(self, first_phase=0, last_phase=None):
        if last_phase is None:
            last_phase = self.num_phases - 1
        L.check(L.lib.bqa_backward(self.handle, first_phase, last_phase, L.current_stream_ptr()), "bqa_backward")

    def phase_grad_begin(self, phase):
        return int(L.lib.bqa_phase_grad_begin(self.handle, phase))

    def profile_begin(self):
        L.check(L.lib.bqa_profile_begin(self.handle), "bqa_profile_begin")

    def profile_end(self):
        """-> {kind: (ms, work, launches)} for the kernels launched since profile_begin; work = matmul FLOPs for
        the GEMM / attention kinds, algorithmic HBM bytes for the row kernels (include/bqa.h)."""
        K = 9  # BQA_PROFILE_KINDS
        ms, fl, n = (ctypes.c_double * K)(), (ctypes.c_double * K)(), (L.c_i64 * K)()
        L.check(L.lib.bqa_profile_end(self.handle, ms, fl, n), "bqa_profile_end")
        names = ("gemm_fwd", "gemm_dgrad", "gemm_wgrad", "attn_fwd", "attn_bwd", "ln_fwd", "ln_bwd", "embed_fwd",
                 "embed_bwd")
        return {k: (ms[i], fl[i], int(n[i])) for i, k in enumerate(names)}

    def launch_count(self):
        return int(L.lib.bqa_launch_count(self.handle))


def make_loss_config(*, w_start=1.0, w_end=1.0, w_start_reg=1.0, w_end_reg=1.0, w_cls=1.0, loss="smooth",
                     smooth_alpha=0.01, cls_weight=None, loss_scale=1.0, focal_alpha=1.0, focal_gamma=2.0):
    lc = L.LossConfig()
    lc.w_start, lc.w_end, lc.w_start_reg, lc.w_end_reg, lc.w_cls = w_start, w_end, w_start_reg, w_end_reg, w_cls
    if loss == "smooth":
        lc.cls_loss = L.CLS_SMOOTH
    elif loss == "ce":
        lc.cls_loss = L.CLS_CE
    elif loss == "focal":
        lc.cls_loss = L.CLS_FOCAL
    else:
        raise NotImplementedError(f"class loss {loss!r} (reference init.py:22-31 offers 'smooth', 'ce', 'focal')")
    lc.focal_alpha, lc.focal_gamma = float(focal_alpha), float(focal_gamma)
    lc.smooth_alpha = smooth_alpha
    lc.cls_weight = cls_weight.data_ptr() if cls_weight is not None else None
    lc.loss_scale = loss_scale
    return lc
