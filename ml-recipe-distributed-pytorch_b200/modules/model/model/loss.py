"""Losses of the recipe with the reference's class names and call signature
(modules/model/model/loss.py:5-106).

`WeightedLoss.__call__(preds, targets, avg_meters=...)` returns a scalar that supports
`.backward()` exactly like the reference's; its value, the five per-key losses reported through
`avg_meters`, and the gradient w.r.t. the five model outputs all come from ONE fused kernel
(bqa_qa_loss) and ONE device->host copy, instead of the reference's per-key ATen graphs and six
blocking `.item()` calls (loss.py:93,98).
"""
import ctypes

import torch
import torch.nn as nn

from .... import _lib as L
from .... import engine as _engine

_KEYS = ('start_class', 'end_class', 'start_reg', 'end_reg', 'cls')


class LabelSmoothingLossWithLogits(nn.Module):
    """Configuration holder: KLDiv(batchmean) against a smoothed one-hot target (loss.py:5-38)."""

    def __init__(self, n_classes, *, smoothing=0.0, ignore_index=-100):
        super().__init__()
        assert 0 <= smoothing <= 1
        if ignore_index != -100:
            raise NotImplementedError("the fused loss kernel implements the reference's only use: ignore_index=-100")
        self.n_classes, self.smoothing, self.ignore_index = n_classes, smoothing, ignore_index


class FocalLossWithLogits(nn.Module):
    """Configuration holder: NLLLoss(ignore_index=-1, mean) of alpha (1 - p)^gamma log p (loss.py:57-71)."""

    def __init__(self, alpha=1, gamma=2, *, ignore_index=-1, reduction='mean'):
        super().__init__()
        if ignore_index != -1 or reduction != 'mean':
            raise NotImplementedError("the fused loss kernel implements the reference's only use: "
                                      "ignore_index=-1, reduction='mean'")
        self.alpha, self.gamma = alpha, gamma


def _run_loss_kernel(outs, targets, lc, with_grad):
    """outs: 5 fp32 CUDA tensors in _KEYS order.  Returns (losses[6] device tensor, grads or None)."""
    start, end, reg_s, reg_e, cls = [o.detach().to(torch.float32).contiguous() for o in outs]
    dev = start.device
    if dev.type != "cuda":
        raise RuntimeError("WeightedLoss runs on sm_100a only (there is no CPU implementation of this path)")
    B, S = start.shape
    C = cls.shape[1]
    ys = targets['start_class'].to(device=dev, dtype=torch.int64).contiguous()
    ye = targets['end_class'].to(device=dev, dtype=torch.int64).contiguous()
    yc = targets['cls'].to(device=dev, dtype=torch.int64).contiguous()
    yrs = targets['start_reg'].to(device=dev, dtype=torch.float32).contiguous()
    yre = targets['end_reg'].to(device=dev, dtype=torch.float32).contiguous()
    losses = torch.empty(6, dtype=torch.float32, device=dev)
    scratch = torch.empty(5 * B, dtype=torch.float32, device=dev)
    grads = None
    if with_grad:
        grads = [torch.empty_like(start), torch.empty_like(end), torch.empty_like(reg_s), torch.empty_like(reg_e),
                 torch.empty_like(cls)]
    g = grads or [None] * 5
    L.check(L.lib.bqa_qa_loss(start.data_ptr(), end.data_ptr(), cls.data_ptr(), reg_s.data_ptr(), reg_e.data_ptr(),
                              ys.data_ptr(), ye.data_ptr(), yc.data_ptr(), yrs.data_ptr(), yre.data_ptr(),
                              L.ptr(g[0]), L.ptr(g[1]), L.ptr(g[4]), L.ptr(g[2]), L.ptr(g[3]), scratch.data_ptr(),
                              losses.data_ptr(), B, S, C, ctypes.byref(lc), int(with_grad), L.current_stream_ptr()),
            "bqa_qa_loss")
    return losses, grads


class _FusedLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, start, end, reg_s, reg_e, cls, targets, lc, holder):
        with_grad = any(ctx.needs_input_grad[:5])
        losses, grads = _run_loss_kernel((start, end, reg_s, reg_e, cls), targets, lc, with_grad)
        holder.append(losses)
        ctx.grads = grads
        return losses[5].clone()

    @staticmethod
    def backward(ctx, g):
        if ctx.grads is None:
            return (None,) * 8
        return tuple(d * g for d in ctx.grads) + (None, None, None)


class WeightedLoss:
    def __init__(self, init_losses):
        self._losses = init_losses
        self._cls_weight = None

    def loss_config(self, loss_scale=1.0, device=None):
        """The (loss_fn, weight) table of init.py:34-38 as the kernel's configuration struct."""
        lo = self._losses
        cls_fn = lo['cls'][0]
        kw = dict(w_start=float(lo['start_class'][1]), w_end=float(lo['end_class'][1]),
                  w_start_reg=float(lo['start_reg'][1]), w_end_reg=float(lo['end_reg'][1]), w_cls=float(lo['cls'][1]),
                  loss_scale=loss_scale)
        if isinstance(cls_fn, LabelSmoothingLossWithLogits):
            kw.update(loss='smooth', smooth_alpha=float(cls_fn.smoothing))
        elif isinstance(cls_fn, nn.CrossEntropyLoss):
            w = cls_fn.weight
            if w is not None:
                self._cls_weight = w.detach().to(device=device or w.device, dtype=torch.float32).contiguous()
            kw.update(loss='ce', cls_weight=self._cls_weight)
        elif isinstance(cls_fn, FocalLossWithLogits):
            kw.update(loss='focal', focal_alpha=float(cls_fn.alpha), focal_gamma=float(cls_fn.gamma))
        else:
            raise NotImplementedError(f"class loss {type(cls_fn).__name__} has no kernel")
        return _engine.make_loss_config(**kw)

    def __call__(self, preds, targets, *, avg_meters=None):
        assert set(preds.keys()).intersection(set(targets.keys())) == set(preds.keys())
        assert set(self._losses.keys()).intersection(set(targets.keys())) == set(self._losses.keys())
        lc = self.loss_config(device=preds['start_class'].device)
        holder = []
        full_loss = _FusedLossFn.apply(preds['start_class'], preds['end_class'], preds['start_reg'],
                                       preds['end_reg'], preds['cls'], targets, lc, holder)
        if avg_meters is not None:
            vals = holder[0].tolist()  # the step's only device->host synchronisation
            for i, key in enumerate(_KEYS):
                avg_meters[key] = vals[i]
            avg_meters['loss'] = vals[5]
        return full_loss

    def to(self, device):
        for key in self._losses.keys():
            self._losses[key][0].to(device)
        return self
