"""Optimizer of the recipe: HF AdamW(correct_bias=False) as constructed by the reference in
modules/init.py:125-138, fused with the global-norm clip of trainer.py:221-225 and
`optimizer.zero_grad()` (trainer.py:287) into two kernel passes over the model's flat buffers
(csrc/optimizer.cu).  The reference file of this name holds AdaMod (optim.py:42-100, selected by
`optimizer=adamod`): it is fused the same way (bqa_clip_adamod), with its third state tensor
`exp_avg_lr`.

The object is a real `torch.optim.Optimizer`: `param_groups[i]['lr']` is what LambdaLR rewrites
and what `Trainer._get_lr` reads (trainer.py:194-195), and `state_dict()` / `load_state_dict()`
produce / accept the per-parameter `{step, exp_avg, exp_avg_sq}` layout of the reference's
checkpoints (trainer.py:366,396); the moment tensors are views into two flat buffers.
"""
import ctypes

import torch

from .... import _lib as L

CHUNK_ELEMS = 65536


class FusedAdamW(torch.optim.Optimizer):
    # per-parameter state tensors (checkpoint names) and the flat buffers they are views of
    _STATE = (('exp_avg', '_m'), ('exp_avg_sq', '_v'))

    def __init__(self, params, model, lr=1e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0, correct_bias=False):
        if correct_bias:
            raise NotImplementedError("the reference always passes correct_bias=False (init.py:137); "
                                      "bias correction has no kernel")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._init_fused(model)

    def _init_fused(self, model):
        if len(self.param_groups) > 8:
            raise ValueError("at most 8 parameter groups")
        self._model = model
        self._max_grad_norm = 0.0
        self._built_for = None
        self._grads_clean = False
        self._sumsq_by_bucket = False

    # ---- flat state ------------------------------------------------------------------------
    def _build(self):
        eng = self._model.engine
        if self._built_for is eng:
            return eng
        dev = eng.device
        for _, attr in self._STATE:
            setattr(self, attr, torch.zeros(eng.total, dtype=torch.float32, device=dev))
        self._sumsq = torch.zeros(1, dtype=torch.float64, device=dev)
        self._scratch = torch.empty(L.lib.bqa_sumsq_scratch_floats(), dtype=torch.float32, device=dev)
        key_of = {id(p): k for k, p in self._model.named_parameters()}
        chunks = []
        self._keys = []
        for gi, group in enumerate(self.param_groups):
            for p in group['params']:
                key = key_of.get(id(p))
                if key is None:
                    raise ValueError("FusedAdamW only optimises parameters of the bound BertForQuestionAnswering")
                off, shape = eng.layout[key]
                n = p.numel()
                self._keys.append((p, key))
                for b in range(0, n, CHUNK_ELEMS):
                    e = min(b + CHUNK_ELEMS, n)
                    e_pad = (e + 3) // 4 * 4  # tensors are 128-element aligned: padding belongs to nobody
                    chunks.append((off + b, off + e_pad, float(group['weight_decay']), gi))
        arr = (L.AdamWChunk * len(chunks))()
        for i, (b, e, wd, gi) in enumerate(chunks):
            arr[i].begin, arr[i].end, arr[i].weight_decay, arr[i].group = b, e, wd, gi
        raw = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
        self._chunks = raw.to(dev)
        self._n_chunks = len(chunks)
        # expose the moments through the standard optimizer state (checkpoint compatibility)
        for p, key in self._keys:
            st = self.state[p]
            for name, attr in self._STATE:
                prev = st.get(name)
                view = eng.view(getattr(self, attr), key)
                # moments that exist already (loaded from a checkpoint before the first build, or living in
                # the buffers of a previous engine) move into the new flat buffers
                if prev is not None and prev.data_ptr() != view.data_ptr():
                    view.copy_(prev.to(device=dev, dtype=torch.float32).reshape(view.shape))
                st[name] = view
            st.setdefault('step', 0)
        self._built_for = eng
        return eng

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        eng = self._build()
        for p, key in self._keys:
            st = self.state[p]
            for name, attr in self._STATE:
                view = eng.view(getattr(self, attr), key)
                if name in st and st[name].data_ptr() != view.data_ptr():
                    view.copy_(st[name])
                st[name] = view
            st.setdefault('step', 0)

    # ---- step --------------------------------------------------------------------------------
    def prepare_clip(self, max_grad_norm):
        """Trainer._clip_grad_norm: launches the sum-of-squares pass; the coefficient is applied by
        the following step() on the device (no host synchronisation).  If the sum was already
        accumulated bucket by bucket during the backward (begin_bucket_sumsq / add_bucket_sumsq), nothing
        is launched here."""
        eng = self._build()
        self._max_grad_norm = float(max_grad_norm) if max_grad_norm else 0.0
        if self._max_grad_norm > 0.0 and not self._sumsq_by_bucket:
            self._sumsq.zero_()
            L.check(L.lib.bqa_grad_sumsq(eng.grads.data_ptr(), eng.total, self._sumsq.data_ptr(),
                                         self._scratch.data_ptr(), L.current_stream_ptr()), "bqa_grad_sumsq")
        self._sumsq_by_bucket = False

    def begin_bucket_sumsq(self):
        """Start of an optimizer step's last backward: the sum of squares of the gradient will be accumulated
        per bucket (add_bucket_sumsq) as the buckets become final, overlapped with the rest of the backward.
        Call with the stream current on which add_bucket_sumsq will run."""
        self._build()
        self._sumsq.zero_()
        self._sumsq_by_bucket = True

    def add_bucket_sumsq(self, begin, end):
        """*sumsq += sum(g[begin:end]^2) on the current stream (stream-ordered with the other buckets)."""
        eng = self._build()
        L.check(L.lib.bqa_grad_sumsq(eng.grads.data_ptr() + 4 * begin, end - begin, self._sumsq.data_ptr(),
                                     self._scratch.data_ptr(), L.current_stream_ptr()), "bqa_grad_sumsq")

    def grad_norm(self):
        """Total gradient norm seen by the last prepare_clip (synchronises)."""
        return float(self._sumsq.sqrt().item())

    @torch.no_grad()
    def step(self, closure=None):
        if closure is not None:
            raise NotImplementedError("closures are not supported")
        eng = self._build()
        g0 = self.param_groups[0]
        lrs = (ctypes.c_float * len(self.param_groups))(*[float(g['lr']) for g in self.param_groups])
        b1, b2 = g0['betas']
        self._launch_update(eng, g0, lrs, b1, b2)
        for p, _ in self._keys:
            self.state[p]['step'] += 1
        self._max_grad_norm = 0.0
        self._grads_clean = True
        self._model.grads_dirty = False  # the kernel zeroed the flat gradient buffer
        self._model.mark_weights_synced()

    def zero_grad(self, set_to_none=False):
        """Gradients live in the engine's flat buffer; step() already cleared it."""
        eng = self._build()
        if not (self._grads_clean and not self._model.grads_dirty):
            eng.grads.zero_()
        self._grads_clean = True
        self._model.grads_dirty = False
        self._model.restore_grad_views()

    def _launch_update(self, eng, g0, lrs, b1, b2):
        L.check(L.lib.bqa_clip_adamw(eng.params.data_ptr(), eng.grads.data_ptr(), self._m.data_ptr(),
                                     self._v.data_ptr(), eng.shadow.data_ptr(), self._chunks.data_ptr(),
                                     self._n_chunks, self._sumsq.data_ptr(), self._max_grad_norm, lrs,
                                     len(self.param_groups), b1, b2, g0['eps'], 1, L.current_stream_ptr()),
                "bqa_clip_adamw")


class AdaMod(FusedAdamW):
    """AdaMod (reference modules/model/trainer/optim.py:6-100) over the model's flat buffers: the
    reference's constructor arguments plus the model whose parameters are optimised (init.py passes
    it).  State per parameter: {step, exp_avg, exp_avg_sq, exp_avg_lr}, as in the reference."""
    _STATE = (('exp_avg', '_m'), ('exp_avg_sq', '_v'), ('exp_avg_lr', '_s'))

    def __init__(self, params, model, lr=1e-3, betas=(0.9, 0.999), beta3=0.999, eps=1e-8, weight_decay=0):
        if not 0.0 <= lr:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if not 0.0 <= eps:
            raise ValueError("Invalid epsilon value: {}".format(eps))
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError("Invalid beta parameter at index 0: {}".format(betas[0]))
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError("Invalid beta parameter at index 1: {}".format(betas[1]))
        if not 0.0 <= beta3 < 1.0:
            raise ValueError("Invalid beta3 parameter: {}".format(beta3))
        torch.optim.Optimizer.__init__(self, params, dict(lr=lr, betas=betas, beta3=beta3, eps=eps,
                                                          weight_decay=weight_decay))
        self._init_fused(model)

    def _launch_update(self, eng, g0, lrs, b1, b2):
        step = self.state[self._keys[0][0]]['step'] + 1  # every parameter is updated on every step
        L.check(L.lib.bqa_clip_adamod(eng.params.data_ptr(), eng.grads.data_ptr(), self._m.data_ptr(),
                                      self._v.data_ptr(), self._s.data_ptr(), eng.shadow.data_ptr(),
                                      self._chunks.data_ptr(), self._n_chunks, self._sumsq.data_ptr(),
                                      self._max_grad_norm, lrs, len(self.param_groups), b1, b2, g0['beta3'],
                                      g0['eps'], int(step), 1, L.current_stream_ptr()), "bqa_clip_adamod")
