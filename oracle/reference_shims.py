"""Makes the UNMODIFIED reference importable in this authoring container.  TEST INFRASTRUCTURE ONLY.

Used by oracle/make_golden.py (golden-vector generation), oracle/ref_runner.py (the timed CPU arm of
bench.py) and by tests that are skipped when no reference tree is available.  The reference's modules
are imported unmodified — from /root/reference where that exists, else from the staged copy
oracle/_ref (oracle/stage_ref.py; git-ignored) — after installing the stand-ins that
SURVEY.md §8c lists for dependencies this image lacks:
  nltk            -> empty module (only used for sentence splitting of real data)
  configargparse  -> the argparse-based shim of the product package (-c files, key=value lines)
  transformers.AdamW -> restatement of transformers 3.5.1 AdamW (class removed upstream)
  BertModel.from_pretrained -> config-only random init (no network, no checkpoint cache)
"""
import os
import sys
import types

def _find_reference_root():
    """/root/reference in the authoring container; on the GPU box the unmodified copy that
    oracle/stage_ref.py put under oracle/_ref (git-ignored, travels with the snapshot)."""
    env = os.environ.get("BQA_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isdir("/root/reference/modules"):
        return "/root/reference"
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


REFERENCE_ROOT = _find_reference_root()
REFERENCE_MODULES = os.path.join(REFERENCE_ROOT, "modules")


def reference_available():
    return os.path.isdir(REFERENCE_MODULES)


class HFAdamW:
    """torch.optim.Optimizer restating transformers 3.5.1 optimization.AdamW (PARITY UNPINNED)."""

    def __new__(cls, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0, correct_bias=True):
        import math

        import torch

        class _AdamW(torch.optim.Optimizer):
            def __init__(self, params, lr, betas, eps, weight_decay, correct_bias):
                super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                                              correct_bias=correct_bias))

            @torch.no_grad()
            def step(self, closure=None):
                for group in self.param_groups:
                    for p in group["params"]:
                        if p.grad is None:
                            continue
                        grad = p.grad
                        state = self.state[p]
                        if len(state) == 0:
                            state["step"] = 0
                            state["exp_avg"] = torch.zeros_like(p)
                            state["exp_avg_sq"] = torch.zeros_like(p)
                        exp_avg, exp_avg_sq = state["exp_avg"], state["exp_avg_sq"]
                        beta1, beta2 = group["betas"]
                        state["step"] += 1
                        exp_avg.mul_(beta1).add_(grad, alpha=1.0 - beta1)
                        exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1.0 - beta2)
                        denom = exp_avg_sq.sqrt().add_(group["eps"])
                        step_size = group["lr"]
                        if group["correct_bias"]:
                            bc1 = 1.0 - beta1 ** state["step"]
                            bc2 = 1.0 - beta2 ** state["step"]
                            step_size = step_size * math.sqrt(bc2) / bc1
                        p.addcdiv_(exp_avg, denom, value=-step_size)
                        if group["weight_decay"] > 0.0:
                            p.add_(p, alpha=-group["lr"] * group["weight_decay"])

        return _AdamW(params, lr, betas, eps, weight_decay, correct_bias)


class _OfflineBert:
    """Stand-in for MODELS['bert']: from_pretrained(name, **kw) -> randomly initialised BertModel."""
    shapes = {}

    @classmethod
    def from_pretrained(cls, name, **kw):
        from transformers import BertConfig, BertModel
        kw = dict(kw)  # num_labels stays: model.py:35 reads config.num_labels for the class head
        shape = dict(cls.shapes.get(name, {}))
        cfg = BertConfig(attn_implementation="eager", **shape, **kw)
        return BertModel(cfg)


class _OfflineRoberta:
    """Stand-in for MODELS['roberta']: from_pretrained(name, **kw) -> randomly initialised RobertaModel."""
    shapes = {}

    @classmethod
    def from_pretrained(cls, name, **kw):
        from transformers import RobertaConfig, RobertaModel
        shape = dict(cls.shapes.get(name, {}))
        cfg = RobertaConfig(attn_implementation="eager", **shape, **dict(kw))
        return RobertaModel(cfg)


class StubTokenizer:
    """What DummyDataset / collate_fun need from a tokenizer (dummy_dataset.py:21-23,39)."""
    model_name = "bert"
    pad_token_id, unk_token_id, cls_token_id, sep_token_id = 0, 100, 101, 102

    def __init__(self, vocab=30522):
        self._n = vocab

    def __len__(self):
        return self._n


def import_reference(extra_shapes=None):
    """Returns a namespace of the reference's modules (model, loss, dummy dataset, collate, trainer, init)."""
    if not reference_available():
        raise RuntimeError(f"{REFERENCE_MODULES} not found")
    if REFERENCE_MODULES not in sys.path:
        sys.path.insert(0, REFERENCE_MODULES)
    sys.modules.setdefault("nltk", types.ModuleType("nltk"))
    if "configargparse" not in sys.modules:
        from __graft_entry__ import load_package
        load_package()
        import importlib
        sys.modules["configargparse"] = importlib.import_module("bqa_b200.configargparse_shim")
    import model.model.model as ref_model_mod
    import model.model.loss as ref_loss_mod
    import model.dataset as ref_dataset
    import model.trainer.trainer as ref_trainer_mod
    if extra_shapes:
        _OfflineBert.shapes.update(extra_shapes)
        _OfflineRoberta.shapes.update(extra_shapes)
    ref_model_mod.MODELS["bert"] = _OfflineBert
    ref_model_mod.MODELS["roberta"] = _OfflineRoberta
    import transformers
    sys.modules["transformers"].AdamW = HFAdamW
    transformers.AdamW = HFAdamW
    import init as ref_init
    ns = types.SimpleNamespace(model=ref_model_mod, loss=ref_loss_mod, dataset=ref_dataset, trainer=ref_trainer_mod,
                               init=ref_init)
    try:
        import model.inference.predictor as ref_pred
        ns.predictor = ref_pred
    except Exception as e:  # pragma: no cover
        ns.predictor = None
        ns.predictor_error = repr(e)
    return ns
