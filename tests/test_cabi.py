"""The C-ABI library loads without a GPU and exports every symbol include/bqa.h declares; the
ctypes binding covers the same set; compute entry points fail loudly without a device."""
import ctypes
import os
import re
import subprocess

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "bqa.h")


def declared_symbols():
    text = open(HEADER).read()
    return sorted(set(re.findall(r"BQA_API\s+[\w\s\*]+?\b(bqa_\w+)\s*\(", text)))


def test_header_declares_entry_points():
    syms = declared_symbols()
    assert len(syms) >= 25
    for must in ("bqa_forward", "bqa_backward", "bqa_loss", "bqa_clip_adamw", "bqa_gemm_bf16", "bqa_decode"):
        assert must in syms


def test_library_exports_every_declared_symbol(pkg):
    out = subprocess.run(["nm", "-D", "--defined-only", pkg._lib.LIB_PATH], capture_output=True, text=True, check=True)
    exported = {line.split()[-1] for line in out.stdout.splitlines() if " T " in line}
    missing = [s for s in declared_symbols() if s not in exported]
    assert not missing, f"declared in bqa.h but not exported: {missing}"
    extra = [s for s in exported if s.startswith("bqa_") and s not in declared_symbols()]
    assert not extra, f"exported but undeclared: {extra}"


def test_binding_covers_header(pkg):
    assert sorted(pkg._lib.SIGNATURES) == declared_symbols()
    assert pkg._lib.lib.bqa_abi_version() == pkg._lib.ABI_VERSION


def test_struct_layouts_match_header(pkg):
    L = pkg._lib
    assert ctypes.sizeof(L.AdamWChunk) == 24
    assert ctypes.sizeof(L.Config) == 52  # 8 shape ints, 3 floats, pad_token_id, position_mode
    assert ctypes.sizeof(L.Layout) == 8 * len(L.LAYOUT_FIELDS)
    assert ctypes.sizeof(L.GemmDesc) % 8 == 0


def test_param_layout_is_pure_host_code(pkg):
    E = pkg.engine
    cfg = E.make_config(E.MODEL_SHAPES["bert-base-uncased"])
    layout, total = E.param_layout(cfg)
    assert len(layout) == 207  # 207 tensors
    assert sum(E._numel(s) for _, s in layout.values()) == 109_489_161
    assert total % 128 == 0
    offs = sorted((o, E._numel(s)) for o, s in layout.values())
    for (o0, n0), (o1, _) in zip(offs, offs[1:]):
        assert o0 + n0 <= o1 and o1 % 128 == 0
    # Q/K/V weights of a layer are adjacent: one [3H, H] operand for the fused projection
    p = "transformer.encoder.layer.3.attention.self."
    q, k, v = (layout[p + n + ".weight"][0] for n in ("query", "key", "value"))
    assert k - q == 768 * 768 and v - k == 768 * 768


def test_bad_config_is_rejected(pkg):
    E = pkg.engine
    bad = dict(E.MODEL_SHAPES["bert-base-uncased"], hidden=770)
    with pytest.raises(RuntimeError, match="unsupported shape"):
        E.param_layout(E.make_config(bad))


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(pkg):
    E = pkg.engine
    with pytest.raises(RuntimeError, match="no CPU implementation"):
        E.Engine(E.make_config(E.MODEL_SHAPES["bert-base-uncased"]), "cpu")
    L = pkg._lib
    d = L.GemmDesc()
    d.M = d.N = d.K = 128
    assert L.lib.bqa_gemm_bf16(ctypes.byref(d), None) != 0
    assert L.lib.bqa_last_error()
