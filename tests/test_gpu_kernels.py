"""Parity of the individual sm_100a kernels (through the C ABI) against torch fp32 on the same
bf16 inputs.  Tolerances: outputs are bf16 (relative rounding 2^-9 = 2e-3 per element), fp32
accumulation; the bounds below are a few times the expected bf16 rounding error."""
import ctypes
import importlib

import numpy as np
import pytest
import torch

from oracle.dropout_hash import keep_mask

pytestmark = pytest.mark.gpu
DEV = "cuda"


def rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-20)).item()


def rnd(*s, scale=0.5):
    return (torch.randn(*s, device=DEV) * scale).to(torch.bfloat16)


GEMM_CASES = [
    # M, N, K, mode, extras
    (128, 256, 64, "tn", {}),
    (1000, 768, 768, "tn", {"bias": True, "residual": True}),            # ragged M
    (256, 320, 128, "tn", {"bias": True}),                                 # partial N tile
    (1024, 768, 768, "tn", {"bias": True, "residual": True, "dropout": 0.1}),
    (1024, 3072, 768, "tn_gelu", {"bias": True}),
    (1024, 768, 3072, "nn", {"residual": True}),
    (1024, 3072, 768, "nn_gelu_bwd", {}),
    (768, 768, 1000, "nt", {"k_splits": 3}),                               # ragged K, split-K reduce-add
    (3072, 768, 12288, "nt", {"k_splits": 8}),                             # wgrad at the BASELINE size
    (12288, 2304, 768, "tn", {"bias": True}),                              # fused QKV at the BASELINE size
    # side operand by TMA with blocks partly / wholly outside the operand, in the one-CTA kernel (M <= 128) and the pair one
    (100, 320, 128, "tn", {"bias": True, "residual": True, "dropout": 0.1}),
    (300, 192, 64, "nn", {"residual": True}),
    (200, 192, 64, "nn_gelu_bwd", {}),
    (96, 64, 64, "nn_gelu_bwd", {}),
]


@pytest.mark.parametrize("M,N,K,mode,ex", GEMM_CASES)
def test_gemm_variants(pkg, M, N, K, mode, ex):
    L = pkg._lib
    torch.manual_seed(0)
    bias = torch.randn(N, device=DEV) * 0.5 if ex.get("bias") else None
    residual = rnd(M, N) if ex.get("residual") else None
    p = ex.get("dropout", 0.0)
    kw, C2, aux = {}, None, None
    if mode.startswith("tn"):
        A, B = rnd(M, K), rnd(N, K)
        ref = A.float() @ B.float().t()
    elif mode.startswith("nn"):
        A, B = rnd(M, K), rnd(K, N)
        ref = A.float() @ B.float()
        kw = {"b_mn_major": True}
    else:
        A, B = rnd(K, M), rnd(K, N)
        ref = A.float().t() @ B.float()
        kw = {"a_mn_major": True, "b_mn_major": True}
    if bias is not None:
        ref = ref + bias
    if mode == "tn_gelu":
        C, C2, epi = torch.empty(M, N, device=DEV, dtype=torch.bfloat16), torch.empty(M, N, device=DEV, dtype=torch.bfloat16), L.EPI_GELU_FWD
    elif mode == "nn_gelu_bwd":
        aux = rnd(M, N, scale=1.0)
        ref = ref * aux.float()
        C, epi = torch.empty(M, N, device=DEV, dtype=torch.bfloat16), L.EPI_GELU_BWD
    elif mode == "nt":
        C = torch.randn(M, N, device=DEV)
        ref = ref + C
        epi = L.EPI_F32_ADD
    else:
        C, epi = torch.empty(M, N, device=DEV, dtype=torch.bfloat16), L.EPI_LINEAR
        if p > 0:
            keep = keep_mask(1234, 7, p, M * N, device=DEV).view(M, N)
            ref = torch.where(keep, ref / (1 - p), torch.zeros_like(ref))
        if residual is not None:
            ref = ref + residual.float()
    L.gemm_bf16(A, B, C, epilogue=epi, C2=C2, bias=bias, residual=residual, aux=aux, k_splits=ex.get("k_splits", 1),
                dropout_p=p, drop_seed=1234, drop_site=7, **kw)
    torch.cuda.synchronize()
    scale = ref.abs().max().item()
    tol = 1e-5 * scale + 1e-3 if epi == L.EPI_F32_ADD else 8e-3 * scale  # bf16 output: 2^-8 of the largest value
    if C2 is None:
        assert (C.float() - ref).abs().max().item() <= tol
    if C2 is not None:  # C holds gelu'(u), C2 gelu(u) for u = ref
        u = ref.clone().requires_grad_(True)
        g = torch.nn.functional.gelu(u)
        g.sum().backward()
        assert (C2.float() - g.detach()).abs().max().item() <= tol
        assert (C.float() - u.grad).abs().max().item() <= 8e-3


@pytest.mark.parametrize("Bt,S,A", [(2, 200, 2), (32, 384, 12)])
def test_gemm_fused_side_outputs(pkg, Bt, S, A):
    """The side outputs that the backward GEMMs' epilogues produce instead of separate passes, at a ragged small
    shape and at the BASELINE shape: (1) EPI_GELU_BWD's column sums of du (bias gradient of the intermediate
    projection), (2) the row-dot D = rowsum(dctx * ctx) per (batch, head, query) of the dgrad GEMM that produces
    dctx (read by the attention backward)."""
    L = pkg._lib
    torch.manual_seed(3)
    H, I = A * 64, A * 256
    M = Bt * S
    # (1) du = (dgm . W2) * gelu'   with W2 [H, I] as the MN-major operand; colsum_out += sum_m du[m, :]
    dgm, W2, gp = rnd(M, H), rnd(H, I, scale=0.05), rnd(M, I, scale=1.0)
    du = torch.empty(M, I, device=DEV, dtype=torch.bfloat16)
    cs = torch.full((I,), 0.5, device=DEV)
    L.gemm_bf16(dgm, W2, du, b_mn_major=True, epilogue=L.EPI_GELU_BWD, aux=gp, colsum_out=cs)
    torch.cuda.synchronize()
    ref = (dgm.float() @ W2.float()) * gp.float()
    assert (du.float() - ref).abs().max().item() <= 8e-3 * ref.abs().max().item()
    want = ref.sum(0) + 0.5
    assert (cs - want).norm().item() <= 3e-3 * want.norm().item() + 1e-3 * ref.abs().max().item() * M ** 0.5
    # (2) dctx = dgm . Wo (MN-major Wo [H, H]); D[b, h, q] = sum_d dctx[b q, 64 h + d] * ctx[b q, 64 h + d]
    Wo, ctx = rnd(H, H, scale=0.05), rnd(M, H)
    dctx = torch.empty(M, H, device=DEV, dtype=torch.bfloat16)
    D = torch.full((Bt, A, S), 7.0, device=DEV)
    L.gemm_bf16(dgm, Wo, dctx, b_mn_major=True, dot_with=ctx, dot_out=D, dot_rows=S)
    torch.cuda.synchronize()
    ref = dgm.float() @ Wo.float()
    assert (dctx.float() - ref).abs().max().item() <= 8e-3 * ref.abs().max().item()
    wantD = (dctx.float() * ctx.float()).view(Bt, S, A, 64).sum(-1).permute(0, 2, 1)
    assert (D - wantD).abs().max().item() <= 2e-3 * wantD.abs().max().item()


def test_gemm_linearity_at_full_size(pkg):
    """Size-independent property at the BASELINE FFN shape: C(A1 + A2) == C(A1) + C(A2) up to rounding."""
    L = pkg._lib
    torch.manual_seed(1)
    M, N, K = 12288, 3072, 768
    A1, A2, B = rnd(M, K), rnd(M, K), rnd(N, K)
    A12 = (A1.float() + A2.float()).to(torch.bfloat16)
    outs = []
    for A in (A1, A2, A12):
        C = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
        L.gemm_bf16(A, B, C)
        outs.append(C.float())
    assert rel(outs[2], outs[0] + outs[1]) < 6e-3


ATTN_CASES = [(2, 128, 2, 0.0, False), (2, 384, 12, 0.0, False), (2, 512, 2, 0.0, False), (3, 200, 2, 0.0, True),
              (2, 384, 4, 0.1, False), (2, 50, 1, 0.0, True), (1, 1, 1, 0.0, False),
              # more tiles than SMs: every CTA of the persistent kernels walks several tiles, crosses
              # (batch, head) boundaries (K/V reload, key-bias refill) and reuses its barriers / TMEM buffers
              (16, 384, 12, 0.1, True), (8, 512, 12, 0.1, False), (40, 200, 2, 0.0, True), (150, 100, 2, 0.1, True)]


@pytest.mark.parametrize("B,S,A,p,use_mask", ATTN_CASES)
def test_attention_forward_backward(pkg, B, S, A, p, use_mask):
    L = pkg._lib
    torch.manual_seed(1)
    H = A * 64
    qkv = torch.randn(B, S, 3 * H, device=DEV).to(torch.bfloat16)
    mask = None
    if use_mask:
        lens = torch.randint(max(1, S // 2), S + 1, (B,), device=DEV)
        mask = torch.arange(S, device=DEV)[None, :] < lens[:, None]
    m8 = mask.to(torch.uint8).contiguous() if mask is not None else None
    ctx = torch.zeros(B, S, H, device=DEV, dtype=torch.bfloat16)
    lse2 = torch.zeros(B, A, S, device=DEV)
    keep_bits = torch.zeros(B, A, S, L.lib.bqa_attention_keep_words(S), dtype=torch.int32, device=DEV)
    st = L.current_stream_ptr()
    L.check(L.lib.bqa_attention_forward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep_bits.data_ptr(), L.ptr(m8), B, S, A, p, 77, 5,
                                        st), "attention forward")
    x = qkv.float().requires_grad_(True)
    q, k, v = [t.view(B, S, A, 64).transpose(1, 2) for t in x.split(H, dim=-1)]
    sc = (q @ k.transpose(-1, -2)) / 8.0
    if mask is not None:
        sc = sc.masked_fill(~mask[:, None, None, :], float("-inf"))
    pr = torch.softmax(sc, -1)
    if p > 0:
        keep = keep_mask(77, 5, p, pr.numel(), device=DEV).view(pr.shape)
        pr = torch.where(keep, pr / (1 - p), torch.zeros_like(pr))
    o = (pr @ v).transpose(1, 2).reshape(B, S, H)
    assert rel(ctx, o) < 8e-3
    assert (lse2 - torch.logsumexp(sc, -1) * 1.4426950408889634).abs().max().item() < 1e-3
    dctx = torch.randn(B, S, H, device=DEV).to(torch.bfloat16)
    o.backward(dctx.float())
    dqkv = torch.zeros(B, S, 3 * H, device=DEV, dtype=torch.bfloat16)
    dsum = torch.zeros(B, A, S, device=DEV)
    dbias = torch.zeros(3 * H, device=DEV)
    L.check(L.lib.bqa_attention_backward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep_bits.data_ptr(), dctx.data_ptr(),
                                         dqkv.data_ptr(), dsum.data_ptr(), L.ptr(m8), B, S, A, p, 77, 5, dbias.data_ptr(),
                                         st),
            "attention backward")
    torch.cuda.synchronize()
    # fused side output: gradient of the Q/K/V bias = column sums of dqkv, accumulated by the kernels' drains
    want_b = dqkv.float().sum(dim=(0, 1))
    assert (dbias - want_b).norm().item() <= 2e-3 * want_b.norm().item() + 1e-4 * max(1.0, dqkv.float().abs().max().item()) * (B * S) ** 0.5
    gscale = x.grad.abs().max().item()
    for i in range(3):
        got, want = dqkv[..., i * H:(i + 1) * H].float(), x.grad[..., i * H:(i + 1) * H]
        # norm-wise 1e-2, with a floor for blocks whose exact gradient is zero (e.g. dQ, dK when S == 1)
        assert (got - want).norm().item() <= 1e-2 * want.norm().item() + 1e-5 * max(1.0, gscale) * want.numel() ** 0.5


def test_attention_rows_sum_to_one_at_full_size(pkg):
    """V = ones => every context row must be exactly the softmax mass (1.0), any B / S / mask."""
    L = pkg._lib
    B, S, A = 32, 384, 12
    H = A * 64
    torch.manual_seed(2)
    qkv = torch.randn(B, S, 3 * H, device=DEV).to(torch.bfloat16)
    qkv[..., 2 * H:] = 1.0
    ctx = torch.zeros(B, S, H, device=DEV, dtype=torch.bfloat16)
    lse2 = torch.zeros(B, A, S, device=DEV)
    keep_bits = torch.zeros(B, A, S, L.lib.bqa_attention_keep_words(S), dtype=torch.int32, device=DEV)
    L.check(L.lib.bqa_attention_forward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep_bits.data_ptr(), None, B, S, A, 0.0, 0, 0,
                                        L.current_stream_ptr()), "attention forward")
    torch.cuda.synchronize()
    assert (ctx.float() - 1.0).abs().max().item() < 1e-2


@pytest.mark.parametrize("T,H", [(1000, 768), (515, 1024), (64, 128), (7, 64)])
def test_layernorm_and_colsum(pkg, T, H):
    L = pkg._lib
    torch.manual_seed(2)
    z, dy = torch.randn(T, H, device=DEV).to(torch.bfloat16), torch.randn(T, H, device=DEV).to(torch.bfloat16)
    g, b = torch.randn(H, device=DEV), torch.randn(H, device=DEV)
    y = torch.empty_like(z)
    st = L.current_stream_ptr()
    L.check(L.lib.bqa_layernorm_forward(z.data_ptr(), g.data_ptr(), b.data_ptr(), y.data_ptr(), T, H, 1e-12, st), "ln")
    zf, gf, bf = z.float().requires_grad_(True), g.clone().requires_grad_(True), b.clone().requires_grad_(True)
    ref = torch.nn.functional.layer_norm(zf, (H,), gf, bf, 1e-12)
    assert rel(y, ref) < 4e-3
    ref.backward(dy.float())
    dz, dzm = torch.empty_like(z), torch.empty_like(z)
    dg, db, dbias = (torch.zeros(H, device=DEV) for _ in range(3))
    L.check(L.lib.bqa_layernorm_backward(dy.data_ptr(), z.data_ptr(), g.data_ptr(), dz.data_ptr(), dzm.data_ptr(),
                                         dg.data_ptr(), db.data_ptr(), dbias.data_ptr(), T, H, 1e-12, 0.0, 0, 0.1, 9,
                                         4321, st), "ln bwd")
    torch.cuda.synchronize()
    assert rel(dz, zf.grad) < 4e-3 and rel(dg, gf.grad) < 1e-4 and rel(db, bf.grad) < 1e-4
    keep = keep_mask(4321, 9, 0.1, T * H, device=DEV).view(T, H)
    want = torch.where(keep, dz.float() / 0.9, torch.zeros_like(dz.float()))
    assert rel(dzm, want) < 4e-3
    assert rel(dbias, dzm.float().sum(0)) < 1e-4
    X = torch.randn(T, 3 * H, device=DEV).to(torch.bfloat16)
    cs = torch.zeros(3 * H, device=DEV)
    L.check(L.lib.bqa_colsum_bf16(X.data_ptr(), 3 * H, cs.data_ptr(), T, 3 * H, st), "colsum")
    torch.cuda.synchronize()
    assert rel(cs, X.float().sum(0)) < 1e-5


def test_dropout_keep_rate_per_site_on_device(pkg):
    """Statistics of the CUDA kernels' own dropout decisions (no restatement of the hash involved): at every
    kind of site the fraction of kept elements is 1 - p within 4 sigma of a binomial, and kept values carry
    the 1 / (1 - p) scale.  Sites: GEMM epilogue (hidden dropout after the attention-output / FFN-output
    projections), attention probabilities (keep bits + the context they produce), LayerNorm-backward mask."""
    L = pkg._lib
    p = 0.1
    st = L.current_stream_ptr()

    def check_rate(kept, n, what):
        sigma = (p * (1 - p) / n) ** 0.5
        assert abs(kept / n - (1 - p)) < 4 * sigma, f"{what}: keep rate {kept / n:.5f} over {n} elements"

    # GEMM epilogue: A = 0, bias = 1  =>  C = dropout(1) in {0, 1/(1-p)}
    M, N, K = 2048, 768, 64
    A, B = torch.zeros(M, K, device=DEV, dtype=torch.bfloat16), torch.zeros(N, K, device=DEV, dtype=torch.bfloat16)
    for site in (2, 3, 35):
        C = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
        L.gemm_bf16(A, B, C, bias=torch.ones(N, device=DEV), dropout_p=p, drop_seed=99, drop_site=site)
        torch.cuda.synchronize()
        vals = C.float()
        kept = int((vals != 0).sum())
        check_rate(kept, M * N, f"gemm epilogue site {site}")
        assert (vals[vals != 0] - 1 / (1 - p)).abs().max().item() < 8e-3
    # attention: Q = K = 0 => uniform probabilities 1/S; V = 1 => ctx[q] = (#kept keys of q) / (S (1-p))
    Bt, S, Ah = 8, 384, 4
    H = Ah * 64
    qkv = torch.zeros(Bt, S, 3 * H, device=DEV, dtype=torch.bfloat16)
    qkv[..., 2 * H:] = 1.0
    ctx = torch.zeros(Bt, S, H, device=DEV, dtype=torch.bfloat16)
    lse2 = torch.zeros(Bt, Ah, S, device=DEV)
    W = L.lib.bqa_attention_keep_words(S)
    keep_bits = torch.zeros(Bt, Ah, S, W, dtype=torch.int32, device=DEV)
    L.check(L.lib.bqa_attention_forward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep_bits.data_ptr(), None, Bt, S,
                                        Ah, p, 99, 4, st), "attention forward")
    torch.cuda.synchronize()
    words = keep_bits.view(Bt, Ah, S, 2, W // 2).to(torch.int64) & 0xFFFFFFFF
    shifts = torch.arange(32, device=DEV, dtype=torch.int64)
    per_layout = ((words[..., None] >> shifts) & 1).sum(dim=(-1, -2))        # [B, A, S, 2]: per query / per key
    n = Bt * Ah * S * S
    check_rate(int(per_layout[..., 0].sum()), n, "attention keep bits (query-major)")
    check_rate(int(per_layout[..., 1].sum()), n, "attention keep bits (key-major)")
    want_ctx = per_layout[..., 0].float() / (S * (1 - p))                      # [B, A, S]
    got_ctx = ctx.float().view(Bt, S, Ah, 64).permute(0, 2, 1, 3)
    assert (got_ctx - want_ctx[..., None]).abs().max().item() < 1e-2
    # LayerNorm backward: dz_masked is dz where kept (scaled), 0 where dropped
    T, Hh = 4096, 768
    z, dy = torch.randn(T, Hh, device=DEV).to(torch.bfloat16), torch.randn(T, Hh, device=DEV).to(torch.bfloat16)
    g = torch.ones(Hh, device=DEV)
    dz, dzm = torch.empty_like(z), torch.empty_like(z)
    dg, db, dbias = (torch.zeros(Hh, device=DEV) for _ in range(3))
    L.check(L.lib.bqa_layernorm_backward(dy.data_ptr(), z.data_ptr(), g.data_ptr(), dz.data_ptr(), dzm.data_ptr(),
                                         dg.data_ptr(), db.data_ptr(), dbias.data_ptr(), T, Hh, 1e-12, 0.0, 0, p, 6, 99,
                                         st), "ln bwd")
    torch.cuda.synchronize()
    live = dz.float() != 0
    check_rate(int(((dzm.float() != 0) & live).sum()), int(live.sum()), "layernorm backward mask")


def test_decode_bit_exact_against_reference_golden(pkg):
    from conftest import load_npz
    fx = load_npz("ref_decode.npz")
    L = pkg._lib
    s, e, c = (torch.from_numpy(fx[k]).to(DEV).contiguous() for k in ("start", "end", "cls"))
    B, S = s.shape
    idx = torch.zeros(3, B, dtype=torch.int64, device=DEV)
    val = torch.zeros(4, B, device=DEV)
    L.check(L.lib.bqa_decode(s.data_ptr(), e.data_ptr(), c.data_ptr(), idx[0].data_ptr(), idx[1].data_ptr(),
                             idx[2].data_ptr(), val[0].data_ptr(), val[1].data_ptr(), val[2].data_ptr(),
                             val[3].data_ptr(), B, S, 5, L.current_stream_ptr()), "decode")
    torch.cuda.synchronize()
    assert (idx[0].cpu().numpy() == fx["start_ids"]).all()   # includes a planted exact tie: first index wins
    assert (idx[1].cpu().numpy() == fx["end_ids"]).all()
    assert (idx[2].cpu().numpy() == fx["cls_ids"]).all()
    assert (val[3].cpu().numpy() == fx["scores"]).all()
    assert abs(val[2].cpu().numpy() - fx["cls_probas"]).max() < 1e-6


def test_decode_full_size_argmax_property(pkg):
    """At the inference batch of BASELINE config 5: kernel arg max == torch.max on the same logits."""
    L = pkg._lib
    B, S = 256, 512
    torch.manual_seed(3)
    s, e, c = torch.randn(B, S, device=DEV), torch.randn(B, S, device=DEV), torch.randn(B, 5, device=DEV)
    s[:, 100] = s[:, 400] = 50.0
    idx = torch.zeros(3, B, dtype=torch.int64, device=DEV)
    val = torch.zeros(4, B, device=DEV)
    L.check(L.lib.bqa_decode(s.data_ptr(), e.data_ptr(), c.data_ptr(), idx[0].data_ptr(), idx[1].data_ptr(),
                             idx[2].data_ptr(), val[0].data_ptr(), val[1].data_ptr(), val[2].data_ptr(),
                             val[3].data_ptr(), B, S, 5, L.current_stream_ptr()), "decode")
    torch.cuda.synchronize()
    assert (idx[0] == 100).all() and torch.equal(idx[1], e.max(-1)[1]) and torch.equal(idx[2], c.max(-1)[1])
    assert torch.equal(val[3], s.max(-1)[0] + e.max(-1)[0] - (s[:, 0] + e[:, 0]))


def test_optimizer_kernels_match_oracle(pkg):
    from oracle import bertqa_oracle as O
    L = pkg._lib
    torch.manual_seed(4)
    n = 4 * 65536 + 1000
    p, g = torch.randn(n, device=DEV), torch.randn(n, device=DEV) * 3
    m, v = torch.randn(n, device=DEV) * 0.1, torch.rand(n, device=DEV) * 0.1
    P = {"w.weight": p[:200000].clone(), "w.bias": p[200000:].clone()}
    G = {"w.weight": g[:200000].clone(), "w.bias": g[200000:].clone()}
    state = {"w.weight": {"step": 0, "exp_avg": m[:200000].clone(), "exp_avg_sq": v[:200000].clone()},
             "w.bias": {"step": 0, "exp_avg": m[200000:].clone(), "exp_avg_sq": v[200000:].clone()}}
    total = O.clip_grad_norm_(G, 1.0)
    O.hf_adamw_step(P, G, state, lr=1e-2, weight_decay=0.1)
    chunks = (L.AdamWChunk * 5)()
    spans = [(0, 65536, 0.1, 0), (65536, 131072, 0.1, 0), (131072, 200000, 0.1, 0), (200000, 262144, 0.0, 1),
             (262144, n, 0.0, 1)]
    for i, (b, e, wd, gi) in enumerate(spans):
        chunks[i].begin, chunks[i].end, chunks[i].weight_decay, chunks[i].group = b, e, wd, gi
    cd = torch.frombuffer(bytearray(bytes(chunks)), dtype=torch.uint8).to(DEV)
    sumsq = torch.zeros(1, dtype=torch.float64, device=DEV)
    scratch = torch.empty(L.lib.bqa_sumsq_scratch_floats(), device=DEV)
    shadow = torch.zeros(n, dtype=torch.bfloat16, device=DEV)
    st = L.current_stream_ptr()
    L.check(L.lib.bqa_grad_sumsq(g.data_ptr(), n, sumsq.data_ptr(), scratch.data_ptr(), st), "sumsq")
    lrs = (ctypes.c_float * 2)(1e-2, 1e-2)
    L.check(L.lib.bqa_clip_adamw(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), shadow.data_ptr(),
                                 cd.data_ptr(), 5, sumsq.data_ptr(), 1.0, lrs, 2, 0.9, 0.999, 1e-6, 1, st), "adamw")
    torch.cuda.synchronize()
    assert abs(sumsq.sqrt().item() - total.item()) < 1e-3 * total.item()
    want = torch.cat([P["w.weight"], P["w.bias"]])
    torch.testing.assert_close(p, want, rtol=2e-5, atol=2e-6)
    torch.testing.assert_close(m, torch.cat([state["w.weight"]["exp_avg"], state["w.bias"]["exp_avg"]]), rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(shadow.float(), want.to(torch.bfloat16).float(), rtol=0, atol=0)
    assert (g == 0).all()


def test_adamod_kernel_matches_reference_golden(pkg):
    """bqa_clip_adamod over a flat buffer against six steps of the reference's AdaMod class
    (tests/golden/ref_adamod.npz): parameters after every step and the final exp_avg_lr state."""
    from conftest import load_npz
    L = pkg._lib
    fx = load_npz("ref_adamod.npz")
    names = [str(n) for n in fx["names"]]
    no_decay = ("bias", "LayerNorm.bias", "LayerNorm.weight")
    offs, total = {}, 0
    for n in names:
        offs[n] = total
        total += (fx[f"p0/{n}"].size + 127) // 128 * 128
    p = torch.zeros(total, device=DEV)
    m, v, s = torch.zeros_like(p), torch.zeros_like(p), torch.zeros_like(p)
    g = torch.zeros_like(p)
    shadow = torch.zeros(total, dtype=torch.bfloat16, device=DEV)
    for n in names:
        t = torch.from_numpy(np.asarray(fx[f"p0/{n}"])).flatten()
        p[offs[n]:offs[n] + t.numel()] = t.to(DEV)
    chunks = (L.AdamWChunk * len(names))()
    for i, n in enumerate(names):
        size = fx[f"p0/{n}"].size
        decay = not any(nd in n for nd in no_decay)
        chunks[i].begin, chunks[i].end = offs[n], offs[n] + (size + 3) // 4 * 4
        chunks[i].weight_decay, chunks[i].group = (float(fx["weight_decay"]) if decay else 0.0), (0 if decay else 1)
    cd = torch.frombuffer(bytearray(bytes(chunks)), dtype=torch.uint8).to(DEV)
    lrs = (ctypes.c_float * 2)(float(fx["lr"]), float(fx["lr"]))
    st = L.current_stream_ptr()
    for t in range(6):
        for n in names:
            gt = torch.from_numpy(np.asarray(fx[f"g{t}/{n}"])).flatten()
            g[offs[n]:offs[n] + gt.numel()] = gt.to(DEV)
        L.check(L.lib.bqa_clip_adamod(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), s.data_ptr(),
                                      shadow.data_ptr(), cd.data_ptr(), len(names), None, 0.0, lrs, 2, 0.9, 0.999,
                                      0.999, 1e-8, t + 1, 1, st), "adamod")
        torch.cuda.synchronize()
        assert (g == 0).all()
        for n in names:
            want = torch.from_numpy(np.asarray(fx[f"p{t + 1}/{n}"])).flatten().to(DEV)
            torch.testing.assert_close(p[offs[n]:offs[n] + want.numel()], want, rtol=2e-5, atol=1e-6,
                                       msg=lambda msg, n=n, t=t: f"step {t} {n}: {msg}")
    for n in names:
        want = torch.from_numpy(np.asarray(fx[f"exp_avg_lr/{n}"])).flatten().to(DEV)
        torch.testing.assert_close(s[offs[n]:offs[n] + want.numel()], want, rtol=2e-5, atol=1e-9)
    torch.testing.assert_close(shadow.float(), p.to(torch.bfloat16).float(), rtol=0, atol=0)


def test_focal_loss_kernel_matches_reference_golden(pkg):
    """The fused loss kernel with cls_loss = focal against the reference's FocalLossWithLogits: value and
    gradient w.r.t. the class logits (the other four losses get dummy inputs)."""
    from conftest import load_npz
    E = pkg.engine
    loss_mod = importlib.import_module("bqa_b200.modules.model.model.loss")
    fx = load_npz("ref_focal.npz")
    logits = torch.from_numpy(np.asarray(fx["logits"])).to(DEV)
    target = torch.from_numpy(np.asarray(fx["target"])).to(DEV)
    B, S = logits.shape[0], 16
    outs = (torch.zeros(B, S, device=DEV), torch.zeros(B, S, device=DEV), torch.full((B,), 0.5, device=DEV),
            torch.full((B,), 0.5, device=DEV), logits)
    targets = {"start_class": torch.zeros(B, dtype=torch.int64), "end_class": torch.zeros(B, dtype=torch.int64),
               "start_reg": torch.full((B,), 0.5), "end_reg": torch.full((B,), 0.5), "cls": target}
    for alpha, gamma in ((1, 2), (0.25, 3), (2, 0), (1, 1)):
        lc = E.make_loss_config(loss="focal", focal_alpha=alpha, focal_gamma=gamma)
        losses, grads = loss_mod._run_loss_kernel(outs, targets, lc, True)
        torch.cuda.synchronize()
        assert abs(losses[4].item() - float(fx[f"loss/{alpha}/{gamma}"])) < 2e-5
        torch.testing.assert_close(grads[4].cpu(), torch.from_numpy(np.asarray(fx[f"grad/{alpha}/{gamma}"])),
                                   rtol=2e-4, atol=2e-6)
