"""Launches each hot kernel twice at the BASELINE shapes (for `ncu --set full`)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from __graft_entry__ import load_package  # noqa: E402

pkg = load_package()
L = pkg._lib
dev = "cuda"
torch.manual_seed(0)
T, H, I = 12288, 768, 3072
which = sys.argv[1:] or ["gemm", "attn", "rows"]


def rnd(*s):
    return (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)


if "gemm" in which:
    x, w1, b1 = rnd(T, H), rnd(I, H), torch.randn(I, device=dev)
    u, g = torch.empty(T, I, device=dev, dtype=torch.bfloat16), torch.empty(T, I, device=dev, dtype=torch.bfloat16)
    w2, res = rnd(H, I), rnd(T, H)
    z2 = torch.empty(T, H, device=dev, dtype=torch.bfloat16)
    dgm = rnd(T, H)
    du = torch.empty(T, I, device=dev, dtype=torch.bfloat16)
    dw = torch.zeros(I, H, device=dev)
    for _ in range(2):
        L.gemm_bf16(x, w1, u, epilogue=L.EPI_GELU_FWD, C2=g, bias=b1)                       # ffn1
        L.gemm_bf16(g, w2, z2, bias=b1[:H].contiguous(), residual=res, dropout_p=0.1, drop_seed=1, drop_site=2)  # ffn2
        L.gemm_bf16(dgm, w2, du, b_mn_major=True, epilogue=L.EPI_GELU_BWD, aux=u)            # dgrad ffn2
        L.gemm_bf16(du, w1, z2, b_mn_major=True, residual=res)                               # dgrad ffn1
        L.gemm_bf16(du, x, dw, a_mn_major=True, b_mn_major=True, epilogue=L.EPI_F32_ADD, k_splits=4)  # wgrad ffn1
    torch.cuda.synchronize()
if "attn" in which:
    B, S, A = 32, 384, 12
    qkv = torch.randn(B, S, 3 * H, device=dev).to(torch.bfloat16)
    ctx = torch.zeros(B, S, H, device=dev, dtype=torch.bfloat16)
    lse2 = torch.zeros(B, A, S, device=dev)
    keep_bits = torch.zeros(B, A, S, L.lib.bqa_attention_keep_words(S), dtype=torch.int32, device=dev)
    dctx = torch.randn(B, S, H, device=dev).to(torch.bfloat16)
    dqkv = torch.zeros(B, S, 3 * H, device=dev, dtype=torch.bfloat16)
    dsum = torch.zeros(B, A, S, device=dev)
    st = L.current_stream_ptr()
    for _ in range(2):
        L.check(L.lib.bqa_attention_forward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep_bits.data_ptr(), None, B, S, A, 0.1, 7, 5,
                                            st), "fwd")
        L.check(L.lib.bqa_attention_backward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep_bits.data_ptr(), dctx.data_ptr(),
                                             dqkv.data_ptr(), dsum.data_ptr(), None, B, S, A, 0.1, 7, 5, None, st), "bwd")
    torch.cuda.synchronize()
if "rows" in which:
    z, dy = rnd(T, H), rnd(T, H)
    gam, bet = torch.randn(H, device=dev), torch.randn(H, device=dev)
    y, dz, dzm = torch.empty_like(z), torch.empty_like(z), torch.empty_like(z)
    dg, db, dbias = (torch.zeros(H, device=dev) for _ in range(3))
    st = L.current_stream_ptr()
    for _ in range(2):
        L.check(L.lib.bqa_layernorm_forward(z.data_ptr(), gam.data_ptr(), bet.data_ptr(), y.data_ptr(), T, H, 1e-12,
                                            st), "ln")
        L.check(L.lib.bqa_layernorm_backward(dy.data_ptr(), z.data_ptr(), gam.data_ptr(), dz.data_ptr(), dzm.data_ptr(),
                                             dg.data_ptr(), db.data_ptr(), dbias.data_ptr(), T, H, 1e-12, 0.0, 0, 0.1,
                                             3, 9, st), "lnb")
        X = rnd(T, I)
        cs = torch.zeros(I, device=dev)
        L.check(L.lib.bqa_colsum_bf16(X.data_ptr(), I, cs.data_ptr(), T, I, st), "colsum")
    torch.cuda.synchronize()
print("done")
