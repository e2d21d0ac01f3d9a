"""CUDA-event timing of the GEMM variants at the BASELINE shapes (L2 flushed between launches).
BQA_LIB_PATH=<other libbqa.so> python tools/time_gemm.py  times another build on the same box."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from __graft_entry__ import load_package
pkg = load_package()
from _ab import maybe_swap_library
maybe_swap_library(pkg)
L = pkg._lib
dev = "cuda"
torch.manual_seed(0)
T, H, I = 12288, 768, 3072
rnd = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)
x, w1, b1 = rnd(T, H), rnd(I, H), torch.randn(I, device=dev)
u, g = torch.empty(T, I, device=dev, dtype=torch.bfloat16), torch.empty(T, I, device=dev, dtype=torch.bfloat16)
w2, res = rnd(H, I), rnd(T, H)
z2 = torch.empty(T, H, device=dev, dtype=torch.bfloat16)
dgm = rnd(T, H)
du = torch.empty(T, I, device=dev, dtype=torch.bfloat16)
dw = torch.zeros(I, H, device=dev)
bH = b1[:H].contiguous()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
cases = {
    "ffn1 fwd  gelu (2 outputs)": lambda: L.gemm_bf16(x, w1, u, epilogue=L.EPI_GELU_FWD, C2=g, bias=b1),
    "ffn2 fwd  bias+dropout+residual": lambda: L.gemm_bf16(g, w2, z2, bias=bH, residual=res, dropout_p=0.1, drop_seed=1, drop_site=2),
    "ffn2 dgrad x gelu'": lambda: L.gemm_bf16(dgm, w2, du, b_mn_major=True, epilogue=L.EPI_GELU_BWD, aux=u),
    "ffn1 dgrad +residual": lambda: L.gemm_bf16(du, w1, z2, b_mn_major=True, residual=res),
    "ffn1 wgrad split-K": lambda: L.gemm_bf16(du, x, dw, a_mn_major=True, b_mn_major=True, epilogue=L.EPI_F32_ADD, k_splits=0),
}
# the attention-side projections of the layer (K = 768 everywhere; N = 2304 / 768)
wq, bq = rnd(3 * H, H), torch.randn(3 * H, device=dev)
qkv = torch.empty(T, 3 * H, device=dev, dtype=torch.bfloat16)
dqkv = rnd(T, 3 * H)
wo = rnd(H, H)
ctx = rnd(T, H)
dsum = torch.zeros(T // 384 * 12 * 384, device=dev)
dwq, dwo, dw2 = torch.zeros(3 * H, H, device=dev), torch.zeros(H, H, device=dev), torch.zeros(H, I, device=dev)
cases.update({
    "qkv fwd   bias": lambda: L.gemm_bf16(x, wq, qkv, bias=bq),
    "ao fwd    bias+dropout+residual": lambda: L.gemm_bf16(ctx, wo, z2, bias=bH, residual=res, dropout_p=0.1, drop_seed=1, drop_site=3),
    "ao dgrad  + rowsum(dctx*ctx)": lambda: L.gemm_bf16(dgm, wo, z2, b_mn_major=True, dot_with=ctx, dot_out=dsum, dot_rows=384),
    "qkv dgrad +residual": lambda: L.gemm_bf16(dqkv, wq, z2, b_mn_major=True, residual=res),
    "qkv wgrad split-K": lambda: L.gemm_bf16(dqkv, x, dwq, a_mn_major=True, b_mn_major=True, epilogue=L.EPI_F32_ADD, k_splits=0),
    "ao wgrad split-K": lambda: L.gemm_bf16(dgm, ctx, dwo, a_mn_major=True, b_mn_major=True, epilogue=L.EPI_F32_ADD, k_splits=0),
    "ffn2 wgrad split-K": lambda: L.gemm_bf16(dgm, g, dw2, a_mn_major=True, b_mn_major=True, epilogue=L.EPI_F32_ADD, k_splits=0),
})
FLOP = {"ffn1": 2.0 * T * H * I, "ffn2": 2.0 * T * H * I, "qkv": 2.0 * T * H * 3 * H, "ao": 2.0 * T * H * H}
L.gemm_bf16(x, w1, u, epilogue=L.EPI_GELU_FWD, C2=g, bias=b1)
print("library:", L.LIB_PATH)
for name, fn in cases.items():
    ts = []
    for _ in range(25):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print(f"  {name:34s} median {ts[12]:6.1f} us   min {ts[0]:6.1f} us   {FLOP[name.split()[0]] / ts[12] / 1e6:6.0f} TFLOP/s")
