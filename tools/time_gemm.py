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
    print(f"  {name:34s} median {ts[12]:6.1f} us   min {ts[0]:6.1f} us")
