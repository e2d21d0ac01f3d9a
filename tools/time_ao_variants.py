"""Attention-output projection (M = 12288, N = K = 768: two waves of tiles) with each part of its epilogue switched on in
turn, cold (L2 flushed) and back to back: shows what bias / dropout / residual cost when the last epilogue is exposed."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, ROOT + "/tools")
import torch
from __graft_entry__ import load_package
pkg = load_package()
L = pkg._lib
dev = "cuda"
torch.manual_seed(0)
T, H, I = 12288, 768, 3072
rnd = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)
ctx, wo, res = rnd(T, H), rnd(H, H), rnd(T, H)
bH = torch.randn(H, device=dev)
z2 = torch.empty(T, H, device=dev, dtype=torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
cases = {
 "ao fwd plain": lambda: L.gemm_bf16(ctx, wo, z2),
 "ao fwd bias": lambda: L.gemm_bf16(ctx, wo, z2, bias=bH),
 "ao fwd bias+res": lambda: L.gemm_bf16(ctx, wo, z2, bias=bH, residual=res),
 "ao fwd bias+drop": lambda: L.gemm_bf16(ctx, wo, z2, bias=bH, dropout_p=0.1, drop_seed=1, drop_site=3),
 "ao fwd bias+drop+res": lambda: L.gemm_bf16(ctx, wo, z2, bias=bH, residual=res, dropout_p=0.1, drop_seed=1, drop_site=3),
}
for M in (12288, 6144, 3072):
  for name, fn0 in cases.items():
    if M != 12288 and name != "ao fwd plain": continue
    fn = fn0 if M == 12288 else (lambda: L.gemm_bf16(ctx[:M], wo, z2[:M]))
    fn()
    ts = []
    for _ in range(25):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print(f"M={M} {name:28s} median {ts[12]:6.1f} us  min {ts[0]:6.1f}")
# back-to-back without flush (launch overhead / warm L2)
for name, fn in cases.items():
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): fn()
    e1.record(); torch.cuda.synchronize()
    print(f"back-to-back {name:28s} {e0.elapsed_time(e1) * 1e3 / 50:6.1f} us")
