"""CUDA-event timing of the row kernels at the BASELINE shapes (T = 12288, H = 768)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from __graft_entry__ import load_package
pkg = load_package()
L = pkg._lib
dev = "cuda"
T, H, I = 12288, 768, 3072
rnd = lambda *s: (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)
z, dy = rnd(T, H), rnd(T, H)
gam, bet = torch.randn(H, device=dev), torch.randn(H, device=dev)
y, dz, dzm = torch.empty_like(z), torch.empty_like(z), torch.empty_like(z)
dg, db, dbias = (torch.zeros(H, device=dev) for _ in range(3))
big = rnd(T, I)
csum = torch.zeros(I, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
st = L.current_stream_ptr()


def timeit(name, fn, n=20):
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print(f"{name}: median {ts[len(ts)//2]:.1f} us  min {ts[0]:.1f} us")


timeit("ln_fwd", lambda: L.check(L.lib.bqa_layernorm_forward(z.data_ptr(), gam.data_ptr(), bet.data_ptr(), y.data_ptr(), T, H, 1e-12, st), "f"))
timeit("ln_bwd(dropout 0.1)", lambda: L.check(L.lib.bqa_layernorm_backward(dy.data_ptr(), z.data_ptr(), gam.data_ptr(), dz.data_ptr(), dzm.data_ptr(), dg.data_ptr(), db.data_ptr(), dbias.data_ptr(), T, H, 1e-12, 0.0, 0, 0.1, 3, 99, st), "b"))
timeit("colsum[T,3072]", lambda: L.check(L.lib.bqa_colsum_bf16(big.data_ptr(), I, csum.data_ptr(), T, I, st), "c"))
