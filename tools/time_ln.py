"""LayerNorm forward / backward kernels alone at the step's shapes, operands rotated through more than L2:
    python tools/time_ln.py            (BQA_LIB_PATH=<other build> for A/B)
Prints us per launch (CUDA events over back-to-back launches) and the algorithmic GB/s."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch
from __graft_entry__ import load_package
pkg = load_package()
from _ab import maybe_swap_library
maybe_swap_library(pkg)
L = pkg._lib
dev = "cuda"
st = L.current_stream_ptr()
for T, H in ((12288, 768), (8192, 1024)):
    nset = 10
    z = [torch.randn(T, H, device=dev).to(torch.bfloat16) for _ in range(nset)]
    dy = [torch.randn(T, H, device=dev).to(torch.bfloat16) for _ in range(nset)]
    out = [torch.empty_like(z[0]) for _ in range(nset)]
    out2 = [torch.empty_like(z[0]) for _ in range(nset)]
    g, b = torch.randn(H, device=dev), torch.randn(H, device=dev)
    dg, db, dbias = (torch.zeros(H, device=dev) for _ in range(3))

    def fwd(i):
        L.check(L.lib.bqa_layernorm_forward(z[i % nset].data_ptr(), g.data_ptr(), b.data_ptr(), out[i % nset].data_ptr(), T, H, 1e-12, st), "ln")

    def bwd(i):
        k = i % nset
        L.check(L.lib.bqa_layernorm_backward(dy[k].data_ptr(), z[k].data_ptr(), g.data_ptr(), out[k].data_ptr(), out2[k].data_ptr(),
                                             dg.data_ptr(), db.data_ptr(), dbias.data_ptr(), T, H, 1e-12, 0.0, 0, 0.1, 9, 4321, st), "ln bwd")

    for name, fn, nbytes in (("ln_fwd", fwd, 2 * T * H * 2), ("ln_bwd", bwd, 4 * T * H * 2)):
        for i in range(20):
            fn(i)
        res = []
        for rep in range(3):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            n = 200
            for i in range(n):
                fn(i)
            e1.record()
            torch.cuda.synchronize()
            res.append(e0.elapsed_time(e1) / n * 1e3)
        us = min(res)
        print(f"{name} T={T} H={H}: {us:.2f} us/launch  {nbytes / us / 1e3:.0f} GB/s  (legs {' '.join(f'{r:.2f}' for r in res)})  lib={L.LIB_PATH}")
