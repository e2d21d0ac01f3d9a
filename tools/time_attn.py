"""CUDA-event timing of the attention kernels at the BASELINE shape (B=32, S=384, A=12, dropout 0.1).
BQA_LIB_PATH=<other libbqa.so> python tools/time_attn.py  times another build on the same box."""
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
B, S, A = 32, 384, 12
H = A * 64
qkv = torch.randn(B, S, 3 * H, device=dev).to(torch.bfloat16)
ctx = torch.zeros(B, S, H, device=dev, dtype=torch.bfloat16)
lse2 = torch.zeros(B, A, S, device=dev)
keep = torch.zeros(B, A, S, L.lib.bqa_attention_keep_words(S), dtype=torch.int32, device=dev)
dctx = torch.randn(B, S, H, device=dev).to(torch.bfloat16)
dqkv = torch.zeros(B, S, 3 * H, device=dev, dtype=torch.bfloat16)
dsum = torch.zeros(B, A, S, device=dev)
mask = torch.ones(B, S, dtype=torch.uint8, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
st = L.current_stream_ptr()


def fwd(m, p):
    L.check(L.lib.bqa_attention_forward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep.data_ptr(), L.ptr(m), B, S, A,
                                        p, 7, 5, st), "fwd")


def bwd(m, p):
    L.check(L.lib.bqa_attention_backward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep.data_ptr(), dctx.data_ptr(),
                                         dqkv.data_ptr(), dsum.data_ptr(), L.ptr(m), B, S, A, p, 7, 5, None, st), "bwd")


print("library:", L.LIB_PATH)
for name, fn in (("forward (keep bits + kernel), p=0.1", lambda: fwd(None, 0.1)),
                 ("forward, p=0", lambda: fwd(None, 0.0)),
                 ("forward, p=0, key mask", lambda: fwd(mask, 0.0)),
                 ("backward (dsum + dq + dkv), p=0.1", lambda: bwd(None, 0.1)),
                 ("backward, p=0.1, key mask", lambda: bwd(mask, 0.1))):
    fwd(None, 0.1)
    ts = []
    for _ in range(25):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print(f"  {name:40s} median {ts[12]:6.1f} us   min {ts[0]:6.1f} us")


# S = 512 (BASELINE configs[3] / [4]): six 96-key chunks, the last one partial (32 keys)
def S512():
    B2, S2, A2 = 16, 512, 16
    H2 = A2 * 64
    qkv2 = torch.randn(B2, S2, 3 * H2, device=dev).to(torch.bfloat16)
    ctx2 = torch.zeros(B2, S2, H2, device=dev, dtype=torch.bfloat16)
    lse22 = torch.zeros(B2, A2, S2, device=dev)
    keep2 = torch.zeros(B2, A2, S2, L.lib.bqa_attention_keep_words(S2), dtype=torch.int32, device=dev)
    ts = []
    for _ in range(20):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        L.check(L.lib.bqa_attention_forward(qkv2.data_ptr(), ctx2.data_ptr(), lse22.data_ptr(), keep2.data_ptr(), None, B2, S2, A2,
                                            0.0, 7, 5, st), "fwd512")
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    fl = 4.0 * B2 * A2 * S2 * S2 * 64
    print(f"  forward S=512 B=16 A=16, p=0           median {ts[10]:6.1f} us   {fl / ts[10] / 1e6:.0f} TFLOP/s")


S512()
