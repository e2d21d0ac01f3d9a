"""Accumulated phase times of thread 0 (query row 0) of every CTA of the attention forward kernel, and how many
CTAs were co-resident per SM (trace build: tools/build_trace.sh)."""
import ctypes, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = ctypes.CDLL(os.path.join(ROOT, "tools/_trace/libbqa_trace.so"))
vp = ctypes.c_void_p
lib.bqa_attention_forward.argtypes = [vp, vp, vp, vp, vp, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_float,
                                      ctypes.c_uint64, ctypes.c_uint32, vp]
lib.bqa_attention_keep_words.argtypes = [ctypes.c_int32]
lib.bqa_trace_read.argtypes = [vp]
B, S, A = 32, 384, 12
H = A * 64
dev = "cuda"
qkv = torch.randn(B, S, 3 * H, device=dev).to(torch.bfloat16)
ctx = torch.zeros(B, S, H, device=dev, dtype=torch.bfloat16)
lse2 = torch.zeros(B, A, S, device=dev)
keep = torch.zeros(B, A, S, lib.bqa_attention_keep_words(S), dtype=torch.int32, device=dev)
st = torch.cuda.current_stream().cuda_stream
names = {1: "wait scores", 2: "TMEM load + chunk max", 3: "rescale", 4: "exp, P stores, arrive", 5: "store release",
         6: "wait O", 7: "epilogue", 8: "chunk prologue (keep words)"}
for p in (0.1, 0.0):
    for _ in range(3):
        assert lib.bqa_attention_forward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep.data_ptr(), None, B, S, A, p, 7, 5, st) == 0
    torch.cuda.synchronize()
    buf = np.zeros(4096 * 16, dtype=np.uint64)
    assert lib.bqa_trace_read(buf.ctypes.data) == 0
    n_cta = 296
    t = buf.reshape(4096, 16)[:n_cta].astype(np.int64)
    total = t[:, 9] - t[:, 0]
    tiles = 1152 / n_cta
    print(f"--- forward, dropout {p}: cycles per CTA {total.mean():.0f}; per tile {total.mean() / tiles:.0f}")
    for k in range(1, 9):
        print(f"  {names[k]:>28s}: {t[:, k].mean() / tiles:8.0f} per tile  {t[:, k].mean() / total.mean() * 100:5.1f}%")
    smid, w0, w1 = t[:, 10], t[:, 11], t[:, 12]
    span = w1.max() - w0.min()
    overlap = 0
    for sm in np.unique(smid):
        idx = np.where(smid == sm)[0]
        if len(idx) >= 2:
            a_, b_ = idx[0], idx[1]
            overlap += max(0, min(w1[a_], w1[b_]) - max(w0[a_], w0[b_])) / max(1, max(w1[a_], w1[b_]) - min(w0[a_], w0[b_]))
    print(f"  SMs used {len(np.unique(smid))}, CTAs per SM max {np.bincount(smid).max()}, kernel span {span / 1e3:.1f} us, "
          f"mean CTA life {(w1 - w0).mean() / 1e3:.1f} us, mean pairwise overlap {overlap / len(np.unique(smid)):.2f}")
