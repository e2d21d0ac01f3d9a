"""Accumulated phase times of the row threads of the attention backward kernels (trace build)."""
import ctypes, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = ctypes.CDLL(os.path.join(ROOT, "tools/_trace/libbqa_trace" + (sys.argv[1] if len(sys.argv) > 1 else "") + ".so"))
vp = ctypes.c_void_p
lib.bqa_attention_forward.argtypes = [vp, vp, vp, vp, vp, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_float,
                                      ctypes.c_uint64, ctypes.c_uint32, vp]
lib.bqa_attention_backward.argtypes = [vp] * 8 + [ctypes.c_int32] * 3 + [ctypes.c_float, ctypes.c_uint64, ctypes.c_uint32, vp, vp]
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
dctx = torch.randn(B, S, H, device=dev).to(torch.bfloat16)
dqkv = torch.zeros(B, S, 3 * H, device=dev, dtype=torch.bfloat16)
dsum = torch.zeros(B, A, S, device=dev)
names = {1: "wait scores", 2: "tmem load", 3: "math", 4: "wait operand buffers", 5: "stores+fence+arrive", 6: "drain",
         7: "tile prologue", 8: "other"}
for p in (0.1, 0.0):
    for _ in range(2):
        assert lib.bqa_attention_forward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep.data_ptr(), None, B, S, A, p, 7, 5, st) == 0
        assert lib.bqa_attention_backward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep.data_ptr(), dctx.data_ptr(),
                                          dqkv.data_ptr(), dsum.data_ptr(), None, B, S, A, p, 7, 5, None, st) == 0
    torch.cuda.synchronize()
    buf = np.zeros(4096 * 16, dtype=np.uint64)
    assert lib.bqa_trace_read(buf.ctypes.data) == 0
    t = buf.reshape(4096, 16)[:148].astype(np.int64)
    total = t[:, 9] - t[:, 0]
    chunks = 1152 * 3 / 148
    print(f"--- dkv kernel (last launched), dropout {p}: cycles per CTA {total.mean():.0f}; per chunk {total.mean() / chunks:.0f}")
    for k in range(1, 9):
        print(f"  {names[k]:>22s}: {t[:, k].mean():9.0f} per CTA  {t[:, k].mean() / chunks:8.0f} per chunk  {t[:, k].mean() / total.mean() * 100:5.1f}%")
    allchunks = 1152 * 3 / 148
    for k, nm in ((13, "score-MMA warp: wait operands (TMA)"), (14, "score-MMA warp: wait S/dP columns free"), (15, "score-MMA warp: issue + commits")):
        print(f"  {nm:>44s}: {t[:, k].mean() / allchunks:8.0f} per chunk (all chunks)  {t[:, k].mean() / total.mean() * 100:5.1f}%")
