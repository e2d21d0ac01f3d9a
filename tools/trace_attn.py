"""Phase timeline of attn_fwd CTAs (debug build with -DBQA_ATTN_TRACE, tools/_trace/libbqa_trace.so)."""
import ctypes, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = ctypes.CDLL(os.path.join(ROOT, "tools/_trace/libbqa_trace.so"))
vp = ctypes.c_void_p
lib.bqa_attention_forward.argtypes = [vp, vp, vp, vp, vp, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_float,
                                      ctypes.c_uint64, ctypes.c_uint32, vp]
lib.bqa_trace_read.argtypes = [vp]
B, S, A = 32, 384, 12
H = A * 64
dev = "cuda"
qkv = torch.randn(B, S, 3 * H, device=dev).to(torch.bfloat16)
ctx = torch.zeros(B, S, H, device=dev, dtype=torch.bfloat16)
lse2 = torch.zeros(B, A, S, device=dev)
keep = torch.zeros(B, A, S, 12, dtype=torch.int32, device=dev)
st = torch.cuda.current_stream().cuda_stream
lib.bqa_attention_backward.argtypes = [vp] * 8 + [ctypes.c_int32] * 3 + [ctypes.c_float, ctypes.c_uint64, ctypes.c_uint32, vp, vp]
dctx = torch.randn(B, S, H, device=dev).to(torch.bfloat16)
dqkv = torch.zeros(B, S, 3 * H, device=dev, dtype=torch.bfloat16)
dsum = torch.zeros(B, A, S, device=dev)


def report(title, names):
    torch.cuda.synchronize()
    buf = np.zeros(4096 * 16, dtype=np.uint64)
    assert lib.bqa_trace_read(buf.ctypes.data) == 0
    t = buf.reshape(4096, 16)[:1152].astype(np.int64)
    print("---", title)
    for i in range(9):
        if names[i] is None or names[i + 1] is None:
            continue
        x = t[:, i + 1] - t[:, i]
        print(f"  {names[i]:>18s} -> {names[i+1]:<18s} {x.mean():9.0f} {np.median(x):9.0f}")
    for (a_, b_) in ((2, 4), (4, 5), (5, 6)):
        x = t[:, b_] - t[:, a_]
        print(f"  slot{a_} -> slot{b_}: {x.mean():9.0f}")
    print("  total per CTA", (t[:, 9] - t[:, 0]).mean())


for p in (0.1, 0.0):
    rc = lib.bqa_attention_forward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep.data_ptr(), None, B, S, A, p, 7, 5, st)
    # backward = dsum kernel, dq kernel, dkv kernel on one stream: the trace holds the LAST kernel (dkv)
    rc = lib.bqa_attention_backward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep.data_ptr(), dctx.data_ptr(),
                                    dqkv.data_ptr(), dsum.data_ptr(), None, B, S, A, p, 7, 5, None, st)
    assert rc == 0
    report(f"dkv kernel, dropout {p}", ["start", "ctrl_loads_issued", "kv_landed", None, "S0_ready", "chunk1_start", "chunks_done", "out_ready", "stored", "end"])
for p in (0.1, 0.0):
    for _ in range(3):
        rc = lib.bqa_attention_forward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep.data_ptr(), None, B, S, A, p, 7, 5, st)
        assert rc == 0
    torch.cuda.synchronize()
    buf = np.zeros(4096 * 16, dtype=np.uint64)
    assert lib.bqa_trace_read(buf.ctypes.data) == 0
    t = buf.reshape(4096, 16)[:1152].astype(np.int64)
    names = ["start", "ctrl_loads_issued", "qk_landed", "v_landed", "S_ready(row0)", "pass1_done", "pass2_done", "O_ready", "stored", "end"]
    d = t[:, 1:10] - t[:, 0:9]
    print(f"--- dropout {p}: per-CTA phase durations (cycles): mean / median / p90")
    for i in range(9):
        x = d[:, i]
        print(f"  {names[i]:>18s} -> {names[i+1]:<18s} {x.mean():9.0f} {np.median(x):9.0f} {np.percentile(x,90):9.0f}")
    tot = t[:, 9] - t[:, 0]
    print("  total per CTA", tot.mean(), np.median(tot))
    # per-SM: gaps between consecutive CTAs
    smid = t[:, 15]
    gaps, busy = [], []
    span = []
    for sm in np.unique(smid):
        idx = np.where(smid == sm)[0]
        o = idx[np.argsort(t[idx, 0])]
        g = t[o[1:], 0] - t[o[:-1], 9]
        gaps += g.tolist()
        span.append(t[o[-1], 9] - t[o[0], 0])
        busy.append((t[o, 9] - t[o, 0]).sum())
    print("  CTAs/SM", 1152 / len(np.unique(smid)), "gap between CTAs on an SM: mean", np.mean(gaps), "median", np.median(gaps))
    print("  per-SM span mean", np.mean(span), "busy mean", np.mean(busy), "kernel span", t[:, 9].max() - t[:, 0].min())
