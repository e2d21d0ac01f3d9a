"""Reference point only (not a product path): cuBLAS bf16 GEMM through torch.matmul at the step's GEMM shapes,
L2 flushed between launches, next to nothing else on the GPU.  Tells how far the hand-written kernel is from the
library at the SAME shapes (MEASURED_PEAKS.json's figure is for 8192^3)."""
import torch
dev = "cuda"
T, H, I = 12288, 768, 3072
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
cases = [("qkv fwd   [T,H]x[H,3H]", T, 3 * H, H), ("ao fwd    [T,H]x[H,H]", T, H, H), ("ffn1 fwd  [T,H]x[H,I]", T, I, H),
         ("ffn2 fwd  [T,I]x[I,H]", T, H, I), ("ffn1 wgrad [I,T]x[T,H]", I, H, T), ("qkv wgrad [3H,T]x[T,H]", 3 * H, H, T),
         ("ao wgrad  [H,T]x[T,H]", H, H, T), ("ffn2 dgrad [T,H]x[H,I]", T, I, H), ("qkv dgrad [T,3H]x[3H,H]", T, H, 3 * H)]
for name, M, N, K in cases:
    a = torch.randn(M, K, device=dev).to(torch.bfloat16)
    b = torch.randn(K, N, device=dev).to(torch.bfloat16)
    c = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        torch.matmul(a, b, out=c)
    ts = []
    for _ in range(20):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b, out=c); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print(f"{name:28s} median {ts[10]:6.1f} us  {2.0 * M * N * K / ts[10] / 1e6:7.0f} TFLOP/s")
