"""Developer check (GPU box): tcgen05 GEMM variants vs torch fp32 matmul of the same bf16 inputs.

Each case runs in its own subprocess with a timeout so that a trapped kernel cannot take the
other cases down.  Results: gpurun_out/gemm_check.json.  Not part of the test-suite; the
pytest versions live in tests/test_gemm_gpu.py.
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = {
    # name: (M, N, K, mode, extras)
    "tn_1tile": (128, 256, 64, "tn", {}),
    "tn_k4": (128, 256, 256, "tn", {}),
    "tn_multi": (512, 768, 768, "tn", {"bias": True}),
    "tn_mtail": (1000, 768, 768, "tn", {"bias": True, "residual": True}),
    "tn_ntail": (256, 320, 128, "tn", {"bias": True}),
    "tn_big_qkv": (12288, 2304, 768, "tn", {"bias": True, "time": True}),
    "tn_gelu": (1024, 3072, 768, "tn_gelu", {"bias": True}),
    "tn_dropout": (1024, 768, 768, "tn", {"bias": True, "residual": True, "dropout": 0.1}),
    "nn_1tile": (128, 256, 64, "nn", {}),
    "nn_multi": (1024, 768, 3072, "nn", {"residual": True}),
    "nn_gelu_bwd": (1024, 3072, 768, "nn_gelu_bwd", {}),
    "nn_big": (12288, 768, 3072, "nn", {"time": True}),
    "nt_1tile": (128, 256, 64, "nt", {}),
    "nt_multi": (768, 768, 1024, "nt", {"k_splits": 4}),
    "nt_ktail": (768, 768, 1000, "nt", {"k_splits": 3}),
    "nt_big": (3072, 768, 12288, "nt", {"k_splits": 8, "time": True}),
    "tn_big_ffn1": (12288, 3072, 768, "tn_gelu", {"bias": True, "time": True}),
    "tn_big_ffn2": (12288, 768, 3072, "tn", {"bias": True, "residual": True, "time": True}),
}


def run_case(name):
    import torch
    from __graft_entry__ import load_package
    pkg = load_package()
    L = pkg._lib
    from oracle.dropout_hash import keep_mask
    M, N, K, mode, ex = CASES[name]
    torch.manual_seed(0)
    dev = "cuda"
    out = {"name": name, "M": M, "N": N, "K": K, "mode": mode}

    def rnd(*s):
        return (torch.randn(*s, device=dev) * 0.5).to(torch.bfloat16)

    bias = (torch.randn(N, device=dev) * 0.5) if ex.get("bias") else None
    residual = rnd(M, N) if ex.get("residual") else None
    p = ex.get("dropout", 0.0)
    if mode in ("tn", "tn_gelu"):
        A, B = rnd(M, K), rnd(N, K)
        ref = A.float() @ B.float().t()
        kw = {}
    elif mode in ("nn", "nn_gelu_bwd"):
        A, B = rnd(M, K), rnd(K, N)
        ref = A.float() @ B.float()
        kw = {"b_mn_major": True}
    elif mode == "nt":
        A, B = rnd(K, M), rnd(K, N)
        ref = A.float().t() @ B.float()
        kw = {"a_mn_major": True, "b_mn_major": True}
    if bias is not None:
        ref = ref + bias
    C2 = None
    aux = None
    if mode == "tn_gelu":
        C = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        C2 = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        epi = L.EPI_GELU_FWD
    elif mode == "nn_gelu_bwd":
        aux = rnd(M, N)
        C = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        epi = L.EPI_GELU_BWD
        u = aux.float().requires_grad_(True)
        g = torch.nn.functional.gelu(u)
        g.backward(torch.ones_like(g))
        ref = ref * u.grad
    elif mode == "nt":
        C = torch.randn(M, N, device=dev)
        ref = ref + C
        epi = L.EPI_F32_ADD
    else:
        C = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        epi = L.EPI_LINEAR
        if p > 0:
            keep = keep_mask(1234, 7, p, M * N, device=dev).view(M, N)
            ref = torch.where(keep, ref / (1 - p), torch.zeros_like(ref))
            out["keep_rate"] = keep.float().mean().item()
        if residual is not None:
            ref = ref + residual.float()

    def call():
        L.gemm_bf16(A, B, C, epilogue=epi, C2=C2, bias=bias, residual=residual, aux=aux,
                    k_splits=ex.get("k_splits", 1), dropout_p=p, drop_seed=1234, drop_site=7, **kw)

    call()
    torch.cuda.synchronize()
    got = C.float()
    err = (got - ref).abs()
    denom = ref.abs().max().item() + 1e-6
    out["max_abs_err"] = err.max().item()
    out["rel_to_max"] = err.max().item() / denom
    out["mean_abs_err"] = err.mean().item()
    out["ref_absmax"] = denom
    # bf16 output rounding: ~2^-8 relative to each value; accumulate differences are tiny
    tol = 2e-2 * denom if epi != L.EPI_F32_ADD else 1e-3 * denom
    out["ok"] = bool(err.max().item() <= tol)
    if mode == "tn_gelu":  # C = gelu'(u), C2 = gelu(u)
        u = ref.clone().requires_grad_(True)
        g_ref = torch.nn.functional.gelu(u)
        g_ref.sum().backward()
        e2 = (C2.float() - g_ref.detach()).abs().max().item()
        e1 = (C.float() - u.grad).abs().max().item()
        out["gelu_max_abs_err"] = e2
        out["gelu_grad_max_abs_err"] = e1
        out["max_abs_err"] = e1
        out["ok"] = bool(e2 <= tol and e1 <= 8e-3)
    if not out["ok"]:
        bad = (err > tol).nonzero()
        out["n_bad"] = int(bad.shape[0])
        out["first_bad"] = bad[:8].tolist()
        out["bad_rows_hist"] = torch.bincount((bad[:, 0] % 128), minlength=128).tolist() if bad.numel() else []
        out["bad_cols_hist"] = torch.bincount((bad[:, 1] % 64), minlength=64).tolist() if bad.numel() else []
        out["sample_got"] = got[:2, :8].tolist()
        out["sample_ref"] = ref[:2, :8].tolist()
    if ex.get("time"):
        if mode == "nt":
            C.zero_()
        for _ in range(3):
            call()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 20
        e0.record()
        for _ in range(iters):
            call()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        out["ms"] = ms
        out["tflops"] = 2.0 * M * N * K / ms / 1e9
        # torch/cuBLAS for comparison
        if mode in ("tn", "tn_gelu"):
            f = lambda: torch.matmul(A, B.t())
        elif mode.startswith("nn"):
            f = lambda: torch.matmul(A, B)
        else:
            f = lambda: torch.matmul(A.t(), B)
        for _ in range(3):
            f()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            f()
        e1.record()
        torch.cuda.synchronize()
        out["cublas_ms"] = e0.elapsed_time(e1) / iters
        out["cublas_tflops"] = 2.0 * M * N * K / out["cublas_ms"] / 1e9
    print("RESULT " + json.dumps(out))


def main():
    if len(sys.argv) > 1 and sys.argv[1] != "--all":
        run_case(sys.argv[1])
        return
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    results = []
    for name in CASES:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), name], capture_output=True, text=True,
                               timeout=120, cwd=ROOT)
            line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
            if line:
                res = json.loads(line[-1][7:])
            else:
                res = {"name": name, "ok": False, "rc": r.returncode, "stdout": r.stdout[-1500:],
                       "stderr": r.stderr[-2500:]}
        except subprocess.TimeoutExpired:
            res = {"name": name, "ok": False, "timeout": True}
        res["wall_s"] = round(time.time() - t0, 1)
        print(json.dumps(res)[:1200], flush=True)
        results.append(res)
        with open(os.path.join(ROOT, "gpurun_out", "gemm_check.json"), "w") as f:
            json.dump(results, f, indent=1)
    print("SUMMARY", sum(1 for r in results if r.get("ok")), "/", len(results), "ok")


if __name__ == "__main__":
    main()
