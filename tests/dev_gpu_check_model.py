"""Developer check (GPU box): attention / row kernels / full step vs the oracle.
Each case runs in its own subprocess (timeout) -> gpurun_out/model_check.json."""
import ctypes
import json
import math
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def relerr(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def maxerr(a, b):
    return (a.float() - b.float()).abs().max().item()


def attn_case(B, S, A, p, use_mask, bwd=True):
    import torch
    from __graft_entry__ import load_package
    from oracle.dropout_hash import keep_mask
    pkg = load_package()
    L = pkg._lib
    dev = "cuda"
    torch.manual_seed(1)
    H = A * 64
    qkv = (torch.randn(B, S, 3 * H, device=dev)).to(torch.bfloat16)
    mask = None
    if use_mask:
        lens = torch.randint(S // 2, S + 1, (B,), device=dev)
        mask = (torch.arange(S, device=dev)[None, :] < lens[:, None])
    ctx = torch.zeros(B, S, H, device=dev, dtype=torch.bfloat16)
    lse2 = torch.zeros(B, A, S, device=dev)
    keep_bits = torch.zeros(B, A, S, L.lib.bqa_attention_keep_words(S), dtype=torch.int32, device=dev)
    st = L.current_stream_ptr()
    m8 = mask.to(torch.uint8).contiguous() if mask is not None else None
    L.check(L.lib.bqa_attention_forward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep_bits.data_ptr(), L.ptr(m8), B, S, A, p, 77, 5,
                                        st), "attn fwd")
    torch.cuda.synchronize()
    # reference
    x = qkv.float().requires_grad_(True)
    q, k, v = [t.view(B, S, A, 64).transpose(1, 2) for t in x.split(H, dim=-1)]
    sc = (q @ k.transpose(-1, -2)) / 8.0
    if mask is not None:
        sc = sc.masked_fill(~mask[:, None, None, :], float("-inf"))
    pr = torch.softmax(sc, -1)
    lse_ref = torch.logsumexp(sc, -1) * 1.4426950408889634
    if p > 0:
        keep = keep_mask(77, 5, p, pr.numel(), device=dev).view(pr.shape)
        pr = torch.where(keep, pr / (1 - p), torch.zeros_like(pr))
    o = (pr @ v).transpose(1, 2).reshape(B, S, H)
    out = {"fwd_rel": relerr(ctx, o), "fwd_max": maxerr(ctx, o), "lse_max": maxerr(lse2, lse_ref)}
    ok = out["fwd_rel"] < 1.5e-2 and out["lse_max"] < 1e-2
    if bwd:
        dctx = torch.randn(B, S, H, device=dev).to(torch.bfloat16)
        o.backward(dctx.float())
        dqkv = torch.zeros(B, S, 3 * H, device=dev, dtype=torch.bfloat16)
        dsum = torch.zeros(B, A, S, device=dev)
        L.check(L.lib.bqa_attention_backward(qkv.data_ptr(), ctx.data_ptr(), lse2.data_ptr(), keep_bits.data_ptr(), dctx.data_ptr(),
                                             dqkv.data_ptr(), dsum.data_ptr(), L.ptr(m8), B, S, A, p, 77, 5, None, st),
                "attn bwd")
        torch.cuda.synchronize()
        g = x.grad
        for i, n in enumerate("qkv"):
            out[f"d{n}_rel"] = relerr(dqkv[..., i * H:(i + 1) * H], g[..., i * H:(i + 1) * H])
            ok = ok and out[f"d{n}_rel"] < 2.5e-2
    out["ok"] = bool(ok)
    return out


def rows_case():
    import torch
    from __graft_entry__ import load_package
    pkg = load_package()
    L = pkg._lib
    dev = "cuda"
    torch.manual_seed(2)
    out = {}
    ok = True
    for (T, H) in ((1000, 768), (515, 1024), (64, 128)):
        z = torch.randn(T, H, device=dev).to(torch.bfloat16)
        g = torch.randn(H, device=dev)
        b = torch.randn(H, device=dev)
        y = torch.empty_like(z)
        st = L.current_stream_ptr()
        L.check(L.lib.bqa_layernorm_forward(z.data_ptr(), g.data_ptr(), b.data_ptr(), y.data_ptr(), T, H, 1e-12, st),
                "ln fwd")
        zf = z.float().requires_grad_(True)
        gf = g.clone().requires_grad_(True)
        bf = b.clone().requires_grad_(True)
        ref = torch.nn.functional.layer_norm(zf, (H,), gf, bf, 1e-12)
        out[f"ln_fwd_{T}x{H}"] = relerr(y, ref)
        ok = ok and out[f"ln_fwd_{T}x{H}"] < 5e-3
        dy = torch.randn(T, H, device=dev).to(torch.bfloat16)
        ref.backward(dy.float())
        dz = torch.empty_like(z)
        dzm = torch.empty_like(z)
        dg = torch.zeros(H, device=dev)
        db = torch.zeros(H, device=dev)
        dbias = torch.zeros(H, device=dev)
        L.check(L.lib.bqa_layernorm_backward(dy.data_ptr(), z.data_ptr(), g.data_ptr(), dz.data_ptr(), dzm.data_ptr(),
                                             dg.data_ptr(), db.data_ptr(), dbias.data_ptr(), T, H, 1e-12, 0.0, 0, 0.0,
                                             0, 0, st), "ln bwd")
        torch.cuda.synchronize()
        out[f"ln_bwd_dz_{T}x{H}"] = relerr(dz, zf.grad)
        out[f"ln_bwd_dg_{T}x{H}"] = relerr(dg, gf.grad)
        out[f"ln_bwd_db_{T}x{H}"] = relerr(db, bf.grad)
        out[f"ln_bwd_dbias_{T}x{H}"] = relerr(dbias, dz.float().sum(0))
        out[f"ln_bwd_dzm_{T}x{H}"] = maxerr(dzm, dz)
        ok = ok and out[f"ln_bwd_dz_{T}x{H}"] < 6e-3 and out[f"ln_bwd_dg_{T}x{H}"] < 3e-3 and \
            out[f"ln_bwd_db_{T}x{H}"] < 3e-3 and out[f"ln_bwd_dbias_{T}x{H}"] < 1e-3 and out[f"ln_bwd_dzm_{T}x{H}"] == 0
        X = torch.randn(T, 3 * H, device=dev).to(torch.bfloat16)
        cs = torch.zeros(3 * H, device=dev)
        L.check(L.lib.bqa_colsum_bf16(X.data_ptr(), 3 * H, cs.data_ptr(), T, 3 * H, st), "colsum")
        torch.cuda.synchronize()
        out[f"colsum_{T}x{3*H}"] = relerr(cs, X.float().sum(0))
        ok = ok and out[f"colsum_{T}x{3*H}"] < 1e-4
    # decode
    B, S, C = 7, 300, 5
    s = torch.randn(B, S, device=dev)
    e = torch.randn(B, S, device=dev)
    s[0, 10] = s[0, 200] = 9.0  # tie -> first index
    c = torch.randn(B, C, device=dev)
    si = torch.zeros(B, dtype=torch.int64, device=dev)
    ei = torch.zeros_like(si)
    ci = torch.zeros_like(si)
    sm = torch.zeros(B, device=dev)
    em = torch.zeros(B, device=dev)
    cp = torch.zeros(B, device=dev)
    sc = torch.zeros(B, device=dev)
    L.check(L.lib.bqa_decode(s.data_ptr(), e.data_ptr(), c.data_ptr(), si.data_ptr(), ei.data_ptr(), ci.data_ptr(),
                             sm.data_ptr(), em.data_ptr(), cp.data_ptr(), sc.data_ptr(), B, S, C,
                             L.current_stream_ptr()), "decode")
    torch.cuda.synchronize()
    from oracle.bertqa_oracle import decode
    d = decode({"start_class": s.cpu(), "end_class": e.cpu(), "cls": c.cpu()})
    out["decode_start"] = bool((si.cpu() == d["start_ids"]).all())
    out["decode_end"] = bool((ei.cpu() == d["end_ids"]).all())
    out["decode_cls"] = bool((ci.cpu() == d["cls_ids"]).all())
    out["decode_score_exact"] = bool((sc.cpu() == d["scores"]).all())
    out["decode_prob_err"] = maxerr(cp.cpu(), d["cls_probas"])
    ok = ok and out["decode_start"] and out["decode_end"] and out["decode_cls"] and out["decode_score_exact"]
    out["ok"] = bool(ok)
    return out


def model_case(shape, B, S, p_drop, use_mask, timing=False, check=True):
    import torch
    from __graft_entry__ import load_package
    from oracle import bertqa_oracle as O
    pkg = load_package()
    L = pkg._lib
    E = pkg.engine
    dev = "cuda"
    cfgo = O.OracleConfig(num_labels=5, hidden_dropout=p_drop, attn_dropout=p_drop, **shape)
    cfg = E.make_config(shape, hidden_dropout=p_drop, attn_dropout=p_drop)
    eng = E.Engine(cfg, dev)
    P = O.init_params(cfgo, 1234, device=dev)
    # make LN params / biases non-trivial so that their gradients are exercised
    g = torch.Generator(device="cpu").manual_seed(5)
    for k in P:
        if k.endswith("LayerNorm.weight"):
            P[k] = (1.0 + 0.1 * torch.randn(P[k].shape, generator=g)).to(dev)
        elif k.endswith(".bias") and k.startswith("transformer."):
            P[k] = (0.02 * torch.randn(P[k].shape, generator=g)).to(dev)
    for k, v in eng.param_views().items():
        v.copy_(P[k])
    eng.sync_weights()
    inputs, labels = O.dummy_batch(B, S, vocab=shape["vocab"], device=dev)
    if use_mask:
        lens = torch.randint(S // 2, S + 1, (B,), device=dev)
        m = torch.arange(S, device=dev)[None, :] < lens[:, None]
        inputs["input_ids"] = inputs["input_ids"] * m
        inputs["attention_mask"] = m
        labels["start_class"][0] = -1
    lco = O.LossConfig()
    lc = E.make_loss_config()
    out = {}
    seed, step = 4242, 3
    outs = eng.forward(inputs["input_ids"], inputs["token_type_ids"], inputs["attention_mask"] if use_mask else None,
                       training=True, seed=seed, step=step)
    losses = eng.loss(outs, labels, lc, with_grad=True).clone()
    eng.backward()
    torch.cuda.synchronize()
    ok = True
    if check:
        drop = O.DropSpec(seed, step) if p_drop > 0 else None
        oin = dict(inputs)
        if not use_mask:
            oin["attention_mask"] = None
        total, parts, grads, oout = O.loss_and_grads(P, cfgo, oin, labels, lco, drop=drop, bf16=True)
        names = ["start_class", "end_class", "cls", "start_reg", "end_reg"]
        for n, t in zip(names, outs):
            out["out_" + n] = relerr(t, oout[n])
            ok = ok and out["out_" + n] < 2e-2
        ref_l = [parts[k].item() for k in ("start_class", "end_class", "start_reg", "end_reg", "cls")] + [total.item()]
        out["losses"] = losses.tolist()
        out["losses_ref"] = ref_l
        ok = ok and all(abs(a - b) <= 2e-3 * max(1.0, abs(b)) for a, b in zip(losses.tolist(), ref_l))
        gv = eng.grad_views()
        worst = []
        for k in grads:
            r = relerr(gv[k], grads[k])
            worst.append((r, k, grads[k].norm().item()))
        worst.sort(reverse=True)
        out["grad_worst"] = [(round(r, 5), k, n) for r, k, n in worst[:8]]
        out["grad_median"] = sorted(w[0] for w in worst)[len(worst) // 2]
        gflat_ref = torch.cat([grads[k].flatten() for k in grads])
        gflat = torch.cat([gv[k].flatten() for k in grads])
        out["grad_global_rel"] = relerr(gflat, gflat_ref)
        ok = ok and out["grad_global_rel"] < 3e-2 and worst[0][0] < 8e-2
    if timing:
        for _ in range(3):
            outs = eng.forward(inputs["input_ids"], inputs["token_type_ids"], None, training=True, seed=seed, step=step)
            eng.loss(outs, labels, lc, with_grad=True)
            eng.backward()
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        iters = 5
        tf = tb = 0.0
        for _ in range(iters):
            ev[0].record()
            outs = eng.forward(inputs["input_ids"], inputs["token_type_ids"], None, training=True, seed=seed, step=step)
            eng.loss(outs, labels, lc, with_grad=True)
            ev[1].record()
            eng.backward()
            ev[2].record()
            torch.cuda.synchronize()
            tf += ev[0].elapsed_time(ev[1])
            tb += ev[1].elapsed_time(ev[2])
        out["fwd_ms"] = tf / iters
        out["bwd_ms"] = tb / iters
        out["samples_per_s_fwd_bwd"] = B / ((tf + tb) / iters / 1e3)
    out["ok"] = bool(ok)
    return out


TINY = dict(num_layers=2, hidden=128, heads=2, intermediate=512, vocab=1000, max_pos=512, type_vocab=2)
BASE = dict(num_layers=12, hidden=768, heads=12, intermediate=3072, vocab=30522, max_pos=512, type_vocab=2)

CASES = {
    "attn_s128": lambda: attn_case(2, 128, 2, 0.0, False),
    "attn_s384": lambda: attn_case(2, 384, 12, 0.0, False),
    "attn_s512": lambda: attn_case(2, 512, 2, 0.0, False),
    "attn_s200_mask": lambda: attn_case(3, 200, 2, 0.0, True),
    "attn_s384_drop": lambda: attn_case(2, 384, 4, 0.1, False),
    "attn_s50": lambda: attn_case(2, 50, 1, 0.0, True),
    "rows": rows_case,
    "tiny_nodrop": lambda: model_case(TINY, 4, 128, 0.0, False),
    "tiny_drop": lambda: model_case(TINY, 4, 128, 0.1, False),
    "tiny_mask_s200": lambda: model_case(TINY, 3, 200, 0.0, True),
    "base_b4_s384": lambda: model_case(BASE, 4, 384, 0.0, False),
    "base_b32_time": lambda: model_case(BASE, 32, 384, 0.1, False, timing=True, check=False),
}


def main():
    if len(sys.argv) > 1 and sys.argv[1] != "--all":
        res = CASES[sys.argv[1]]()
        res["name"] = sys.argv[1]
        print("RESULT " + json.dumps(res))
        return
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    results = []
    names = sys.argv[2:] if len(sys.argv) > 2 else list(CASES)
    for name in names:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), name], capture_output=True, text=True,
                               timeout=300, cwd=ROOT)
            line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
            if line:
                res = json.loads(line[-1][7:])
            else:
                res = {"name": name, "ok": False, "rc": r.returncode, "stdout": r.stdout[-1500:],
                       "stderr": r.stderr[-3000:]}
        except subprocess.TimeoutExpired:
            res = {"name": name, "ok": False, "timeout": True}
        res["wall_s"] = round(time.time() - t0, 1)
        print(json.dumps(res)[:2500], flush=True)
        results.append(res)
        with open(os.path.join(ROOT, "gpurun_out", "model_check.json"), "w") as f:
            json.dump(results, f, indent=1)
    print("SUMMARY", sum(1 for r in results if r.get("ok")), "/", len(results), "ok")


if __name__ == "__main__":
    main()
