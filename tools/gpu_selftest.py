"""Run every kernel check in its own subprocess (a hang or a fault in one kernel must not hide
the others) and write ``gpurun_out/selftest.json``.

    python tools/gpu_selftest.py            # all checks
    python tools/gpu_selftest.py gemm attn  # only checks whose name contains one of the words
    python tools/gpu_selftest.py --one NAME # (internal) run a single check in this process
"""
from __future__ import annotations

import json
import math
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _rel_err(a, b):
    import torch
    # element-wise bound: every element's error is measured against its OWN magnitude plus the tensor's rms (the
    # rounding error of a bf16 dot product scales with the summands, which the rms stands for) — one bad element in a
    # tensor of large values can no longer hide behind max|b|
    a, b = a.float(), b.float()
    scale = b.abs() + b.pow(2).mean().sqrt().clamp_min(1e-6)
    return ((a - b).abs() / scale).max().item()


def _time(fn, iters=10, warm=3):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


CHECKS = {}


def check(name):
    def deco(f):
        CHECKS[name] = f
        return f
    return deco


def _gemm_case(layout, M, N, K, out_dtype=None, accumulate=False):
    import torch
    from shuffle_exchange_b200.ops import functional as F
    torch.manual_seed(0)
    dev = "cuda"
    if layout == 0:
        a = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        b = torch.randn(N, K, device=dev, dtype=torch.bfloat16)
        ref = a.float() @ b.float().t()
    elif layout == 1:
        a = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        b = torch.randn(K, N, device=dev, dtype=torch.bfloat16)
        ref = a.float() @ b.float()
    else:
        a = torch.randn(K, M, device=dev, dtype=torch.bfloat16)
        b = torch.randn(K, N, device=dev, dtype=torch.bfloat16)
        ref = a.float().t() @ b.float()
    out = None
    if accumulate or out_dtype is not None:
        out = torch.randn(M, N, device=dev, dtype=out_dtype or torch.bfloat16)
        if accumulate:
            ref = ref + out.float()
    got = F.gemm(a, b, layout, out=out, accumulate=accumulate)
    torch.cuda.synchronize()
    return _rel_err(got, ref)


@check("gemm_nt_small")
def _():
    return {"err": _gemm_case(0, 128, 256, 64)}


@check("gemm_nt")
def _():
    return {"err": max(_gemm_case(0, 256, 512, 256), _gemm_case(0, 1024, 4096, 4096))}


@check("gemm_nn")
def _():
    return {"err": max(_gemm_case(1, 128, 256, 64), _gemm_case(1, 1024, 4096, 1024))}


@check("gemm_tn")
def _():
    return {"err": max(_gemm_case(2, 128, 256, 64), _gemm_case(2, 4096, 1024, 2048))}


@check("gemm_tails")
def _():
    errs = [_gemm_case(0, 200, 328, 136), _gemm_case(1, 200, 328, 136), _gemm_case(2, 200, 328, 136),
            _gemm_case(0, 64, 96, 64)]
    return {"err": max(errs)}


@check("gemm_fp32_acc")
def _():
    import torch
    return {"err": max(_gemm_case(2, 512, 768, 1024, out_dtype=torch.float32, accumulate=True),
                       _gemm_case(0, 512, 768, 1024, accumulate=True))}


@check("gemm_perf")
def _():
    import torch
    from shuffle_exchange_b200.ops import functional as F
    res = {}
    for (M, N, K) in [(8192, 8192, 8192), (8192, 6144, 4096), (8192, 28672, 4096), (8192, 4096, 14336)]:
        a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
        b = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        ms = _time(lambda: F.gemm(a, b, 0, out=out))
        ms_ref = _time(lambda: torch.matmul(a, b.t(), out=out))
        res[f"{M}x{N}x{K}"] = {"ms": ms, "tflops": 2 * M * N * K / ms / 1e9, "cublas_ms": ms_ref,
                               "cublas_tflops": 2 * M * N * K / ms_ref / 1e9}
    for layout, name in ((1, "nn"), (2, "tn")):
        M = N = K = 8192
        a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
        b = torch.randn(K, N, device="cuda", dtype=torch.bfloat16)
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        ms = _time(lambda: F.gemm(a, b, layout, out=out))
        res[f"{name}_8192"] = {"ms": ms, "tflops": 2 * M * N * K / ms / 1e9}
    return res


def _with_gemm_mode(mode, fn):
    from shuffle_exchange_b200.native import lib
    lib.call("sxb_gemm_set_mode", mode)
    try:
        return fn()
    finally:
        lib.call("sxb_gemm_set_mode", 0)


@check("gemm_pair")
def _():
    """CTA-pair (cta_group::2) kernel forced on: every layout, ragged edges, fp32 accumulate."""
    import torch
    from shuffle_exchange_b200.native import lib

    def run():
        errs = {"nt": _gemm_case(0, 1024, 4096, 4096), "nn": _gemm_case(1, 1024, 4096, 1024),
                "tn": _gemm_case(2, 4096, 1024, 2048), "nt_small": _gemm_case(0, 256, 256, 64),
                "tails": max(_gemm_case(0, 200, 328, 136), _gemm_case(1, 200, 328, 136), _gemm_case(2, 200, 328, 136)),
                "acc": max(_gemm_case(2, 512, 768, 1024, out_dtype=torch.float32, accumulate=True),
                           _gemm_case(0, 512, 768, 1024, accumulate=True)),
                "many_tiles": _gemm_case(0, 8192, 6144, 512)}
        return errs
    errs = _with_gemm_mode(2, run)
    return {"err": max(errs.values()), "cases": errs, "watchdog": lib.hang_info("gemm")}


@check("gemm_pair_perf")
def _():
    import torch
    from shuffle_exchange_b200.ops import functional as F
    res = {}
    shapes = [(0, 8192, 8192, 8192), (0, 8192, 6144, 4096), (0, 8192, 28672, 4096), (0, 8192, 4096, 14336),
              (0, 8192, 4096, 4096), (1, 8192, 4096, 28672), (1, 8192, 14336, 4096), (2, 28672, 4096, 8192),
              (2, 4096, 14336, 8192), (0, 8192, 128256, 4096)]
    for (layout, M, N, K) in shapes:
        if layout == 0:
            a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16); b = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
        elif layout == 1:
            a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16); b = torch.randn(K, N, device="cuda", dtype=torch.bfloat16)
        else:
            a = torch.randn(K, M, device="cuda", dtype=torch.bfloat16); b = torch.randn(K, N, device="cuda", dtype=torch.bfloat16)
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        fl = 2 * M * N * K / 1e9
        ms1 = _with_gemm_mode(1, lambda: _time(lambda: F.gemm(a, b, layout, out=out)))
        ms2 = _with_gemm_mode(2, lambda: _time(lambda: F.gemm(a, b, layout, out=out)))
        if layout == 0:
            msr = _time(lambda: torch.matmul(a, b.t(), out=out))
        elif layout == 1:
            msr = _time(lambda: torch.matmul(a, b, out=out))
        else:
            msr = _time(lambda: torch.matmul(a.t(), b, out=out))
        res[f"{'NT NN TN'.split()[layout]}_{M}x{N}x{K}"] = {"single_tflops": fl / ms1, "pair_tflops": fl / ms2,
                                                             "cublas_tflops": fl / msr}
        del a, b, out
    return res


@check("rmsnorm")
def _():
    import torch
    from shuffle_exchange_b200.ops import functional as F
    torch.manual_seed(0)
    errs = {}
    for H in (512, 4096, 8192):
        x = torch.randn(300, H, device="cuda", dtype=torch.bfloat16, requires_grad=True)
        r = torch.randn(300, H, device="cuda", dtype=torch.bfloat16, requires_grad=True)
        w = (1 + 0.1 * torch.randn(H, device="cuda")).to(torch.bfloat16).requires_grad_()
        y, h = F.rms_norm(x, w, 1e-5, r)
        gy, gh = torch.randn_like(y), torch.randn_like(h)
        (y.float() * gy.float()).sum().add((h.float() * gh.float()).sum()).backward()
        xr, rr, wr = (t.detach().float().requires_grad_() for t in (x, r, w))
        hr = xr + rr
        hr_q = hr + (hr.to(torch.bfloat16).float() - hr).detach()  # straight-through the bf16 rounding
        yr = hr_q * torch.rsqrt(hr_q.pow(2).mean(-1, keepdim=True) + 1e-5) * wr
        ((yr * gy.float()).sum() + (hr * gh.float()).sum()).backward()
        errs[H] = {"y": _rel_err(y, yr), "dx": _rel_err(x.grad, xr.grad), "dr": _rel_err(r.grad, rr.grad),
                   "dw": _rel_err(w.grad, wr.grad)}
        # no-residual variant
        x2 = torch.randn(64, H, device="cuda", dtype=torch.bfloat16, requires_grad=True)
        y2, _ = F.rms_norm(x2, w.detach(), 1e-5)
        y2.float().sum().backward()
        x2r = x2.detach().float().requires_grad_()
        (x2r * torch.rsqrt(x2r.pow(2).mean(-1, keepdim=True) + 1e-5) * w.detach().float()).sum().backward()
        errs[H]["dx_nores"] = _rel_err(x2.grad, x2r.grad)
    return {"err": max(v for d in errs.values() for v in d.values()), "detail": errs}


@check("layernorm")
def _():
    import torch
    from shuffle_exchange_b200.ops import functional as F
    torch.manual_seed(0)
    H = 1024
    x = torch.randn(200, H, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    w = (1 + 0.1 * torch.randn(H, device="cuda")).to(torch.bfloat16).requires_grad_()
    b = (0.1 * torch.randn(H, device="cuda")).to(torch.bfloat16).requires_grad_()
    y, _ = F.layer_norm(x, w, b, 1e-5)
    g = torch.randn_like(y)
    (y.float() * g.float()).sum().backward()
    xr, wr, br = (t.detach().float().requires_grad_() for t in (x, w, b))
    yr = torch.nn.functional.layer_norm(xr, (H,), wr, br, 1e-5)
    (yr * g.float()).sum().backward()
    return {"err": max(_rel_err(y, yr), _rel_err(x.grad, xr.grad), _rel_err(w.grad, wr.grad),
                       _rel_err(b.grad, br.grad))}


@check("rope")
def _():
    import torch
    from shuffle_exchange_b200.ops import functional as F
    torch.manual_seed(0)
    T, Hq, Hk, D = 512, 8, 2, 128
    C = (Hq + 2 * Hk) * D
    x = torch.randn(T, C, device="cuda", dtype=torch.bfloat16)
    cos, sin = F.rope_tables(1024, D, 500000.0, "cuda")
    pos = torch.randint(0, 1024, (T,), device="cuda")
    ref = F.rope_ref(x, pos, cos, sin, Hq + Hk, D)
    xin = x.clone().requires_grad_()
    y = F.rope(xin * 1.0, pos, cos, sin, Hq + Hk, D)
    g = torch.randn_like(y)
    y.backward(g)
    gref = F.rope_ref(g, pos, cos, sin, Hq + Hk, D, inverse=True)
    return {"err": max(_rel_err(y, ref), _rel_err(xin.grad, gref))}


@check("swiglu")
def _():
    import torch
    from shuffle_exchange_b200.ops import functional as F
    torch.manual_seed(0)
    x = torch.randn(333, 2 * 1024, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    y = F.swiglu(x)
    g = torch.randn_like(y)
    y.backward(g)
    xr = x.detach().float().requires_grad_()
    yr = torch.nn.functional.silu(xr[:, :1024]) * xr[:, 1024:]
    yr.backward(g.float())
    e1 = max(_rel_err(y, yr), _rel_err(x.grad, xr.grad))
    xb = torch.randn(100, 512, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    bias = torch.randn(512, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    yb = F.bias_act(xb, bias, "gelu")
    yb.backward(torch.ones_like(yb))
    xbr, br = xb.detach().float().requires_grad_(), bias.detach().float().requires_grad_()
    ybr = torch.nn.functional.gelu(xbr + br, approximate="tanh")
    ybr.sum().backward()
    e2 = max(_rel_err(yb, ybr), _rel_err(xb.grad, xbr.grad), _rel_err(bias.grad, br.grad))
    return {"err": max(e1, e2)}


@check("cross_entropy")
def _():
    import torch
    from shuffle_exchange_b200.ops import functional as F
    torch.manual_seed(0)
    T, V = 257, 32008
    logits = (3 * torch.randn(T, V, device="cuda")).to(torch.bfloat16)
    labels = torch.randint(0, V, (T,), device="cuda")
    labels[::7] = -100
    lr = logits.float().requires_grad_()
    ref = torch.nn.functional.cross_entropy(lr, labels, ignore_index=-100)
    ref.backward()
    lin = logits.clone().requires_grad_()
    loss = F.cross_entropy(lin * 1.0, labels)
    loss.backward()
    return {"err": max(abs(loss.item() - ref.item()) / abs(ref.item()), _rel_err(lin.grad, lr.grad))}


@check("embedding")
def _():
    import torch
    from shuffle_exchange_b200.ops import functional as F
    torch.manual_seed(0)
    table = torch.randn(1000, 256, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    ids = torch.randint(0, 1000, (4, 77), device="cuda")
    out = F.embedding(ids, table)
    g = torch.randn_like(out)
    out.backward(g)
    tr = table.detach().float().requires_grad_()
    outr = torch.nn.functional.embedding(ids, tr)
    outr.backward(g.float())
    return {"err": max(_rel_err(out, outr), _rel_err(table.grad, tr.grad))}


@check("adam")
def _():
    import torch
    from shuffle_exchange_b200.native import lib as n
    torch.manual_seed(0)
    N = 1_000_003
    N4 = N - N % 4
    p = torch.randn(N4, device="cuda")
    g = torch.randn(N4, device="cuda").to(torch.bfloat16)
    m = torch.zeros(N4, device="cuda")
    v = torch.zeros(N4, device="cuda")
    lp = torch.empty(N4, device="cuda", dtype=torch.bfloat16)
    pr = p.clone().requires_grad_()
    opt = torch.optim.AdamW([pr], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    for step in (1, 2, 3):
        pr.grad = g.float()
        opt.step()
        n.call("sxb_adam_flat", p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), lp.data_ptr(), N4,
               n.BF16, n.BF16, 1e-2, 0.9, 0.95, 1e-8, 0.1, step, 1, 1, 1.0, None, None, n.stream_ptr())
    torch.cuda.synchronize()
    err = max(_rel_err(p, pr.detach()), _rel_err(lp, pr.detach()))
    big = 1 << 28
    P = torch.randn(big, device="cuda"); G = torch.randn(big, device="cuda").to(torch.bfloat16)
    Mm = torch.zeros(big, device="cuda"); Vv = torch.zeros(big, device="cuda")
    LP = torch.empty(big, device="cuda", dtype=torch.bfloat16)
    ms_big = _time(lambda: n.call("sxb_adam_flat", P.data_ptr(), G.data_ptr(), Mm.data_ptr(), Vv.data_ptr(),
                                  LP.data_ptr(), big, n.BF16, n.BF16, 1e-2, 0.9, 0.95, 1e-8, 0.1, 4, 1, 1,
                                  1.0, None, None, n.stream_ptr()), iters=5)
    gbs = big * (4 * 3 * 2 + 2 + 2) / ms_big / 1e6
    return {"err": err, "adam_268M_ms": ms_big,
            "adam_GBps": gbs}


def _attn_case(B, S, Hq, Hk, D, causal, bwd=True):
    import torch
    from shuffle_exchange_b200.ops import functional as F
    torch.manual_seed(0)
    qkv = torch.randn(B, S, (Hq + 2 * Hk) * D, device="cuda", dtype=torch.bfloat16)
    qkv.requires_grad_()
    q = qkv[..., : Hq * D].view(B, S, Hq, D)
    k = qkv[..., Hq * D: (Hq + Hk) * D].view(B, S, Hk, D)
    v = qkv[..., (Hq + Hk) * D:].view(B, S, Hk, D)
    from shuffle_exchange_b200.native import lib as n
    out = F.flash_attention(q, k, v, causal)
    hi = n.hang_info("attn")
    if hi is not None:
        return {"hang_fwd": hi}
    qr = qkv.detach().float().requires_grad_()
    q2 = qr[..., : Hq * D].view(B, S, Hq, D)
    k2 = qr[..., Hq * D: (Hq + Hk) * D].view(B, S, Hk, D)
    v2 = qr[..., (Hq + Hk) * D:].view(B, S, Hk, D)
    ref = F.attention_ref(q2, k2, v2, causal)
    res = {"fwd": _rel_err(out, ref)}
    if bwd:
        g = torch.randn_like(out)
        out.backward(g)
        hi = n.hang_info("attn")
        if hi is not None:
            return {"hang_bwd": hi}
        ref.backward(g.float())
        dq, dqr = qkv.grad[..., : Hq * D], qr.grad[..., : Hq * D]
        dk, dkr = qkv.grad[..., Hq * D: (Hq + Hk) * D], qr.grad[..., Hq * D: (Hq + Hk) * D]
        dv, dvr = qkv.grad[..., (Hq + Hk) * D:], qr.grad[..., (Hq + Hk) * D:]
        res.update(dq=_rel_err(dq, dqr), dk=_rel_err(dk, dkr), dv=_rel_err(dv, dvr))
    return res


def _attn_sweep(cfgs, bwd):
    out = {}
    worst = 0.0
    for cfg in cfgs:
        print("CFG", cfg, flush=True)
        r = _attn_case(*cfg, bwd=bwd)
        print("  ->", r, flush=True)
        out[str(cfg)] = r
        for v in r.values():
            worst = max(worst, v if isinstance(v, float) else 1e9)
    return {"err": worst, "detail": out}


_ATTN_CFGS = [(1, 128, 1, 1, 128, False), (1, 256, 1, 1, 128, True), (2, 256, 1, 1, 128, True),
              (1, 512, 4, 2, 128, True), (2, 512, 4, 2, 128, True), (1, 1024, 2, 1, 128, False),
              (1, 300, 2, 1, 128, True), (1, 128, 1, 1, 64, False), (1, 384, 2, 2, 64, True)]


def _mk_attn_check(cfg, bwd):
    def f():
        r = _attn_case(*cfg, bwd=bwd)
        worst = max((v if isinstance(v, float) else 1e9) for v in r.values())
        return {"err": worst, "detail": r}
    return f


for _cfg in _ATTN_CFGS:
    _tag = "x".join(str(int(v)) for v in _cfg)
    CHECKS[f"attn_fwd_{_tag}"] = _mk_attn_check(_cfg, False)
for _cfg in _ATTN_CFGS:
    _tag = "x".join(str(int(v)) for v in _cfg)
    CHECKS[f"attn_bwd_{_tag}"] = _mk_attn_check(_cfg, True)


@check("attn_fused_qkv")
def _():
    """Fused-QKV entry point (gradient returned as one tensor) against the fp32 reference, GQA and MHA."""
    import torch
    from shuffle_exchange_b200.ops import functional as F
    worst, detail = 0.0, {}
    for (B, S, Hq, Hk, D) in ((2, 384, 4, 2, 128), (1, 300, 2, 2, 64)):
        torch.manual_seed(0)
        qkv = torch.randn(B, S, (Hq + 2 * Hk) * D, device="cuda", dtype=torch.bfloat16, requires_grad=True)
        out = F.flash_attention_qkv(qkv, Hq, Hk, D, True)
        g = torch.randn_like(out)
        out.backward(g)
        qr = qkv.detach().float().requires_grad_()
        q2, k2, v2 = F._split_qkv(qr, Hq, Hk, D)
        ref = F.attention_ref(q2, k2, v2, True)
        ref.backward(g.float())
        r = {"fwd": _rel_err(out, ref), "dqkv": _rel_err(qkv.grad, qr.grad)}
        detail[f"{B}x{S}x{Hq}x{Hk}x{D}"] = r
        worst = max(worst, *r.values())
    return {"err": worst, "detail": detail}


@check("attn_perf")
def _():
    import torch
    from shuffle_exchange_b200.ops import functional as F
    B, S, Hq, Hk, D = 1, 8192, 32, 8, 128
    qkv = torch.randn(B, S, (Hq + 2 * Hk) * D, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    q = qkv[..., : Hq * D].view(B, S, Hq, D)
    k = qkv[..., Hq * D: (Hq + Hk) * D].view(B, S, Hk, D)
    v = qkv[..., (Hq + Hk) * D:].view(B, S, Hk, D)
    flops = 4 * B * Hq * S * S * D / 2
    ms_f = _time(lambda: F.flash_attention(q.detach(), k.detach(), v.detach(), True))
    out = F.flash_attention(q, k, v, True)
    g = torch.randn_like(out)
    ms_b = _time(lambda: torch.autograd.grad(out, qkv, g, retain_graph=True), iters=5)
    qs, ks, vs = (t.detach().transpose(1, 2).contiguous().requires_grad_() for t in (q, k, v))
    sd = lambda: torch.nn.functional.scaled_dot_product_attention(qs, ks, vs, is_causal=True, enable_gqa=True)
    ms_sf = _time(sd)
    o2 = sd()
    g2 = torch.randn_like(o2)
    ms_sb = _time(lambda: torch.autograd.grad(o2, (qs, ks, vs), g2, retain_graph=True), iters=5)
    return {"fwd_ms": ms_f, "fwd_tflops": flops / ms_f / 1e9, "bwd_ms": ms_b,
            "bwd_tflops": 2.5 * flops / ms_b / 1e9, "sdpa_fwd_ms": ms_sf, "sdpa_bwd_ms": ms_sb}


@check("llama_tiny")
def _():
    import torch
    from shuffle_exchange_b200.models.llama import LlamaForCausalLM, llama_config, reference_forward
    from shuffle_exchange_b200.native import lib as n
    torch.manual_seed(0)
    out = {}
    for name, kw in (("d64", {}), ("d128", dict(hidden_size=512, num_attention_heads=4, num_key_value_heads=2))):
        cfg = llama_config("llama-tiny", **kw)
        m = LlamaForCausalLM(cfg).to("cuda", torch.bfloat16).train()
        ids = torch.randint(0, cfg.vocab_size, (2, 256), device="cuda")
        loss = m(ids, labels=ids)
        loss.backward()
        hi = n.hang_info("attn") or n.hang_info("gemm")
        if hi:
            return {"ok": False, "hang": hi}
        mr = LlamaForCausalLM(cfg).to("cuda", torch.float32)
        mr.load_state_dict({k: v.float() for k, v in m.state_dict().items()})
        mr.cfg.activation_checkpointing = False
        # autograd through the fp32 oracle
        params = {k: v.detach().clone().requires_grad_() for k, v in mr.state_dict().items()}
        mr.load_state_dict(params, strict=True)
        for k, v in mr.named_parameters():
            v.requires_grad_()
        import torch.nn.functional as TF
        from shuffle_exchange_b200.ops import functional as F
        # run the composite (CPU-path) implementation on GPU tensors by calling the reference pieces
        lref = reference_forward(mr, ids, ids)
        out[name] = {"loss": loss.item(), "ref": lref.item(), "rel": abs(loss.item() - lref.item()) / abs(lref.item())}
    return {"err": max(v["rel"] for v in out.values()), "detail": out}


@check("llama_layer_perf")
def _():
    import torch
    from shuffle_exchange_b200.models.llama import LlamaDecoderLayer, llama_config
    from shuffle_exchange_b200.ops import functional as F
    cfg = llama_config("llama3-8b")
    layer = LlamaDecoderLayer(cfg, 0, device="cuda", dtype=torch.bfloat16)
    layer.reset_parameters(0.02)
    B, S = 1, 8192
    h = torch.randn(B, S, cfg.hidden_size, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    res = torch.randn(B, S, cfg.hidden_size, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    pos = torch.arange(S, device="cuda").repeat(B)
    cos, sin = F.rope_tables(S, cfg.head_dim, cfg.rope_theta, "cuda")

    def fwd():
        with torch.no_grad():
            return layer(h, res, pos, cos, sin)

    def fwdbwd():
        o, r = layer(h, res, pos, cos, sin)
        (o.float().mean() + r.float().mean()).backward()
    ms_f = _time(fwd, iters=5)
    ms_fb = _time(fwdbwd, iters=5)
    return {"fwd_ms": ms_f, "fwd_bwd_ms": ms_fb, "est_step_ms_32_layers_recompute": 32 * (ms_f + ms_fb)}


@check("attn_long_vs_sdpa")
def _():
    import torch
    from shuffle_exchange_b200.native import lib as n
    from shuffle_exchange_b200.ops import functional as F
    torch.manual_seed(0)
    out = {}
    for (B, S, Hq, Hk, D) in [(1, 8192, 4, 2, 128), (2, 4096, 4, 1, 64)]:
        qkv = torch.randn(B, S, (Hq + 2 * Hk) * D, device="cuda", dtype=torch.bfloat16, requires_grad=True)
        q = qkv[..., : Hq * D].view(B, S, Hq, D)
        k = qkv[..., Hq * D: (Hq + Hk) * D].view(B, S, Hk, D)
        v = qkv[..., (Hq + Hk) * D:].view(B, S, Hk, D)
        o = F.flash_attention(q, k, v, True)
        g = torch.randn_like(o)
        (dqkv,) = torch.autograd.grad(o, qkv, g)
        hi = n.hang_info("attn")
        if hi:
            return {"ok": False, "hang": hi}
        q2, k2, v2 = (t.detach().float().transpose(1, 2).contiguous().requires_grad_() for t in (q, k, v))
        o2 = torch.nn.functional.scaled_dot_product_attention(q2, k2, v2, is_causal=True, enable_gqa=True)
        o2.backward(g.float().transpose(1, 2))
        ref = torch.cat([q2.grad.transpose(1, 2).reshape(B, S, -1), k2.grad.transpose(1, 2).reshape(B, S, -1),
                         v2.grad.transpose(1, 2).reshape(B, S, -1)], -1)
        out[f"{B}x{S}x{Hq}x{Hk}x{D}"] = {"fwd": _rel_err(o, o2.transpose(1, 2)), "bwd": _rel_err(dqkv, ref)}
    return {"err": max(v for d in out.values() for v in d.values()), "detail": out}


@check("llama_train_parity")
def _():
    """5 optimizer steps through the engine (bf16, ZeRO-3, fused AdamW in backward) vs the fp32
    PyTorch oracle with torch.optim.AdamW: loss curves must track."""
    import os
    import torch
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT="29577", RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
    import shuffle_exchange_b200 as sxb
    from shuffle_exchange_b200.models.llama import LlamaForCausalLM, llama_config, reference_forward
    torch.manual_seed(0)
    cfg = llama_config("llama-tiny", hidden_size=512, num_attention_heads=4, num_key_value_heads=2)
    model = LlamaForCausalLM(cfg, device="cuda", dtype=torch.bfloat16)
    ref = LlamaForCausalLM(cfg, device="cuda", dtype=torch.float32)
    ref.load_state_dict({k: v.float() for k, v in model.state_dict().items()})
    ref.cfg.activation_checkpointing = False
    hp = dict(lr=2e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    engine, _, _, _ = sxb.initialize(model=model, model_parameters=model.parameters(), config={
        "train_micro_batch_size_per_gpu": 2, "bf16": {"enabled": True}, "steps_per_print": 1 << 30,
        "optimizer": {"type": "AdamW", "params": dict(lr=hp["lr"], betas=list(hp["betas"]), eps=hp["eps"],
                                                        weight_decay=hp["weight_decay"])},
        "zero_optimization": {"stage": 3}})
    params = list(ref.parameters())
    opt = torch.optim.AdamW(params, **hp)
    g = torch.Generator(device="cpu").manual_seed(1)
    ours, theirs = [], []
    for step in range(6):
        ids = torch.randint(0, cfg.vocab_size, (2, 256), generator=g).cuda()
        loss = engine(ids, labels=ids)
        engine.backward(loss)
        engine.step()
        ours.append(float(loss))
        opt.zero_grad()
        # differentiable oracle: run reference_forward on live parameters
        sd_live = dict(ref.named_parameters())
        lref = _oracle_loss(ref, ids)
        lref.backward()
        opt.step()
        theirs.append(float(lref))
    err = max(abs(a - b) / abs(b) for a, b in zip(ours, theirs))
    return {"err": err, "ours": ours, "oracle": theirs}


def _oracle_loss(model, ids):
    """fp32 autograd version of models.llama.reference_forward (same math, live parameters)."""
    import torch
    from shuffle_exchange_b200.ops import functional as F
    cfg = model.cfg
    P = dict(model.named_parameters())
    B, S = ids.shape
    d, nq, nk, f = cfg.head_dim, cfg.num_attention_heads, cfg.num_key_value_heads, cfg.intermediate_size
    h = P["model.embed_tokens"][ids]
    inv = 1.0 / (cfg.rope_theta ** (torch.arange(0, d, 2, dtype=torch.float64, device=h.device) / d))
    ang = torch.outer(torch.arange(S, dtype=torch.float64, device=h.device), inv)
    cos, sin = ang.cos().float(), ang.sin().float()

    def rms(x, w):
        return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + cfg.rms_norm_eps) * w

    def rot(x):
        a, b = x[..., : d // 2], x[..., d // 2:]
        c, s = cos[None, :, None, :], sin[None, :, None, :]
        return torch.cat([a * c - b * s, b * c + a * s], -1)

    for i in range(cfg.num_hidden_layers):
        p = f"model.layers.{i}."
        x = rms(h, P[p + "input_layernorm"])
        qkv = x @ P[p + "wqkv"].t()
        q = rot(qkv[..., : nq * d].view(B, S, nq, d))
        k = rot(qkv[..., nq * d: (nq + nk) * d].view(B, S, nk, d))
        v = qkv[..., (nq + nk) * d:].view(B, S, nk, d)
        o = F.attention_ref(q, k, v, True)
        h = h + o.reshape(B, S, nq * d) @ P[p + "wo"].t()
        x = rms(h, P[p + "post_attention_layernorm"])
        gu = x @ P[p + "w_gate_up"].t()
        h = h + (torch.nn.functional.silu(gu[..., :f]) * gu[..., f:]) @ P[p + "w_down"].t()
    x = rms(h, P["model.norm"])
    logits = x @ P["lm_head"].t()
    tgt = torch.full_like(ids, -100)
    tgt[:, :-1] = ids[:, 1:]
    return torch.nn.functional.cross_entropy(logits.view(-1, logits.shape[-1]), tgt.view(-1), ignore_index=-100)


def main():
    if len(sys.argv) >= 3 and sys.argv[1] == "--one":
        name = sys.argv[2]
        import faulthandler
        faulthandler.dump_traceback_later(45, exit=False)
        import torch  # noqa
        try:
            res = CHECKS[name]()
            print("RESULT " + json.dumps({"ok": True, **res}, default=float))
        except Exception as e:  # noqa
            import traceback
            traceback.print_exc()
            print("RESULT " + json.dumps({"ok": False, "error": f"{type(e).__name__}: {e}"[:500]}))
        return
    words = [a for a in sys.argv[1:] if not a.startswith("-")]
    names = [n for n in CHECKS if not words or any(w in n for w in words)]
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    results = {}
    for name in names:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, __file__, "--one", name], capture_output=True, text=True,
                               timeout=60)
            line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
            if line:
                results[name] = json.loads(line[-1][7:])
            else:
                results[name] = {"ok": False, "error": "no result", "rc": r.returncode,
                                 "tail": (r.stdout + r.stderr)[-800:]}
            if not results[name].get("ok"):
                results[name]["tail"] = (r.stdout + r.stderr)[-1500:]
        except subprocess.TimeoutExpired as e:
            def _s(b):
                return (b.decode(errors="replace") if isinstance(b, bytes) else (b or ""))[-1800:]
            results[name] = {"ok": False, "error": "timeout (hang?)", "stdout": _s(e.stdout), "stderr": _s(e.stderr)}
        results[name]["secs"] = round(time.time() - t0, 1)
        err = results[name].get("err")
        tol = 2e-2
        if results[name].get("ok") and err is not None and not (err == err and err < tol):
            results[name]["ok"] = False
            results[name]["error"] = f"numerical error {err} >= {tol}"
        print(name, json.dumps(results[name], default=float)[:600], flush=True)
        with open(os.path.join(ROOT, "gpurun_out", "selftest.json"), "w") as f:
            json.dump(results, f, indent=1, default=float)
    bad = [n for n, r in results.items() if not r.get("ok")]
    print("FAILED:", bad)


if __name__ == "__main__":
    main()
