"""GPU numerics of the extended tcgen05 GEMM (csrc/gemm_ext.cu): gated B operand (K2), grouped ragged-M / ragged-K
problems (K4/K5) and fp8 e4m3 operands (K14), each against a plain PyTorch fp32 statement of the same op."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda")


def _close(got, ref, K, what=""):
    """bf16 output of an fp32-accumulated product: relative error of one bf16 rounding plus an absolute term that
    covers fp32 summation-order noise for |ref| << sqrt(K)."""
    got, ref = got.float(), ref.float()
    err = (got - ref).abs()
    tol = 8e-3 * ref.abs() + 2e-3 * math.sqrt(K)
    bad = err > tol
    assert not bool(bad.any()), f"{what}: {int(bad.sum())}/{got.numel()} off, max err {err.max().item():.4f}"


def _mm(a, b, layout):
    fa, fb = a.float(), b.float()
    return fa @ fb.t() if layout == 0 else (fa @ fb if layout == 1 else fa.t() @ fb)


@pytest.mark.parametrize("layout,M,N,K", [(0, 512, 768, 1024), (1, 640, 1024, 512), (2, 768, 512, 1088),
                                          (0, 300, 520, 136), (1, 8192, 4096, 4096)])
def test_plain_matches_fp32(layout, M, N, K):
    from shuffle_exchange_b200.ops.gemm_ext import gemm_gated
    dev = _dev()
    torch.manual_seed(layout * 7 + M)
    a = torch.randn((K, M) if layout == 2 else (M, K), device=dev).to(torch.bfloat16)
    b = torch.randn((N, K) if layout == 0 else (K, N), device=dev).to(torch.bfloat16)
    out = gemm_gated(a, b, layout, None)
    _close(out, _mm(a, b, layout), K, f"layout {layout}")
    # fp32 accumulate-into output (main-grad style)
    acc = torch.randn(M, N, device=dev)
    want = acc + _mm(a, b, layout)
    gemm_gated(a, b, layout, None, out=acc, accumulate=True)
    _close(acc, want, K, "accumulate")


@pytest.mark.parametrize("layout", [0, 1])
def test_gated_operand_waits_for_arrival_flags(layout):
    """The weight matrix sits inside a larger 'gathered' buffer.  The GEMM is launched while only the first flags are
    set; the rest are published later by stream memory operations on another stream (behind a sleep), as a copy-engine
    gather would.  The result must be exact and the GEMM must not finish before the last flag."""
    from shuffle_exchange_b200.native import lib as nlib
    from shuffle_exchange_b200.ops.gemm_ext import Gate, gemm_gated, stream_write_u32
    dev = _dev()
    torch.manual_seed(3)
    M, N, K = 1024, 1536, 1024
    lead = 4096 + 64                                   # elements in front of the matrix inside the flat buffer
    wshape = (N, K) if layout == 0 else (K, N)
    flat = torch.randn(lead + wshape[0] * wshape[1] + 333, device=dev).to(torch.bfloat16)
    w = flat[lead: lead + wshape[0] * wshape[1]].view(wshape)
    a = torch.randn(M, K, device=dev).to(torch.bfloat16)
    chunk = 1 << 18
    nchunks = (flat.numel() * 2 + chunk - 1) // chunk
    flags = torch.zeros(nchunks, dtype=torch.int32, device=dev)
    value = 5
    gate = Gate(flags, value, chunk, flat.data_ptr())
    want = _mm(a, w, layout)
    side = torch.cuda.Stream()
    done_gemm, done_flags = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    start.record()
    with torch.cuda.stream(side):
        side.wait_event(start)
        for c in range(nchunks):
            torch.cuda._sleep(400_000)                 # ~0.2 ms between arrivals; launched before the GEMM
            stream_write_u32(flags, c, value)
        done_flags.record(side)
    out = gemm_gated(a, w, layout, gate)
    done_gemm.record()
    torch.cuda.synchronize()
    assert nlib.hang_info("gemmx") is None
    _close(out, want, K, "gated")
    assert start.elapsed_time(done_gemm) >= 0.95 * start.elapsed_time(done_flags), "GEMM finished before its operand"
    # flags already set (steady state: next gather uses a larger sequence number, older ones still pass)
    out2 = gemm_gated(a, w, layout, Gate(flags, value - 1, chunk, flat.data_ptr()))
    _close(out2, want, K, "gated, flags preset")


def _make_groups(dev, counts, K, N, seed):
    from shuffle_exchange_b200.ops.gemm_ext import aligned_offsets
    torch.manual_seed(seed)
    cnt = torch.tensor(counts, device=dev, dtype=torch.int64)
    off = aligned_offsets(cnt)
    R = int(off[-1])
    x = torch.zeros(R, K, device=dev, dtype=torch.bfloat16)
    for g, c in enumerate(counts):
        x[int(off[g]): int(off[g]) + c] = torch.randn(c, K, device=dev).to(torch.bfloat16)
    w = (torch.randn(len(counts), N, K, device=dev) * 0.5).to(torch.bfloat16)
    return x, w, off, R


@pytest.mark.parametrize("counts,K,N", [([700, 0, 256, 1300], 512, 768), ([5000, 3000], 4096, 1024), ([1] * 8, 256, 256)])
def test_grouped_forward_dgrad_wgrad(counts, K, N):
    from shuffle_exchange_b200.native import lib as nlib
    from shuffle_exchange_b200.ops.gemm_ext import grouped_gemm_nn, grouped_gemm_nt, grouped_gemm_tn
    dev = _dev()
    x, w, off, R = _make_groups(dev, counts, K, N, 11)
    y = grouped_gemm_nt(x, w, off)
    o = off.tolist()
    for g in range(len(counts)):
        lo, hi = o[g], o[g + 1]
        if hi > lo:
            _close(y[lo:hi], x[lo:hi].float() @ w[g].float().t(), K, f"fwd g{g}")
    dy = torch.zeros(R, N, device=dev, dtype=torch.bfloat16)
    for g, c in enumerate(counts):
        dy[o[g]: o[g] + c] = torch.randn(c, N, device=dev).to(torch.bfloat16)
    dx = grouped_gemm_nn(dy, w, off)
    for g in range(len(counts)):
        lo, hi = o[g], o[g + 1]
        if hi > lo:
            _close(dx[lo:hi], dy[lo:hi].float() @ w[g].float(), N, f"dgrad g{g}")
    dw = torch.full_like(w, 7.0)
    grouped_gemm_tn(dy, x, off, dw)
    for g in range(len(counts)):
        lo, hi = o[g], o[g + 1]
        _close(dw[g], dy[lo:hi].float().t() @ x[lo:hi].float(), max(hi - lo, 1), f"wgrad g{g}")
    # accumulate into an fp32 main-grad stack
    acc = torch.randn(len(counts), N, K, device=dev)
    want = acc.clone()
    for g in range(len(counts)):
        want[g] += dy[o[g]:o[g + 1]].float().t() @ x[o[g]:o[g + 1]].float()
    grouped_gemm_tn(dy, x, off, acc, accumulate=True)
    _close(acc, want, max(counts), "wgrad accumulate")
    assert nlib.hang_info("gemmx") is None


def test_fp8_gemm_rowwise_scaled():
    from shuffle_exchange_b200.ops.gemm_ext import fp8_gemm_nt, fp8_rowwise_quant
    dev = _dev()
    torch.manual_seed(5)
    M, N, K = 1024, 1280, 2048
    a = (torch.randn(M, K, device=dev) * torch.rand(M, 1, device=dev) * 4).to(torch.bfloat16)
    b = (torch.randn(N, K, device=dev) * 0.1).to(torch.bfloat16)
    qa, sa = fp8_rowwise_quant(a)
    qb, sb = fp8_rowwise_quant(b)
    # the quantiser: same result as PyTorch's e4m3 cast of x / scale
    ref_scale = a.float().abs().amax(1) / 448.0
    assert torch.allclose(sa, ref_scale, rtol=1e-6)
    ref_q = (a.float() / ref_scale[:, None]).to(torch.float8_e4m3fn).view(torch.uint8)
    assert (qa != ref_q).float().mean().item() < 1e-3          # ties on the division only
    out = fp8_gemm_nt(qa, sa, qb, sb)
    deq = lambda q, s: q.view(torch.float8_e4m3fn).float() * s[:, None]
    _close(out, deq(qa, sa) @ deq(qb, sb).t(), K, "fp8 vs dequantised operands")
    full = a.float() @ b.float().t()
    rel = ((out.float() - full).norm() / full.norm()).item()
    assert rel < 0.06, rel                                      # e4m3 (3 mantissa bits) quantisation bound


def test_grouped_linear_autograd_bf16_and_fp8():
    from shuffle_exchange_b200.ops.gemm_ext import grouped_linear
    dev = _dev()
    counts, K, N = [900, 300, 0, 2000], 1024, 512
    x, w, off, R = _make_groups(dev, counts, K, N, 23)
    o = off.tolist()
    xr = x.clone().float().requires_grad_()
    wr = w.clone().float().requires_grad_()
    yr = torch.zeros(R, N, device=dev)
    parts = [xr[o[g]:o[g + 1]] @ wr[g].t() for g in range(len(counts))]
    yr = torch.cat(parts)
    gy = torch.zeros(R, N, device=dev)
    for g, c in enumerate(counts):
        gy[o[g]: o[g] + c] = torch.randn(c, N, device=dev)
    gy = gy.to(torch.bfloat16)
    yr.backward(gy.float())
    xg = x.clone().requires_grad_()
    wg = w.clone().requires_grad_()
    y = grouped_linear(xg, wg, off)
    y.backward(gy)
    _close(y, yr, K, "y")
    _close(xg.grad, xr.grad, N, "dx")
    _close(wg.grad, wr.grad, max(counts), "dw")
    # fp8 forward: within the quantisation bound of the bf16 result; backward stays bf16
    x8 = x.clone().requires_grad_()
    w8 = w.clone().requires_grad_()
    y8 = grouped_linear(x8, w8, off, fp8=True)
    rel = ((y8.float() - yr).norm() / yr.norm()).item()
    assert rel < 0.06, rel
    y8.backward(gy)
    _close(x8.grad, xr.grad, N, "dx (fp8 fwd)")
