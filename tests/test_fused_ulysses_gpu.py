"""Fused Ulysses attention block (K6 / K7) on ONE GPU: ``P`` logical sequence-parallel ranks run stage by stage, every
"peer" buffer on this device.  The column-scatter GEMM epilogue, the row-scatter attention epilogues and the head /
sequence bookkeeping must reproduce the ordinary attention block on the concatenated sequence, forward and backward."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _plain_block(x, wqkv, wo, cos, sin, Hq, Hk, D):
    from shuffle_exchange_b200.ops import functional as F
    qkv = F.linear(x, wqkv)
    qkv = F.rope(qkv, None, cos, sin, Hq + Hk, D)
    ctx = F.flash_attention_qkv(qkv.view(1, x.shape[0], -1), Hq, Hk, D, causal=True)
    return F.linear(ctx.reshape(x.shape[0], Hq * D), wo)


@pytest.mark.parametrize("P,D", [(2, 128), (4, 128), (2, 64)])
def test_emulated_block_matches_plain_block(P, D):
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from shuffle_exchange_b200.ops import functional as F
    from shuffle_exchange_b200.sequence.fused_ulysses import UlyssesPeers, ulysses_attention_emulated
    dev = torch.device("cuda")
    torch.manual_seed(P)
    H, Hq, Hk, Sl = 1024, 8, 4, 256
    S = Sl * P
    wqkv = (torch.randn((Hq + 2 * Hk) * D, H, device=dev) / math.sqrt(H)).to(torch.bfloat16)
    wo = (torch.randn(H, Hq * D, device=dev) / math.sqrt(Hq * D)).to(torch.bfloat16)
    cos, sin = F.rope_tables(S, D, 10000.0, dev)
    x = torch.randn(S, H, device=dev).to(torch.bfloat16)
    dy = torch.randn(S, H, device=dev).to(torch.bfloat16)
    # ---- plain block on the whole sequence (same kernels, no sequence parallelism)
    xr, wqr, wor = x.clone().requires_grad_(), wqkv.clone().requires_grad_(), wo.clone().requires_grad_()
    yr = _plain_block(xr, wqr, wor, cos, sin, Hq, Hk, D)
    yr.backward(dy)
    # ---- fused, P logical ranks
    up = UlyssesPeers(None, Sl, Hq, Hk, D, emulate_world=P)
    xs = [x[r * Sl:(r + 1) * Sl].contiguous() for r in range(P)]
    dys = [dy[r * Sl:(r + 1) * Sl].contiguous() for r in range(P)]
    outs, dxs, gwqkv, gwo = ulysses_attention_emulated(xs, wqkv, wo, cos, sin, up, True, dys)
    y = torch.cat(outs)
    dx = torch.cat(dxs)
    rel = lambda a, b: ((a.float() - b.float()).norm() / b.float().norm()).item()
    assert torch.allclose(y.float(), yr.float(), rtol=2e-2, atol=2e-2), (y.float() - yr.float()).abs().max()
    assert rel(dx, xr.grad) < 1e-2, rel(dx, xr.grad)
    assert rel(gwqkv, wqr.grad) < 1e-2, rel(gwqkv, wqr.grad)
    assert rel(gwo, wor.grad) < 1e-2, rel(gwo, wor.grad)
    # fp32 statement of the forward, independent of the kernels
    xf, wqf, wof = x.float(), wqkv.float(), wo.float()
    qkv = F.rope_ref(xf @ wqf.t(), torch.arange(S, device=dev), cos, sin, Hq + Hk, D)
    q = qkv[:, :Hq * D].view(1, S, Hq, D)
    k = qkv[:, Hq * D:(Hq + Hk) * D].view(1, S, Hk, D)
    v = qkv[:, (Hq + Hk) * D:].view(1, S, Hk, D)
    ctx = F.attention_ref(q, k, v, causal=True).reshape(S, Hq * D)
    want = ctx.float() @ wof.t()
    assert rel(y, want) < 2e-2, rel(y, want)
