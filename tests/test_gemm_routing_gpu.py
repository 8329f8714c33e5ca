"""Call sites that used library matmuls now run on the framework's tcgen05 GEMM (AutoTP linears, FPDT chunked FFN and
chunked logits loss).  Oracle: the same op in PyTorch fp32."""
import pytest
import torch
from torch import nn

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-9)).item()


def test_autotp_linears_forward_backward_on_tcgen05():
    from shuffle_exchange_b200.module_inject.layers import LinearAllreduce, LinearLayer, set_autotp_mode
    from shuffle_exchange_b200.native import lib
    set_autotp_mode(training=True)
    try:
        torch.manual_seed(0)
        for cls in (LinearLayer, LinearAllreduce):
            base = nn.Linear(256, 384, bias=True).cuda().bfloat16()
            layer = cls(base, mp_group=None)
            x = torch.randn(4, 64, 256, device="cuda", dtype=torch.bfloat16, requires_grad=True)
            before = lib.launch_count
            y = layer(x)
            g = torch.randn_like(y)
            y.backward(g)
            xf = x.detach().float().requires_grad_(True)
            wf = base.weight.detach().float().requires_grad_(True)
            yf = torch.nn.functional.linear(xf, wf, base.bias.detach().float())
            yf.backward(g.float())
            assert _rel(y, yf) < 1e-2 and _rel(x.grad, xf.grad) < 1e-2 and _rel(layer.weight.grad, wf.grad) < 1e-2
            assert lib.launch_count >= before + 3          # forward, dgrad and wgrad went through the native library
    finally:
        set_autotp_mode(training=False)


def test_fpdt_chunked_ffn_and_logits_loss_match_fp32():
    from shuffle_exchange_b200.sequence.fpdt_layer import FPDT_FFN, FPDT_LogitsLoss
    torch.manual_seed(1)
    s, b, h, f, V = 64, 2, 128, 512, 1024
    bf = dict(device="cuda", dtype=torch.bfloat16)
    x = torch.randn(s, b, h, **bf, requires_grad=True)
    w1 = (torch.randn(f, h, **bf) * 0.05).requires_grad_(True)
    b1 = (torch.randn(f, **bf) * 0.05).requires_grad_(True)
    w2 = (torch.randn(h, f, **bf) * 0.05).requires_grad_(True)
    b2 = (torch.randn(h, **bf) * 0.05).requires_grad_(True)
    y, _ = FPDT_FFN.apply(x, w1, b1, w2, b2, True, 16)
    g = torch.randn_like(y)
    y.backward(g)
    ref = [t.detach().float().requires_grad_(True) for t in (x, w1, b1, w2, b2)]
    yr = torch.nn.functional.linear(
        torch.nn.functional.gelu(torch.nn.functional.linear(ref[0], ref[1], ref[2]), approximate="tanh"), ref[3], ref[4])
    yr.backward(g.float())
    assert _rel(y, yr) < 2e-2
    for got, want in zip((x, w1, b1, w2, b2), ref):
        assert _rel(got.grad, want.grad) < 3e-2

    lm = torch.randn(s, b, h, **bf, requires_grad=True)
    lw = (torch.randn(V, h, **bf) * 0.05).requires_grad_(True)
    labels = torch.randint(0, V, (b, s), device="cuda")
    loss = FPDT_LogitsLoss.apply(lm, labels, lw, 0, 1, None, 4)
    loss.sum().backward()
    lmf, lwf = lm.detach().float().requires_grad_(True), lw.detach().float().requires_grad_(True)
    logits = torch.nn.functional.linear(lmf, lwf)
    lr = torch.nn.functional.cross_entropy(logits.view(-1, V), labels.t().reshape(-1), reduction="none").view(s, b)
    lr.sum().backward()
    assert _rel(loss, lr) < 1e-2 and _rel(lm.grad, lmf.grad) < 3e-2 and _rel(lw.grad, lwf.grad) < 3e-2
