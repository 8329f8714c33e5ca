"""Host optimizers (csrc/host/cpu_optim.cpp, AVX-512 + OpenMP) against plain PyTorch implementations of the same update
rule; sizes are chosen to hit the vector body, the scalar tail and the multi-thread split."""
import pytest
import torch

from shuffle_exchange_b200.ops.adagrad import DeepSpeedCPUAdagrad
from shuffle_exchange_b200.ops.adam import DeepSpeedCPUAdam
from shuffle_exchange_b200.ops.lion import DeepSpeedCPULion

SIZES = [1, 17, 1024, 100003]


def _params(seed):
    g = torch.Generator().manual_seed(seed)
    ps = [torch.randn(n, generator=g).requires_grad_() for n in SIZES]
    return ps, [p.detach().clone().requires_grad_() for p in ps]


def _set_grads(a, b, seed):
    g = torch.Generator().manual_seed(100 + seed)
    for p, q in zip(a, b):
        p.grad = torch.randn(p.shape, generator=g)
        q.grad = p.grad.clone()


@pytest.mark.parametrize("adamw", [True, False])
def test_cpu_adam_matches_torch(adamw):
    ours, ref = _params(0)
    kw = dict(lr=3e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    opt = DeepSpeedCPUAdam(ours, adamw_mode=adamw, **kw)
    topt = (torch.optim.AdamW if adamw else torch.optim.Adam)(ref, **kw)
    for step in range(5):
        _set_grads(ours, ref, step)
        opt.step()
        topt.step()
    for p, q in zip(ours, ref):
        assert torch.allclose(p, q, rtol=2e-5, atol=2e-6), (p - q).abs().max()
    st = opt.state[ours[-1]]
    assert st["step"] == 5 and torch.allclose(st["exp_avg"], topt.state[ref[-1]]["exp_avg"], rtol=1e-5, atol=1e-7)


def test_cpu_adam_flat_step_writes_low_precision_copy_and_honours_scale_and_noop():
    hp = torch.nn.Parameter(torch.randn(4099))          # the flat fp32 master shard of a group, as ZeRO passes it
    ref = hp.detach().clone().requires_grad_()
    grad = torch.randn(4099)
    opt = DeepSpeedCPUAdam([hp], lr=1e-2, weight_decay=0.0)
    topt = torch.optim.AdamW([ref], lr=1e-2, weight_decay=0.0)
    group = opt.param_groups[0]
    lp = torch.zeros(4099, dtype=torch.bfloat16)
    opt.step_flat(hp, grad * 8.0, group, lp_out=lp, grad_scale=1.0 / 8.0)        # loss-scaled gradient
    ref.grad = grad.clone()
    topt.step()
    assert torch.allclose(hp.data, ref.data, rtol=2e-5, atol=2e-6)
    assert torch.equal(lp, hp.data.to(torch.bfloat16))                                 # emitted with round-to-nearest-even
    before = hp.data.clone()
    opt.step_flat(hp, grad, group, noop_flag=torch.tensor([1]))                  # overflow step: nothing moves
    assert torch.equal(hp.data, before) and opt.state[hp]["step"] == 1


def test_cpu_adagrad_matches_torch():
    ours, ref = _params(1)
    opt = DeepSpeedCPUAdagrad(ours, lr=1e-2, eps=1e-10, weight_decay=0.01)
    topt = torch.optim.Adagrad(ref, lr=1e-2, eps=1e-10, weight_decay=0.01)
    for step in range(4):
        _set_grads(ours, ref, step)
        opt.step()
        topt.step()
    for p, q in zip(ours, ref):
        assert torch.allclose(p, q, rtol=2e-5, atol=2e-6), (p - q).abs().max()


def test_cpu_lion_matches_the_update_rule():
    ours, ref = _params(2)
    lr, (b1, b2), wd = 1e-3, (0.9, 0.99), 0.1
    opt = DeepSpeedCPULion(ours, lr=lr, betas=(b1, b2), weight_decay=wd)
    m = [torch.zeros_like(p) for p in ref]
    for step in range(4):
        _set_grads(ours, ref, step)
        opt.step()
        with torch.no_grad():
            for p, mom in zip(ref, m):
                p.mul_(1 - lr * wd)
                p.add_(torch.sign(mom * b1 + p.grad * (1 - b1)), alpha=-lr)
                mom.mul_(b2).add_(p.grad, alpha=1 - b2)
    for p, q in zip(ours, ref):
        assert torch.allclose(p, q, rtol=1e-5, atol=1e-6), (p - q).abs().max()
