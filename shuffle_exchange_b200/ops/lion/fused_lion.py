"""FusedLion (reference: deepspeed/ops/lion/fused_lion.py:17-131, ``multi_tensor_lion`` :118-128) on
``sxb_lion_flat``: p <- p(1 - lr*wd) - lr*sign(b1*m + (1-b1)*g);  m <- b2*m + (1-b2)*g."""
import torch

from ...native import lib as _n


class FusedLion(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), weight_decay=0.0, set_grad_none=True):
        super().__init__(params, dict(lr=lr, betas=betas, weight_decay=weight_decay))
        self.set_grad_none = set_grad_none

    def zero_grad(self, set_to_none=None):
        super().zero_grad(set_to_none=self.set_grad_none if set_to_none is None else set_to_none)

    def flat_state(self, hp):
        st = self.state[hp]
        if "exp_avg" not in st:
            st["step"] = 0
            st["exp_avg"] = torch.zeros_like(hp)
        return st

    @torch.no_grad()
    def step_flat(self, hp, grad, group, lp_out=None, grad_scale=1.0, noop_flag=None, dyn_scale=None,
                  advance_step=True):
        st = self.flat_state(hp)
        if advance_step:
            st["step"] += 1
        if dyn_scale is not None:
            grad_scale = grad_scale * float(dyn_scale.item())
        b1, b2 = group["betas"]
        if hp.is_cuda:
            _n.call("sxb_lion_flat", hp.data_ptr(), grad.data_ptr(), st["exp_avg"].data_ptr(), _n.ptr(lp_out),
                    hp.numel(), _n.dtype_code(grad.dtype), _n.dtype_code(lp_out.dtype) if lp_out is not None else _n.BF16,
                    float(group["lr"]), float(b1), float(b2), float(group["weight_decay"]), float(grad_scale),
                    _n.ptr(noop_flag), _n.stream_ptr())
        else:
            _lion_reference(hp, grad.float() * grad_scale, st["exp_avg"], group)
            if lp_out is not None:
                lp_out.copy_(hp)

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            group["step"] = group.get("step", 0) + 1
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise NotImplementedError("FusedLion does not support sparse gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
                if p.is_cuda and p.dtype == torch.float32:
                    b1, b2 = group["betas"]
                    g = p.grad.contiguous()
                    _n.call("sxb_lion_flat", p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(), None, p.numel(),
                            _n.dtype_code(g.dtype), _n.BF16, float(group["lr"]), float(b1), float(b2),
                            float(group["weight_decay"]), 1.0, None, _n.stream_ptr())
                else:
                    pf = p.float()
                    _lion_reference(pf, p.grad.float(), st["exp_avg"], group)
                    p.copy_(pf.to(p.dtype))
        return loss


def _lion_reference(p, g, m, group):
    b1, b2 = group["betas"]
    lr, wd = group["lr"], group["weight_decay"]
    c = m * b1 + g * (1 - b1)
    p.mul_(1 - lr * wd).add_(torch.sign(c), alpha=-lr)
    m.mul_(b2).add_(g, alpha=1 - b2)
