"""FusedLamb (reference: deepspeed/ops/lamb/fused_lamb.py:12-173; native ``lamb(...)`` :165 returning
the trust coefficient).  Two launches per tensor: (1) moments + raw update + ||p||², ||u||² ;
(2) clamp(||p||/||u||, min_coeff, max_coeff) and apply."""
import torch

from ...native import lib as _n


class FusedLamb(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, eps_inside_sqrt=False,
                 weight_decay=0.0, max_grad_norm=0.0, max_coeff=10.0, min_coeff=0.01, amsgrad=False):
        if amsgrad:
            raise RuntimeError("FusedLamb does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        max_grad_norm=max_grad_norm, max_coeff=max_coeff, min_coeff=min_coeff)
        super().__init__(params, defaults)
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.lamb_coeffs = []

    def get_lamb_coeffs(self):
        return [float(c) for c in self.lamb_coeffs]

    @torch.no_grad()
    def step(self, closure=None, grads=None, output_params=None, scale=1.0, grad_norms=None):
        loss = closure() if closure is not None else None
        self.lamb_coeffs = []
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("FusedLamb does not support sparse gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p, dtype=torch.float32)
                    st["exp_avg_sq"] = torch.zeros_like(p, dtype=torch.float32)
                st["step"] += 1
                combined = scale
                if group["max_grad_norm"] > 0 and grad_norms is not None:
                    pass
                if p.is_cuda and p.dtype == torch.float32:
                    g = p.grad.contiguous()
                    upd = torch.empty_like(p)
                    norms = torch.zeros(2, dtype=torch.float32, device=p.device)
                    coeff = torch.zeros(1, dtype=torch.float32, device=p.device)
                    _n.call("sxb_lamb", p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(),
                            st["exp_avg_sq"].data_ptr(), None, upd.data_ptr(), norms.data_ptr(), coeff.data_ptr(),
                            p.numel(), _n.dtype_code(g.dtype), _n.BF16, float(group["lr"]), float(b1), float(b2),
                            float(group["eps"]), float(group["weight_decay"]), int(st["step"]),
                            int(bool(group["bias_correction"])), float(group["min_coeff"]), float(group["max_coeff"]),
                            1.0 / combined, _n.stream_ptr())
                    self.lamb_coeffs.append(coeff)
                else:
                    self.lamb_coeffs.append(_lamb_reference(p, p.grad.float() / combined, st, group))
        return loss


def _lamb_reference(p, g, st, group):
    b1, b2 = group["betas"]
    m, v, step = st["exp_avg"], st["exp_avg_sq"], st["step"]
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    bc1 = 1 - b1 ** step if group["bias_correction"] else 1.0
    bc2 = 1 - b2 ** step if group["bias_correction"] else 1.0
    pf = p.float()
    u = (m / bc1) / ((v / bc2).sqrt() + group["eps"]) + group["weight_decay"] * pf
    pn, un = pf.norm(), u.norm()
    coeff = 1.0
    if pn > 0 and un > 0:
        coeff = float(min(max(pn / un, group["min_coeff"]), group["max_coeff"]))
    p.copy_((pf - group["lr"] * coeff * u).to(p.dtype))
    return torch.tensor(coeff)
