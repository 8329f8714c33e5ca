"""FusedAdam — Adam/AdamW on the sm_100a fused kernels.

API of the reference class (deepspeed/ops/adam/fused_adam.py:18-195: ``FusedAdam(params, lr,
bias_correction, betas, eps, adam_w_mode, weight_decay, amsgrad, set_grad_none)``; one multi-tensor
launch per dtype list).  Two entry points:

* :meth:`step` — ``torch.optim.Optimizer`` contract; parameters are bucketed by dtype and each bucket
  is one ``sxb_multi_tensor_adam`` launch (state dtype == param dtype, as in the reference).
* :meth:`step_flat` — what the ZeRO optimizers call: one launch over a flat fp32 master shard with the
  gradient in bf16/fp16/fp32, loss-scale/clip folded into ``grad_scale``, overflow skip through a
  device flag, and the low-precision parameter shard written in the same pass.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from ...native import lib as _n


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, adam_w_mode=True,
                 weight_decay=0.0, amsgrad=False, set_grad_none=True):
        if amsgrad:
            raise RuntimeError("FusedAdam does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.set_grad_none = set_grad_none

    def zero_grad(self, set_to_none: Optional[bool] = None):
        if set_to_none is None:
            set_to_none = self.set_grad_none
        super().zero_grad(set_to_none=set_to_none)

    # ---------------------------------------------------------------------------------- flat
    def flat_state(self, hp: torch.Tensor):
        st = self.state[hp]
        if "exp_avg" not in st:
            st["step"] = 0
            st["exp_avg"] = torch.zeros_like(hp)
            st["exp_avg_sq"] = torch.zeros_like(hp)
        return st

    @torch.no_grad()
    def step_flat(self, hp: torch.Tensor, grad: torch.Tensor, group: dict, lp_out: Optional[torch.Tensor] = None,
                  grad_scale: float = 1.0, noop_flag: Optional[torch.Tensor] = None,
                  dyn_scale: Optional[torch.Tensor] = None, advance_step: bool = True):
        st = self.flat_state(hp)
        if advance_step:
            st["step"] += 1
        beta1, beta2 = group["betas"]
        n = hp.numel()
        if hp.is_cuda:
            _n.call("sxb_adam_flat", hp.data_ptr(), grad.data_ptr(), st["exp_avg"].data_ptr(),
                    st["exp_avg_sq"].data_ptr(), _n.ptr(lp_out), n, _n.dtype_code(grad.dtype),
                    _n.dtype_code(lp_out.dtype) if lp_out is not None else _n.BF16, float(group["lr"]), float(beta1),
                    float(beta2), float(group["eps"]), float(group["weight_decay"]), int(st["step"]),
                    self.adam_w_mode, int(bool(group.get("bias_correction", True))), float(grad_scale),
                    _n.ptr(noop_flag), _n.ptr(dyn_scale), _n.stream_ptr())
        else:
            _adam_reference(hp, grad, st["exp_avg"], st["exp_avg_sq"], group, st["step"], self.adam_w_mode,
                            grad_scale, noop_flag, dyn_scale)
            if lp_out is not None:
                lp_out.copy_(hp)

    # ------------------------------------------------------------------------------- per-tensor
    @torch.no_grad()
    def step(self, closure=None, grads=None, output_params=None, scale=None, grad_norms=None, grad_scaler=None):
        if any(a is not None for a in (grads, output_params, scale, grad_norms)):
            raise RuntimeError("FusedAdam has been updated; the (grads, output_params, scale, grad_norms) step "
                               "arguments are no longer supported.")
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            if len(group["params"]) == 0:
                continue
            bias_correction = 1 if group.get("bias_correction", True) else 0
            beta1, beta2 = group["betas"]
            group["step"] = group.get("step", 0) + 1
            buckets = {}
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
                st["step"] = group["step"]
                buckets.setdefault((p.dtype, p.is_cuda), []).append((p.grad, p, st["exp_avg"], st["exp_avg_sq"]))
            for (dt, is_cuda), items in buckets.items():
                if is_cuda:
                    if dt not in (torch.float32, torch.float16, torch.bfloat16):
                        raise RuntimeError("FusedAdam only support fp16, bf16 and fp32.")
                    n = len(items)
                    ptrs = (ctypes.c_void_p * (4 * n))()
                    numels = (ctypes.c_longlong * n)()
                    for i, (g, p, m, v) in enumerate(items):
                        g = g if g.is_contiguous() else g.contiguous()
                        ptrs[i], ptrs[n + i], ptrs[2 * n + i], ptrs[3 * n + i] = (g.data_ptr(), p.data_ptr(),
                                                                                 m.data_ptr(), v.data_ptr())
                        numels[i] = p.numel()
                    _n.call("sxb_multi_tensor_adam", ctypes.cast(ptrs, ctypes.c_void_p),
                            ctypes.cast(numels, ctypes.c_void_p), n, _n.dtype_code(dt), float(group["lr"]),
                            float(beta1), float(beta2), float(group["eps"]), float(group["weight_decay"]),
                            int(group["step"]), self.adam_w_mode, bias_correction, None, _n.stream_ptr())
                else:
                    for g, p, m, v in items:
                        _adam_reference(p, g, m, v, group, group["step"], self.adam_w_mode, 1.0, None, None)
        return loss


def _adam_reference(p, g, m, v, group, step, adamw, grad_scale, noop_flag, dyn_scale):
    """fp32 PyTorch statement of the kernel math (CPU path and test oracle)."""
    if noop_flag is not None and int(noop_flag.item()) != 0:
        return
    beta1, beta2 = group["betas"]
    lr, eps, wd = group["lr"], group["eps"], group["weight_decay"]
    gs = grad_scale * (float(dyn_scale.item()) if dyn_scale is not None else 1.0)
    pf = p.float()
    gf = g.float() * gs
    if not adamw:
        gf = gf + wd * pf
    mf = m.float().mul(beta1).add(gf, alpha=1 - beta1)
    vf = v.float().mul(beta2).addcmul(gf, gf, value=1 - beta2)
    if group.get("bias_correction", True):
        bc1, bc2 = 1 - beta1 ** step, 1 - beta2 ** step
    else:
        bc1 = bc2 = 1.0
    upd = (mf / bc1) / ((vf / bc2).sqrt() + eps)
    if adamw:
        upd = upd + wd * pf
    p.copy_((pf - lr * upd).to(p.dtype))
    m.copy_(mf.to(m.dtype))
    v.copy_(vf.to(v.dtype))
