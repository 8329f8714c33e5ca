"""DeepSpeedCPUAdam — host Adam/AdamW for ZeRO-Offload on the AVX-512 + OpenMP kernel in
``csrc/host/cpu_optim.cpp``.

Constructor and ``step`` follow the reference (deepspeed/ops/adam/cpu_adam.py:18-187: ``model_params,
lr, bias_correction, betas, eps, weight_decay, amsgrad, adamw_mode, fp32_optimizer_states``; native
calls ``create_adam`` :94-102, ``adam_update`` :163-165, ``destroy_adam`` :104-108).  ``step_flat`` is
the entry the ZeRO optimizers use: it can emit the updated parameters directly as bf16/fp16 into a
pinned staging buffer so the H2D copy moves 2 bytes per element.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from ...native import host as _h
from ...utils.logging import logger


def _lp_kind(t: Optional[torch.Tensor]) -> int:
    if t is None:
        return 0
    return {torch.bfloat16: 1, torch.float16: 2}[t.dtype]


class DeepSpeedCPUAdam(torch.optim.Optimizer):
    optimizer_id = 0

    def __init__(self, model_params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, weight_decay=0,
                 amsgrad=False, adamw_mode=True, fp32_optimizer_states=True):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, bias_correction=bias_correction,
                        amsgrad=amsgrad)
        super().__init__(model_params, defaults)
        self.opt_id = DeepSpeedCPUAdam.optimizer_id
        DeepSpeedCPUAdam.optimizer_id += 1
        self.adam_w_mode = adamw_mode
        self.fp32_optimizer_states = fp32_optimizer_states
        self._lib = _h.load()
        self._lib.sxbh_create_adam(self.opt_id, lr, betas[0], betas[1], eps, weight_decay, int(adamw_mode), 0)

    def __del__(self):
        try:
            self._lib.sxbh_destroy_adam(self.opt_id)
        except Exception:
            pass

    def __setstate__(self, state):
        super().__setstate__(state)
        for g in self.param_groups:
            g.setdefault("amsgrad", False)

    def _update(self, p, g, m, v, group, step, lp=None, grad_scale=1.0):
        assert p.device.type == "cpu" and p.dtype == torch.float32, "CPUAdam param is on {} and dtype {}; expected cpu fp32".format(p.device, p.dtype)
        g = g if (g.dtype == torch.float32 and g.is_contiguous()) else g.float().contiguous()
        b1, b2 = group["betas"]
        self._lib.sxbh_adam_update(self.opt_id, int(step), float(group["lr"]), float(b1), float(b2), float(group["eps"]),
                                   float(group["weight_decay"]), int(bool(group["bias_correction"])), p.data_ptr(),
                                   g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(),
                                   None if lp is None else lp.data_ptr(), _lp_kind(lp), float(grad_scale))

    def flat_state(self, hp):
        st = self.state[hp]
        if "exp_avg" not in st:
            st["step"] = 0
            st["exp_avg"] = torch.zeros_like(hp, device="cpu")
            st["exp_avg_sq"] = torch.zeros_like(hp, device="cpu")
        return st

    @torch.no_grad()
    def step_flat(self, hp, grad, group, lp_out=None, grad_scale=1.0, noop_flag=None, dyn_scale=None,
                  advance_step=True):
        if noop_flag is not None and int(noop_flag.item()) != 0:
            return
        if dyn_scale is not None:
            grad_scale = grad_scale * float(dyn_scale.item())
        st = self.flat_state(hp)
        if advance_step:
            st["step"] += 1
        self._update(hp, grad, st["exp_avg"], st["exp_avg_sq"], group, st["step"], lp_out, grad_scale)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p.data, dtype=torch.float32, device="cpu")
                    st["exp_avg_sq"] = torch.zeros_like(p.data, dtype=torch.float32, device="cpu")
                st["step"] += 1
                self._update(p.data, p.grad.data, st["exp_avg"], st["exp_avg_sq"], group, st["step"])
        return loss
