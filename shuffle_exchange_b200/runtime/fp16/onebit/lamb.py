"""1-bit LAMB (https://arxiv.org/abs/2104.06069).

Reference: ``deepspeed/runtime/fp16/onebit/lamb.py:15-443``.  Warm-up runs exact LAMB and records, per parameter,
an exponential average of the trust ratio (``lamb_coeff_freeze``).  In the compression phase the variance is frozen,
momenta of all parameters are scaled to a common magnitude (``scaling_coeff``), flattened and averaged with ONE
compressed all-reduce, and the frozen trust ratio is corrected by the ratio between the frozen variance and a
freshly tracked one (``exp_avg_sq_fresh``), clipped to ``[factor_min, factor_max]`` and rate-limited by
``factor_threshold`` — a local estimate of how LAMB's coefficient would have moved.
"""
from __future__ import annotations

import torch

from .... import comm as dist
from ._base import OnebitBase


class OnebitLamb(OnebitBase):
    def __init__(self, params, deepspeed=None, lr=1e-3, freeze_step=100000, bias_correction=True, betas=(0.9, 0.999),
                 eps=1e-8, eps_inside_sqrt=False, weight_decay=0.0, max_grad_norm=0.0, max_coeff=10.0, min_coeff=0.01,
                 amsgrad=False, cuda_aware=False, comm_backend_name="nccl", coeff_beta=0.9, factor_max=4.0,
                 factor_min=0.5, factor_threshold=0.1):
        if amsgrad:
            raise RuntimeError("1-bit Lamb does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        max_grad_norm=max_grad_norm, max_coeff=max_coeff, min_coeff=min_coeff)
        super().__init__(params, defaults)
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.freeze_step = freeze_step
        self.lamb_freeze_key = False
        self.initialize = False
        self.coeff_beta, self.factor_max, self.factor_min = coeff_beta, factor_max, factor_min
        self.factor_threshold = factor_threshold
        self.lamb_coeffs = []
        self._setup_comm(deepspeed, comm_backend_name, cuda_aware)
        self._flat_state = {}

    def get_lamb_coeffs(self):
        return self.lamb_coeffs

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        self.lamb_coeffs = []
        live = [(g, p) for g in self.param_groups for p in g["params"] if p.grad is not None]
        for group, p in live:
            state = self.state[p]
            if len(state) == 0:
                state["step"] = 0
                state["lamb_coeff_freeze"] = 0.0
                state["last_factor"] = 1.0
                state["exp_avg"] = torch.zeros_like(p)
                state["exp_avg_sq"] = torch.zeros_like(p)
                state["exp_avg_sq_fresh"] = torch.zeros_like(p)
            state["step"] += 1
        if not live:
            return loss
        if not self.lamb_freeze_key:
            for group, p in live:
                state = self.state[p]
                beta1, beta2 = group["betas"]
                m, v = state["exp_avg"], state["exp_avg_sq"]
                m.mul_(beta1).add_(p.grad, alpha=1 - beta1)
                v.mul_(beta2).addcmul_(p.grad, p.grad, value=1 - beta2)
                if state["step"] == self.freeze_step:
                    state["exp_avg_sq_fresh"].copy_(v)
                update = m / (v.sqrt() + group["eps"])
                if group["weight_decay"] > 0.0:
                    update = update + group["weight_decay"] * p
                w_norm, u_norm = p.norm(), update.norm()
                coeff = 1.0
                if w_norm != 0 and u_norm != 0:
                    coeff = min(max((w_norm / u_norm).item(), group["min_coeff"]), group["max_coeff"])
                    if state["step"] > 1:
                        state["lamb_coeff_freeze"] = self.coeff_beta * state["lamb_coeff_freeze"] + (1 - self.coeff_beta) * coeff
                    else:
                        state["lamb_coeff_freeze"] = coeff
                self.lamb_coeffs.append(coeff)
                p.add_(update, alpha=-group["lr"] * coeff)
            if self.state[live[0][1]]["step"] >= self.freeze_step:
                print("OnebitLamb - starting compressed communication")
                self.lamb_freeze_key = True
                self._set_engine_allreduce(False)
                # common magnitude for the fused momentum: scale every momentum to the mean RMS
                rms = [self.state[p]["exp_avg"].norm() / (p.numel() ** 0.5) for _, p in live]
                united = sum(rms) / len(rms)
                for (_, p), r in zip(live, rms):
                    self.state[p]["scaling_coeff"] = (united / r).item() if r > 0 else 1.0
            self.initialize = True
            return loss
        # ---------------------------------------------------------------- compression phase
        flats = []
        for group, p in live:
            state = self.state[p]
            beta1, _ = group["betas"]
            state["exp_avg"].mul_(beta1).add_(p.grad, alpha=1 - beta1)
            state["grad_snapshot"] = p.grad.clone()
            flats.append((state["exp_avg"] * state.get("scaling_coeff", 1.0)).reshape(-1))
        fused = torch.cat(flats)
        fs = self._flat_state
        if fs.get("n") != fused.numel():
            fs.clear()
            fs["n"] = fused.numel()
            class _P:  # shape carrier for the buffer allocator
                def __init__(s, t): s.t = t
                def numel(s): return s.t.numel()
                device = fused.device
                dtype = fused.dtype
            self._error_buffers(fs, _P(fused))
        fused = self._compressed_allreduce(fused, fs)
        off = 0
        for group, p in live:
            state = self.state[p]
            beta1, beta2 = group["betas"]
            n = p.numel()
            m = state["exp_avg"]
            prev = m.clone()
            m.copy_(fused[off: off + n].view_as(p) / state.get("scaling_coeff", 1.0))
            off += n
            # reconstruct the averaged gradient implied by the momentum change to keep a fresh variance estimate
            g_hat = (m - prev * 1.0) / (1 - beta1) + state["grad_snapshot"]
            state.pop("grad_snapshot")
            fresh = state["exp_avg_sq_fresh"]
            fresh.mul_(beta2).addcmul_(g_hat, g_hat, value=1 - beta2)
            v = state["exp_avg_sq"]
            denom = v.sqrt() + group["eps"]
            update_prelim = m / denom
            update = update_prelim + group["weight_decay"] * p if group["weight_decay"] > 0.0 else update_prelim
            factor = (denom / (fresh.sqrt() + group["eps"])).max().item()
            if group["weight_decay"] > 0.0:
                ratio = min(1.0, (update_prelim.norm() / update.norm()).item()) if update.norm() > 0 else 1.0
                factor = factor * ratio + (1.0 - ratio)
            factor = min(max(factor, self.factor_min), self.factor_max)
            lo, hi = state["last_factor"] * (1 - self.factor_threshold), state["last_factor"] * (1 + self.factor_threshold)
            factor = min(max(factor, lo), hi)
            state["last_factor"] = factor
            coeff = state["lamb_coeff_freeze"] * factor
            self.lamb_coeffs.append(coeff)
            p.add_(update, alpha=-group["lr"] * coeff)
        return loss

    def load_state_dict(self, state_dict):
        for i, group in enumerate(self.param_groups):
            if "exp_avg_mask" in group:
                state_dict["param_groups"][i]["exp_avg_mask"] = group["exp_avg_mask"]
            elif "exp_avg_mask" in state_dict["param_groups"][i]:
                state_dict["param_groups"][i].pop("exp_avg_mask")
        super().load_state_dict(state_dict)
        first = self.param_groups[0]["params"][0]
        compressed = self.state[first].get("step", 0) >= self.freeze_step
        self.lamb_freeze_key = compressed
        self._set_engine_allreduce(not compressed)
        self._flat_state = {}
        for group in self.param_groups:
            for p in group["params"]:
                self._drop_error_buffers(self.state[p])
                if not compressed:
                    self.state[p].pop("scaling_coeff", None)
