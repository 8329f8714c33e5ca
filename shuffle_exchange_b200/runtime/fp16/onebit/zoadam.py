"""0/1 Adam (https://arxiv.org/abs/2202.06009).

Reference: ``deepspeed/runtime/fp16/onebit/zoadam.py:14-365``.  Compared with 1-bit Adam it (a) never needs a
full-precision warm-up: the variance is refreshed from all-reduced gradients only at steps spaced by an interval that
doubles every ``var_update_scaler`` refreshes, until ``var_freeze_step``; in between, momentum moves with 1-bit
all-reduce; (b) after the variance is frozen it takes LOCAL steps: parameters move with the local momentum, and
only every ``local_step_interval`` steps (doubling every ``local_step_scaler`` steps, capped at
``local_step_clipper``) the accumulated update is reconciled across ranks through one compressed all-reduce.
"""
from __future__ import annotations

import torch

from .... import comm as dist
from ._base import OnebitBase


class ZeroOneAdam(OnebitBase):
    def __init__(self, params, deepspeed=None, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8,
                 eps_inside_sqrt=False, weight_decay=0.0, max_grad_norm=0.0, var_freeze_step=100000,
                 var_update_scaler=16, local_step_scaler=32678, local_step_clipper=16, amsgrad=False, cuda_aware=False,
                 comm_backend_name="nccl"):
        if amsgrad:
            raise RuntimeError("0/1 Adam does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        max_grad_norm=max_grad_norm)
        super().__init__(params, defaults)
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.var_freeze_step, self.var_update_scaler = var_freeze_step, var_update_scaler
        self.local_step_scaler, self.local_step_clipper = local_step_scaler, local_step_clipper
        self.freeze_key = False
        self.initialize = False
        self._setup_comm(deepspeed, comm_backend_name, cuda_aware)
        self._set_engine_allreduce(False)        # this optimizer owns all gradient communication

    def _init_state(self, state, p):
        state["step"] = 0
        state["exp_avg"] = torch.zeros_like(p)
        state["exp_avg_sq"] = torch.zeros_like(p)
        state["var_interval"], state["var_counter"] = 1, 0
        state["local_step_interval"], state["local_step_counter"] = 1, 0
        state["lrs"] = 0.0
        state["momentum_accumulator"] = torch.zeros_like(p)

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        world = dist.get_world_size() if dist.is_initialized() else 1
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                grad = p.grad
                if grad.is_sparse:
                    raise RuntimeError("0/1 Adam does not support sparse gradients")
                state = self.state[p]
                if len(state) == 0:
                    self._init_state(state, p)
                self._error_buffers(state, p)
                state["step"] += 1
                exp_avg, exp_avg_sq = state["exp_avg"], state["exp_avg_sq"]
                if not self.freeze_key:
                    if state["step"] % state["var_interval"] == 0:
                        # variance refresh step: exact averaged gradient
                        if world > 1:
                            dist.all_reduce(grad)
                            grad.mul_(1.0 / world)
                        exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                        exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                        state["var_counter"] += 1
                        if state["var_counter"] == self.var_update_scaler:
                            state["var_counter"] = 0
                            state["var_interval"] *= 2
                    else:
                        # 1-bit step: the momentum absorbs the locally compressed, globally averaged gradient
                        g = self._compressed_allreduce(grad.clone(), state)
                        exp_avg.mul_(beta1).add_(g, alpha=1 - beta1)
                    update = exp_avg / (exp_avg_sq.sqrt() + group["eps"])
                    if group["weight_decay"] > 0.0:
                        update = update + group["weight_decay"] * p
                    p.add_(update, alpha=-group["lr"])
                else:
                    # local steps with frozen variance
                    exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                    denom = exp_avg_sq.sqrt() + group["eps"]
                    update = exp_avg / denom
                    if group["weight_decay"] > 0.0:
                        update = update + group["weight_decay"] * p
                    p.add_(update, alpha=-group["lr"])
                    state["lrs"] += group["lr"]
                    state["momentum_accumulator"].add_(update, alpha=-group["lr"])
                    state["local_step_counter"] += 1
                    if state["local_step_counter"] % state["local_step_interval"] == 0:
                        # reconcile: replace the sum of local moves by the (compressed) average of everyone's moves,
                        # and realign the momentum with the averaged direction
                        acc = state["momentum_accumulator"]
                        p.sub_(acc)
                        avg = self._compressed_allreduce(acc.clone(), state)
                        p.add_(avg)
                        if state["lrs"] > 0:
                            exp_avg.copy_(-(avg * denom) / state["lrs"])
                        acc.zero_()
                        state["lrs"] = 0.0
                        state["local_step_counter"] = 0
                    if state["step"] % self.local_step_scaler == 0:
                        state["local_step_interval"] = min(self.local_step_clipper, state["local_step_interval"] * 2)
                if not self.freeze_key and state["step"] >= self.var_freeze_step:
                    pass
        if not self.freeze_key:
            any_state = next((self.state[p] for g in self.param_groups for p in g["params"] if p in self.state), None)
            if any_state is not None and any_state["step"] >= self.var_freeze_step:
                self.freeze_key = True
                print("ZeroOneAdam - variance frozen, starting local steps")
        self.initialize = True
        return loss

    def load_state_dict(self, state_dict):
        for i, group in enumerate(self.param_groups):
            if "exp_avg_mask" in group:
                state_dict["param_groups"][i]["exp_avg_mask"] = group["exp_avg_mask"]
            elif "exp_avg_mask" in state_dict["param_groups"][i]:
                state_dict["param_groups"][i].pop("exp_avg_mask")
        super().load_state_dict(state_dict)
        first = self.param_groups[0]["params"][0]
        self.freeze_key = self.state[first].get("step", 0) >= self.var_freeze_step
        self._set_engine_allreduce(False)
        for group in self.param_groups:
            for p in group["params"]:
                self._drop_error_buffers(self.state[p])
