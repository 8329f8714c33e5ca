"""1-bit Adam (https://arxiv.org/abs/2102.02888).

Reference: ``deepspeed/runtime/fp16/onebit/adam.py:14-306``.  Two phases:

* warm-up (``step < freeze_step``): plain Adam; gradients arrive already averaged by the engine's all-reduce;
* compression: the variance ``exp_avg_sq`` is frozen, the engine's gradient all-reduce is switched off, each rank
  folds its LOCAL gradient into the momentum and the momentum is averaged with the error-compensated 1-bit
  all-reduce (``runtime/comm/compressed.py``).

``exp_avg_mask`` in a param group zeroes momentum entries that must not be touched by compression noise (e.g.
position embeddings beyond the trained length), as in the reference.
"""
from __future__ import annotations

import torch

from .... import comm as dist
from ._base import OnebitBase


class OnebitAdam(OnebitBase):
    def __init__(self, params, deepspeed=None, lr=1e-3, freeze_step=100000, bias_correction=True, betas=(0.9, 0.999),
                 eps=1e-8, eps_inside_sqrt=False, weight_decay=0.0, max_grad_norm=0.0, amsgrad=False, cuda_aware=False,
                 comm_backend_name="nccl"):
        if amsgrad:
            raise RuntimeError("1-bit Adam does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        max_grad_norm=max_grad_norm)
        super().__init__(params, defaults)
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.freeze_step = freeze_step
        self.adam_freeze_key = False
        self.initialize = False
        self._setup_comm(deepspeed, comm_backend_name, cuda_aware)

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        last_step = 0
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                grad = p.grad
                if grad.is_sparse:
                    raise RuntimeError("1-bit Adam does not support sparse gradients")
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = 0
                    state["exp_avg"] = torch.zeros_like(p)
                    state["exp_avg_sq"] = torch.zeros_like(p)
                state["step"] += 1
                last_step = state["step"]
                exp_avg, exp_avg_sq = state["exp_avg"], state["exp_avg_sq"]
                if not self.adam_freeze_key:
                    exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                    exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                elif group.get("non_freeze", False):
                    if dist.is_initialized():
                        dist.all_reduce(grad)
                        grad.mul_(1.0 / dist.get_world_size())
                    exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                    exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                else:
                    self._error_buffers(state, p)
                    exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                    exp_avg.copy_(self._compressed_allreduce(exp_avg, state))
                    if "exp_avg_mask" in group:
                        mask = group["exp_avg_mask"].to(exp_avg.device)
                        group["exp_avg_mask"] = mask
                        exp_avg.mul_(mask)
                update = exp_avg / (exp_avg_sq.sqrt() + group["eps"])
                if group["weight_decay"] > 0.0:
                    update = update + group["weight_decay"] * p
                p.add_(update, alpha=-group["lr"])
        if not self.adam_freeze_key and last_step >= self.freeze_step:
            print("OnebitAdam - starting compressed communication")
            self.adam_freeze_key = True
            self._set_engine_allreduce(False)
        self.initialize = True
        return loss

    def load_state_dict(self, state_dict):
        """Masks are configuration, not state; error-feedback buffers are dropped because they only make sense for the
        momentum they were accumulated against (reference :241-306)."""
        for i, group in enumerate(self.param_groups):
            if "exp_avg_mask" in group:
                state_dict["param_groups"][i]["exp_avg_mask"] = group["exp_avg_mask"]
            elif "exp_avg_mask" in state_dict["param_groups"][i]:
                state_dict["param_groups"][i].pop("exp_avg_mask")
        super().load_state_dict(state_dict)
        first = self.param_groups[0]["params"][0]
        compressed = self.state[first].get("step", 0) >= self.freeze_step
        if dist.is_initialized() and dist.get_rank() == 0:
            print("Checkpoint loaded and OnebitAdam " + ("compression" if compressed else "warmup") + " stage starts/continues.")
        self.adam_freeze_key = compressed
        self._set_engine_allreduce(not compressed)
        for group in self.param_groups:
            for p in group["params"]:
                self._drop_error_buffers(self.state[p])
