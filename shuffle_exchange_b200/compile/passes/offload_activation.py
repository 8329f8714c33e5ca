"""Activation offload (reference ``compile/passes/offload_activation.py``): saved activations of early units are the
last to be consumed in backward, so they are the ones worth parking in host memory.  The pass marks the earliest units
whose activations exceed the budget; the engine applies the mark through ``torch.autograd.graph.save_on_cpu`` around
those units' forward."""
import torch

from ..plan import COMPUTE


def offload_activation_fwd(plan, ctx):
    out = plan.clone()
    act = ctx.activation_bytes or {}
    total = sum(act.values())
    over = total + out.simulate(ctx.resident_bytes)["peak_bytes"] - ctx.total_mem * 0.92
    chosen = []
    if over > 0:
        for n in out.computes("fwd"):
            if over <= 0:
                break
            b = act.get(n.unit, 0)
            if b > 0:
                chosen.append(n.unit)
                over -= b
    out.meta["offload_activation"] = chosen
    eng = ctx.engine
    if eng is not None and chosen:
        opt = eng.optimizer
        for u in getattr(opt, "units", []):
            if u.index in chosen and u.module is not None and not hasattr(u.module, "_sxb_act_offload"):
                _wrap_forward(u.module)
    return out


def _wrap_forward(module):
    orig = module.forward

    def fwd(*a, **k):
        if torch.is_grad_enabled():
            with torch.autograd.graph.save_on_cpu(pin_memory=True):
                return orig(*a, **k)
        return orig(*a, **k)

    module.forward = fwd
    module._sxb_act_offload = True
