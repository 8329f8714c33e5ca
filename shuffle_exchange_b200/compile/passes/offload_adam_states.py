"""Optimizer-state offload (reference ``compile/passes/offload_adam_states.py``): when the iteration's peak does not
fit next to the fp32 master + Adam moments, those states leave the GPU for the duration of forward/backward and come
back before the step.  ``offload_adam_states`` decides from the simulated peak; ``*_sync`` always offloads (reference
keeps both variants)."""
from ..plan import OFFLOAD, RELOAD, PlanNode


def _install(ctx, plan, always: bool):
    eng = ctx.engine
    need = always or plan.simulate(ctx.resident_bytes)["peak_bytes"] > ctx.total_mem * 0.92
    plan.meta["offload_adam_states"] = bool(need)
    if not need:
        return plan
    state_bytes = ctx.opt_state_bytes
    plan.nodes.insert(0, PlanNode(OFFLOAD, -1, "fwd", state_bytes, note="adam states -> host"))
    plan.nodes.append(PlanNode(RELOAD, -1, "bwd", state_bytes, note="adam states <- host"))
    if eng is not None and hasattr(eng, "offload_states"):
        hooks = eng.__dict__.setdefault("_compile_hooks", {"pre_forward": [], "pre_step": []})
        hooks["pre_forward"].append(lambda e=eng: _offload(e))
        hooks["pre_step"].append(lambda e=eng: _reload(e))
    return plan


def _offload(engine):
    if not getattr(engine, "_adam_states_on_host", False):
        from ...runtime.zero.offload_config import OffloadStateTypeEnum
        engine.offload_states(include=[OffloadStateTypeEnum.optim_states], pin_memory=True, non_blocking=True)
        engine._adam_states_on_host = True


def _reload(engine):
    if getattr(engine, "_adam_states_on_host", False):
        engine.reload_states(non_blocking=True)
        engine._adam_states_on_host = False


def offload_adam_states(plan, ctx):
    return _install(ctx, plan.clone(), always=False)


def offload_adam_states_sync(plan, ctx):
    return _install(ctx, plan.clone(), always=True)
