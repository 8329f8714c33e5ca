"""Pass registry, pass schedule and application of a compiled plan to the engine.

Reference: ``deepspeed/compile/backend.py`` — ``register_compile_pass`` (:72), ``init_schedule`` (:76: a list of
``(step, [pass names])``), ``launch_compile_passes(global_steps)`` (:88: when training reaches a scheduled step, the
next group of passes is applied and the graphs are re-optimised), ``run_opt_passes`` (:179) and ``make_backend`` (:217).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch

from ..utils.logging import log_dist
from .passes import PASSES
from .plan import ExecutionPlan, build_plan_from_engine

opt_passes: Dict[str, Callable] = dict(PASSES)
_schedule: List[Tuple[int, List[str]]] = []
_next_pass_step = -1
_next_passes: Optional[List[str]] = None
_state: Dict[str, object] = {}


@dataclass
class PassContext:
    engine: object = None
    config: object = None
    available_mem: int = 0          # bytes the gathered parameters may occupy
    total_mem: int = 0              # device memory
    resident_bytes: int = 0         # shards + optimizer states + persistent buffers
    opt_state_bytes: int = 0
    output_size: int = 0
    world_size: int = 1
    activation_bytes: Dict[int, int] = field(default_factory=dict)
    profile: Dict[str, Dict[int, float]] = field(default_factory=dict)


def register_compile_pass(name: str, opt_pass_fn: Callable) -> None:
    opt_passes[name] = opt_pass_fn


def init_schedule(schedule: Sequence[Tuple[int, Sequence[str]]]) -> None:
    """``schedule`` = ``[(step, [pass, ...]), ...]`` with strictly increasing steps."""
    global _schedule, _next_pass_step, _next_passes
    assert isinstance(schedule, (list, tuple)), f"schedule should be a list, but got {type(schedule)}"
    last = -1
    for step, passes in schedule:
        assert isinstance(step, int) and step > last, "schedule steps must be increasing integers"
        assert isinstance(passes, (list, tuple)), f"passes should be a list, but got {type(passes)}"
        for p in passes:
            assert p in opt_passes, f"unknown compile pass {p!r}; registered: {sorted(opt_passes)}"
        last = step
    _schedule = [(int(s), list(p)) for s, p in schedule]
    _next_pass_step, _next_passes = -1, None


def default_schedule(stage: int, cfg) -> List[Tuple[int, List[str]]]:
    """Step 0 is used for tracing/profiling; scheduling passes start once a profile exists (reference
    ``init_z3.py:78-90`` / ``init_z1.py``)."""
    if stage == 3:
        sched = [(0, ["zero3_compile"]), (1, ["prefetch"]), (2, ["selective_gather"])]
        if cfg is not None and cfg.offload_parameters:
            sched[0][1].append("offload_parameters")
    else:
        sched = [(0, ["zero1_compile"])]
    if cfg is not None and cfg.offload_opt_states:
        sched.append((sched[-1][0] + 1, ["offload_adam_states"]))
    if cfg is not None and cfg.offload_activation:
        sched.append((sched[-1][0] + 1, ["offload_activation"]))
    return sched


def run_opt_passes(names: Sequence[str], plan: ExecutionPlan, ctx: PassContext, debug_log: bool = False):
    for name in names:
        before = plan.simulate(0)
        plan = opt_passes[name](plan, ctx)
        plan.validate()
        after = plan.simulate(0)
        if debug_log:
            log_dist(f"[compile] pass {name}: peak {before['peak_bytes'] / 2**30:.2f} -> "
                     f"{after['peak_bytes'] / 2**30:.2f} GiB, exposed comm {before['exposed_comm_us']:.0f} -> "
                     f"{after['exposed_comm_us']:.0f} us", ranks=[0])
    return plan


def make_context(engine, cfg) -> PassContext:
    dev_total = free = 0
    if torch.cuda.is_available():
        free, dev_total = torch.cuda.mem_get_info()
        resident = torch.cuda.memory_allocated()
    else:
        dev_total, free, resident = 1 << 40, 1 << 39, 0
    opt = engine.optimizer
    state_bytes = 0
    for g in getattr(opt, "groups", []) or []:
        n = getattr(g, "shard_numel", 0) or 0
        state_bytes += 12 * int(n)           # fp32 master + two moments
    import torch.distributed as td
    world = td.get_world_size() if td.is_available() and td.is_initialized() else 1
    return PassContext(engine=engine, config=cfg, available_mem=int(free * 0.8), total_mem=int(dev_total),
                       resident_bytes=int(resident), opt_state_bytes=state_bytes, world_size=world)


def apply_plan(engine, plan: ExecutionPlan) -> None:
    """Hand the compiled schedule to the ZeRO-3 optimizer: explicit prefetch lists + persistent set."""
    opt = engine.optimizer
    if not hasattr(opt, "units"):
        return
    for u in opt.units:
        if u.index in plan.persistent and not u.persistent:
            if hasattr(opt, "make_persistent"):
                opt.make_persistent(u)
            else:
                u.persistent = True
    opt._compiled_prefetch = plan.prefetch_table()
    opt._compiled_plan = plan


def launch_compile_passes(global_steps: int, engine=None) -> Optional[ExecutionPlan]:
    """Call once per training step (the engine does, from ``forward``): when ``global_steps`` reaches the next
    scheduled step, run that group of passes on the current plan and apply the result."""
    global _next_pass_step, _next_passes
    engine = engine or _state.get("engine")
    if engine is None or not _schedule:
        return None
    due = [(s, p) for s, p in _schedule if s <= global_steps]
    if not due:
        return None
    names: List[str] = []
    for s, p in due:
        names.extend(p)
        _schedule.remove((s, p))
    cfg = _state.get("config")
    plan = _state.get("plan")
    if plan is None:
        opt = engine.optimizer
        if hasattr(opt, "units") and getattr(opt, "_trace", None):
            from .profilers.comm_profile import estimate_us_per_byte
            g_us, r_us = estimate_us_per_byte(engine)
            plan = build_plan_from_engine(engine, _state.get("profile"), g_us, r_us)
        else:
            # ZeRO-1/2 (nothing to gather) or no trace yet: put the passes back and try next step
            if not hasattr(opt, "units"):
                plan = ExecutionPlan()
            else:
                _schedule[0:0] = due
                return None
    ctx = make_context(engine, cfg)
    ctx.profile = _state.get("profile") or {}
    plan = run_opt_passes(names, plan, ctx, bool(cfg and cfg.debug_log))
    _state["plan"] = plan
    apply_plan(engine, plan)
    log_dist(f"[compile] step {global_steps}: applied passes {names}; persistent units "
             f"{len(plan.persistent)}; simulated {plan.simulate(0)}", ranks=[0])
    return plan


def make_backend(backend, compile_config, compile_kwargs=None, engine=None, schedule=None, stage: int = 3):
    """Bind the schedule compiler to an engine; returns a callable with the reference's backend signature that simply
    runs the module (no graph rewriting happens at call time — plans are applied through the optimizer hooks)."""
    _state.clear()
    _state.update(engine=engine, config=compile_config, backend=backend, kwargs=dict(compile_kwargs or {}))
    init_schedule(schedule if schedule is not None else default_schedule(stage, compile_config))

    def backend_fn(gm, example_inputs=None):
        return gm.forward if hasattr(gm, "forward") else gm

    return backend_fn
