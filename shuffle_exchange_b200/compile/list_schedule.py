"""Memory-bounded scheduling of gather nodes.

Reference: ``deepspeed/compile/list_schedule.py`` — ``simple_prefetch`` (:171-209, hoist each all-gather as early as the
memory budget allows) and ``fast_free_schedule`` (:275-, schedule so parameters are released as early as possible, then
place all-gathers greedily).  Same two policies, on :class:`~.plan.ExecutionPlan` instead of FX graphs."""
from __future__ import annotations

import os
from typing import List, Optional

from .plan import COMPUTE, GATHER, RELEASE, ExecutionPlan, PlanNode


def fast_free_schedule(plan: ExecutionPlan, available_mem: int, output_size: int = 0,
                       debug_log: bool = False) -> ExecutionPlan:
    """Canonical low-memory schedule: every GATHER sits immediately before its first COMPUTE of the phase and every
    RELEASE immediately after the last one (duplicates within a phase removed)."""
    out = plan.clone()
    nodes: List[PlanNode] = []
    for phase in ("fwd", "bwd"):
        seq = [n for n in out.nodes if n.phase == phase]
        last_use = {}
        for i, n in enumerate(seq):
            if n.kind == COMPUTE:
                last_use[n.unit] = i
        gathered = set()
        for i, n in enumerate(seq):
            if n.kind == COMPUTE:
                if n.unit >= 0 and n.unit not in gathered and n.unit not in out.persistent:
                    b = out.unit_bytes[n.unit]
                    g = next((m for m in seq if m.kind == GATHER and m.unit == n.unit), None)
                    nodes.append(PlanNode(GATHER, n.unit, phase, b, g.time_us if g else 0.0))
                    gathered.add(n.unit)
                nodes.append(n)
                if last_use.get(n.unit) == i and n.unit in gathered:
                    nodes.append(PlanNode(RELEASE, n.unit, phase, out.unit_bytes[n.unit]))
                    gathered.discard(n.unit)
            elif n.kind not in (GATHER, RELEASE):
                nodes.append(n)
    out.nodes = nodes
    out.validate()
    return out


def simple_prefetch(plan: ExecutionPlan, available_mem: int, output_size: int = 0,
                    debug_log: bool = False, use_native: Optional[bool] = None) -> ExecutionPlan:
    """Starting from the low-memory schedule, move each GATHER earlier, one COMPUTE at a time, while the simulated
    peak of gathered bytes stays within ``available_mem - output_size``.  Gathers keep their relative order so the
    communication stream never reorders (NCCL requires identical order on all ranks)."""
    out = fast_free_schedule(plan, available_mem, output_size, debug_log)
    budget = max(0, int(available_mem) - int(output_size))
    if use_native is None:
        use_native = os.environ.get("SXB_NATIVE_SCHED", "1") != "0"
    if use_native:
        from . import native_sched
        if native_sched.available():         # same loop in C++ (csrc/host/plan_sched.cpp)
            perm = native_sched.hoist_gathers(out, budget)
            out.nodes = [out.nodes[j] for j in perm]
            out.validate()
            if debug_log:
                out.meta["schedule_log"] = [f"{n.phase}:{n.kind}:{n.unit}" for n in out.nodes]
            return out
    nodes = out.nodes
    i = 0
    while i < len(nodes):
        n = nodes[i]
        if n.kind != GATHER or n.unit in out.persistent:
            i += 1
            continue
        pos = i
        while pos > 0:
            prev = nodes[pos - 1]
            if prev.phase != n.phase and prev.kind != COMPUTE:
                break
            if prev.kind == GATHER:                      # never overtake another gather
                break
            if prev.kind == RELEASE and prev.unit == n.unit:
                break
            # try the swap, keep it only when the budget holds
            nodes[pos - 1], nodes[pos] = nodes[pos], nodes[pos - 1]
            if out.simulate()["peak_bytes"] > budget:
                nodes[pos - 1], nodes[pos] = nodes[pos], nodes[pos - 1]
                break
            pos -= 1
        i += 1
    out.validate()
    if debug_log:
        out.meta["schedule_log"] = [f"{n.phase}:{n.kind}:{n.unit}" for n in nodes]
    return out


def list_schedule(plan: ExecutionPlan) -> ExecutionPlan:
    """Reference name for the default policy."""
    return fast_free_schedule(plan, available_mem=0)
