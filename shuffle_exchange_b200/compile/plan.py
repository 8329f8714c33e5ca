"""Execution-plan IR for the schedule compiler.

One training iteration of a ZeRO-3 engine is a fixed sequence over *units* (modules whose parameters travel together):
forward visits them in recorded order, backward in reverse.  A plan makes that sequence explicit as nodes

    GATHER(u)  — all-gather the bf16 shards of unit u        (bytes: full unit size)
    COMPUTE(u) — run the unit's forward or backward           (time measured by the profiler)
    RELEASE(u) — drop the gathered copy
    REDUCE(u)  — reduce-scatter the unit's gradients (backward only)

so passes can move GATHERs earlier (prefetch), delete GATHER/RELEASE pairs (persistent units) or insert OFFLOAD/RELOAD
nodes, and the memory simulator can say what a candidate schedule peaks at.  The reference manipulates the same four
ops inside FX graphs (``compile/passes/zero3_compile.py``, ``compile/list_schedule.py:24-275``)."""
from __future__ import annotations

import copy
from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Tuple

GATHER, COMPUTE, RELEASE, REDUCE, OFFLOAD, RELOAD = "gather", "compute", "release", "reduce", "offload", "reload"


@dataclass
class PlanNode:
    kind: str
    unit: int                      # unit index (or -1 for non-unit work such as the loss)
    phase: str = "fwd"             # "fwd" | "bwd"
    bytes: int = 0                 # memory acquired (gather / reload / compute activations) or freed (release / offload)
    time_us: float = 0.0
    note: str = ""

    def key(self) -> Tuple[str, int, str]:
        return (self.kind, self.unit, self.phase)


@dataclass
class ExecutionPlan:
    nodes: List[PlanNode] = field(default_factory=list)
    unit_bytes: Dict[int, int] = field(default_factory=dict)
    unit_names: Dict[int, str] = field(default_factory=dict)
    persistent: set = field(default_factory=set)
    meta: Dict[str, object] = field(default_factory=dict)

    def clone(self) -> "ExecutionPlan":
        return copy.deepcopy(self)

    def index_of(self, kind: str, unit: int, phase: str) -> int:
        for i, n in enumerate(self.nodes):
            if n.kind == kind and n.unit == unit and n.phase == phase:
                return i
        raise KeyError((kind, unit, phase))

    def computes(self, phase: Optional[str] = None) -> List[PlanNode]:
        return [n for n in self.nodes if n.kind == COMPUTE and (phase is None or n.phase == phase)]

    # ------------------------------------------------------------------------------------------ simulation
    def simulate(self, base_bytes: int = 0) -> Dict[str, float]:
        """Walk the plan: gathered-parameter memory profile and a two-stream (compute / comm) time estimate in which a
        COMPUTE waits for its unit's GATHER and gathers run back-to-back on the communication stream."""
        live = base_bytes
        peak = live
        gathered: Dict[int, bool] = {u: True for u in self.persistent}
        t_comm = t_comp = 0.0
        ready: Dict[int, float] = {}
        exposed = 0.0
        for n in self.nodes:
            if n.kind in (GATHER, RELOAD):
                if n.kind == GATHER and gathered.get(n.unit):
                    continue
                live += n.bytes
                peak = max(peak, live)
                gathered[n.unit] = True
                t_comm = max(t_comm, t_comp) + n.time_us      # issued when the compute stream reaches this node
                ready[n.unit] = t_comm
            elif n.kind in (RELEASE, OFFLOAD):
                if n.kind == RELEASE and (n.unit in self.persistent or not gathered.get(n.unit)):
                    continue
                live -= n.bytes
                gathered[n.unit] = False
            elif n.kind == COMPUTE:
                need = ready.get(n.unit, 0.0)
                if need > t_comp:
                    exposed += need - t_comp
                    t_comp = need
                t_comp += n.time_us
            elif n.kind == REDUCE:
                t_comm = max(t_comm, t_comp) + n.time_us
        return {"peak_bytes": float(peak), "time_us": max(t_comp, t_comm), "exposed_comm_us": exposed}

    def validate(self) -> None:
        """Every non-persistent COMPUTE must see its unit gathered and not yet released."""
        have = set(self.persistent)
        for n in self.nodes:
            if n.kind in (GATHER, RELOAD):
                have.add(n.unit)
            elif n.kind in (RELEASE, OFFLOAD):
                if n.unit not in self.persistent:
                    have.discard(n.unit)
            elif n.kind == COMPUTE and n.unit >= 0:
                if n.unit not in have:
                    raise ValueError(f"unit {n.unit} computed in {n.phase} without being gathered")

    def prefetch_table(self) -> Dict[Tuple[int, bool], List[int]]:
        """``(unit, is_forward) -> units to start gathering right after that unit's COMPUTE is issued`` — the form the
        ZeRO-3 optimizer consumes (``DeepSpeedZeroOptimizer_Stage3._compiled_prefetch``)."""
        table: Dict[Tuple[int, bool], List[int]] = {}
        last_compute: Optional[PlanNode] = None
        pending_first: List[int] = []
        for n in self.nodes:
            if n.kind == COMPUTE and n.unit >= 0:
                last_compute = n
                table.setdefault((n.unit, n.phase == "fwd"), [])
            elif n.kind == GATHER and n.unit not in self.persistent:
                if last_compute is None:
                    pending_first.append(n.unit)
                else:
                    table.setdefault((last_compute.unit, last_compute.phase == "fwd"), []).append(n.unit)
        # a gather is "attached" to the compute before it; the compute that needs it directly is excluded because the
        # blocking fetch in the pre-forward hook covers it anyway
        for (u, f), lst in table.items():
            table[(u, f)] = [x for x in lst if x != u]
        self.meta["first_gathers"] = pending_first
        return table


def build_plan(order: Iterable[int], unit_bytes: Dict[int, int], unit_names: Optional[Dict[int, str]] = None,
               persistent: Iterable[int] = (), fwd_us: Optional[Dict[int, float]] = None,
               bwd_us: Optional[Dict[int, float]] = None, gather_us_per_byte: float = 0.0,
               reduce_us_per_byte: float = 0.0) -> ExecutionPlan:
    """The naive (no-prefetch) plan: gather right before use, release right after."""
    order = list(order)
    plan = ExecutionPlan(unit_bytes=dict(unit_bytes), unit_names=dict(unit_names or {}), persistent=set(persistent))
    fwd_us, bwd_us = fwd_us or {}, bwd_us or {}
    for phase, seq in (("fwd", order), ("bwd", list(reversed(order)))):
        for u in seq:
            b = unit_bytes[u]
            plan.nodes.append(PlanNode(GATHER, u, phase, b, b * gather_us_per_byte))
            plan.nodes.append(PlanNode(COMPUTE, u, phase, 0, (fwd_us if phase == "fwd" else bwd_us).get(u, 0.0)))
            plan.nodes.append(PlanNode(RELEASE, u, phase, b))
            if phase == "bwd":
                plan.nodes.append(PlanNode(REDUCE, u, phase, b, b * reduce_us_per_byte))
    return plan


def build_plan_from_engine(engine, profile: Optional[Dict[str, Dict[int, float]]] = None,
                           gather_us_per_byte: float = 0.0, reduce_us_per_byte: float = 0.0) -> ExecutionPlan:
    """Plan of the engine's ZeRO-3 optimizer from its recorded unit trace (run one iteration first)."""
    opt = engine.optimizer
    if not hasattr(opt, "units") or not getattr(opt, "_trace", None):
        raise RuntimeError("build_plan_from_engine needs a ZeRO-3 optimizer that has recorded one iteration")
    seen, order = set(), []
    for idx in opt._trace:          # a unit used twice in forward keeps its first position
        if idx not in seen:
            seen.add(idx)
            order.append(idx)
    elem = 2
    unit_bytes = {u.index: int(u.numel) * elem for u in opt.units}
    names = {u.index: u.name for u in opt.units}
    persistent = [u.index for u in opt.units if u.persistent]
    profile = profile or {}
    return build_plan(order, unit_bytes, names, persistent, profile.get("fwd_us"), profile.get("bwd_us"),
                      gather_us_per_byte, reduce_us_per_byte)
