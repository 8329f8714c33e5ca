"""ctypes front end of ``csrc/host/plan_sched.cpp`` (plan simulation and gather hoisting in C++).

``simulate(plan)`` and ``hoist_gathers(plan, budget)`` mirror ``ExecutionPlan.simulate`` and the inner loop of
``list_schedule.simple_prefetch``; ``available()`` is False when the host library cannot be loaded, in which case the
callers keep to the Python implementation."""
from __future__ import annotations

import ctypes
from typing import Dict, List

from ..native import host as _host
from .plan import COMPUTE, GATHER, OFFLOAD, REDUCE, RELEASE, RELOAD, ExecutionPlan

_KIND = {GATHER: 0, COMPUTE: 1, RELEASE: 2, REDUCE: 3, OFFLOAD: 4, RELOAD: 5}
_P = ctypes.c_void_p
_host.register({
    "sxbh_plan_simulate": [ctypes.c_int, _P, _P, _P, _P, _P, _P, ctypes.c_int, ctypes.c_longlong, _P],
    "sxbh_plan_hoist_gathers": [ctypes.c_int, _P, _P, _P, _P, _P, _P, ctypes.c_int, ctypes.c_longlong, _P],
})


def available() -> bool:
    try:
        return hasattr(_host.load(), "sxbh_plan_hoist_gathers")
    except Exception:
        return False


def _arrays(plan: ExecutionPlan):
    n = len(plan.nodes)
    kind = (ctypes.c_int * n)(*[_KIND[x.kind] for x in plan.nodes])
    unit = (ctypes.c_int * n)(*[int(x.unit) for x in plan.nodes])
    phase = (ctypes.c_int * n)(*[0 if x.phase == "fwd" else 1 for x in plan.nodes])
    nbytes = (ctypes.c_longlong * n)(*[int(x.bytes) for x in plan.nodes])
    time_us = (ctypes.c_double * n)(*[float(x.time_us) for x in plan.nodes])
    pers = sorted(int(u) for u in plan.persistent)
    parr = (ctypes.c_int * max(1, len(pers)))(*pers) if pers else (ctypes.c_int * 1)()
    return n, kind, unit, phase, nbytes, time_us, parr, len(pers)


def simulate(plan: ExecutionPlan, base_bytes: int = 0) -> Dict[str, float]:
    n, kind, unit, phase, nbytes, time_us, parr, npers = _arrays(plan)
    out = (ctypes.c_double * 3)()
    rc = _host.load().sxbh_plan_simulate(n, kind, unit, phase, nbytes, time_us, parr, npers, int(base_bytes), out)
    if rc != 0:
        raise RuntimeError(f"sxbh_plan_simulate failed ({rc})")
    return {"peak_bytes": out[0], "time_us": out[1], "exposed_comm_us": out[2]}


def hoist_gathers(plan: ExecutionPlan, budget: int) -> List[int]:
    """Permutation of ``plan.nodes`` after hoisting (``new_nodes[i] = plan.nodes[perm[i]]``)."""
    n, kind, unit, phase, nbytes, time_us, parr, npers = _arrays(plan)
    perm = (ctypes.c_int * max(1, n))()
    rc = _host.load().sxbh_plan_hoist_gathers(n, kind, unit, phase, nbytes, time_us, parr, npers, int(budget), perm)
    if rc != 0:
        raise RuntimeError(f"sxbh_plan_hoist_gathers failed ({rc})")
    return [perm[i] for i in range(n)]
