"""Selective gather (reference ``compile/passes/selective_gather.py``): keep the parameters that are cheapest to hold
and most expensive to re-fetch resident for the whole iteration.  Units are ranked by gather time saved per byte held;
they are made persistent while the simulated peak stays inside the budget."""
from ..plan import COMPUTE, GATHER


def selective_gather(plan, ctx):
    out = plan.clone()
    budget = ctx.available_mem
    base_peak = out.simulate()["peak_bytes"]
    headroom = budget - base_peak
    if headroom <= 0:
        out.meta["selective_gather"] = []
        return out
    uses = {}
    for n in out.nodes:
        if n.kind == COMPUTE and n.unit >= 0:
            uses[n.unit] = uses.get(n.unit, 0) + 1
    gather_t = {}
    for n in out.nodes:
        if n.kind == GATHER:
            gather_t[n.unit] = gather_t.get(n.unit, 0.0) + max(n.time_us, 1e-3)
    cands = [u for u in out.unit_bytes if u not in out.persistent and out.unit_bytes[u] > 0]
    cands.sort(key=lambda u: (-(gather_t.get(u, 0.0) * uses.get(u, 1)) / out.unit_bytes[u], out.unit_bytes[u]))
    chosen = []
    for u in cands:
        b = out.unit_bytes[u]
        if b <= headroom:
            out.persistent.add(u)
            headroom -= b
            chosen.append(u)
    out.meta["selective_gather"] = chosen
    return out
