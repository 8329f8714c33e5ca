"""ZeRO-1/2 pass: gradients are reduced as soon as the owning unit's backward finished (reference
``compile/passes/zero1_compile.py`` ``add_z1_reduce``).  There is nothing to gather, so the plan only carries REDUCE
nodes; applying it turns on the engine's overlapped reduction."""
from ..plan import COMPUTE, REDUCE, PlanNode


def add_z1_reduce(plan, ctx):
    out = plan.clone()
    nodes = []
    for n in out.nodes:
        if n.kind in (COMPUTE, REDUCE):
            nodes.append(n)
    out.nodes = nodes
    out.persistent = set(out.unit_bytes)          # parameters are never partitioned
    out.meta["z1"] = True
    opt = getattr(ctx.engine, "optimizer", None) if ctx.engine is not None else None
    if opt is not None and hasattr(opt, "overlap_comm"):
        opt.overlap_comm = True
    return out
