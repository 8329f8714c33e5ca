"""Parameter offload (reference ``compile/passes/offload_parameters.py`` ``offload_parameter_fwd``): the bf16 SHARDS of
units far from their next use live in pinned host memory; a RELOAD is placed one gather ahead of the GATHER that needs
them so PCIe latency hides behind the previous unit's compute."""
from ..plan import GATHER, OFFLOAD, RELOAD, PlanNode


def offload_parameter_fwd(plan, ctx):
    out = plan.clone()
    world = max(1, ctx.world_size)
    nodes, out_nodes = out.nodes, []
    offloaded = []
    for i, n in enumerate(nodes):
        if n.kind == GATHER and n.unit not in out.persistent and n.phase == "fwd":
            shard = out.unit_bytes[n.unit] // world
            out_nodes.append(PlanNode(RELOAD, n.unit, n.phase, shard, note="shard <- host"))
            offloaded.append(n.unit)
        out_nodes.append(n)
    for u in dict.fromkeys(offloaded):
        out_nodes.append(PlanNode(OFFLOAD, u, "bwd", out.unit_bytes[u] // world, note="shard -> host"))
    out.nodes = out_nodes
    out.meta["offload_parameters"] = list(dict.fromkeys(offloaded))
    return out
