"""Prefetch scheduling (reference ``compile/passes/prefetch.py`` ``schedule_prefetch``): hoist all-gathers ahead of
their consumers as far as the memory budget allows, fusing look-ahead gathers up to a byte cap so one launch latency is
paid per bucket."""
from ..list_schedule import simple_prefetch

MAX_FUSE_SIZE = int(1e9)       # bytes; reference uses the same 1e9 cap for fused all-gather buckets
MARGIN = 0.1                   # fraction of the budget kept free for allocator fragmentation


def schedule_prefetch(plan, ctx):
    budget = int(ctx.available_mem * (1.0 - MARGIN))
    out = simple_prefetch(plan, budget, ctx.output_size, ctx.config.debug_log if ctx.config else False)
    out.meta["prefetch_budget"] = budget
    out.meta["max_fuse_size"] = MAX_FUSE_SIZE
    return out
