"""Base ZeRO-3 pass: canonical gather-before-use / release-after-use placement (reference
``compile/passes/zero3_compile.py`` ``add_z3_gather_release``)."""
from ..list_schedule import fast_free_schedule


def add_z3_gather_release(plan, ctx):
    out = fast_free_schedule(plan, ctx.available_mem)
    out.meta["z3"] = True
    return out
