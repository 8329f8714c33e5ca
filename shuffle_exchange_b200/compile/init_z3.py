"""ZeRO-3 entry of the schedule compiler (reference ``deepspeed/compile/init_z3.py:21-98``)."""
from __future__ import annotations

from .backend import default_schedule, make_backend


def init_z3(engine, backend, compile_config, compile_kwargs, schedule=None):
    opt = engine.optimizer
    if compile_config.symmetric_memory and hasattr(opt, "use_symmetric_memory"):
        opt.use_symmetric_memory = True          # gathers go through the peer-memory all-gather kernels
    if hasattr(opt, "empty_partition_cache"):
        opt.empty_partition_cache()
    return make_backend(backend, compile_config, compile_kwargs, engine=engine,
                        schedule=schedule if schedule is not None else default_schedule(3, compile_config), stage=3)
