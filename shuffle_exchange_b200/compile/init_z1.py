"""ZeRO-1/2 entry of the schedule compiler (reference ``deepspeed/compile/init_z1.py:18-``)."""
from __future__ import annotations

from .backend import default_schedule, make_backend


def init_z1(engine, backend, compile_config, compile_kwargs, schedule=None, use_z2=False):
    return make_backend(backend, compile_config, compile_kwargs, engine=engine,
                        schedule=schedule if schedule is not None else default_schedule(2 if use_z2 else 1,
                                                                                        compile_config),
                        stage=2 if use_z2 else 1)
