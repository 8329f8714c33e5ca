"""Optimisation passes of the schedule compiler.  Every pass is ``fn(plan, ctx) -> plan`` where ``ctx`` is a
:class:`~..backend.PassContext` (engine, compile config, memory budget, profiles)."""
from . import (offload_activation, offload_adam_states, offload_parameters, prefetch, selective_gather,
               zero1_compile, zero3_compile)

PASSES = {
    "zero3_compile": zero3_compile.add_z3_gather_release,
    "zero1_compile": zero1_compile.add_z1_reduce,
    "prefetch": prefetch.schedule_prefetch,
    "selective_gather": selective_gather.selective_gather,
    "offload_adam_states": offload_adam_states.offload_adam_states,
    "offload_adam_states_sync": offload_adam_states.offload_adam_states_sync,
    "offload_parameters": offload_parameters.offload_parameter_fwd,
    "offload_activation": offload_activation.offload_activation_fwd,
}
