"""Schedule compiler ("DeepCompile" equivalent).

The reference (``deepspeed/compile/**``) is a ``torch.compile`` backend that rewrites FX graphs to insert
``dc.allgather_param / wait_allgather / release_param / reduce_grad`` ops and then runs optimisation passes (prefetch
scheduling, selective gather, optimizer-state / parameter / activation offload) driven by graph and communication
profiles.  A tracing compiler is not this framework's hot path: kernels are hand-written and launch-bound regions are
captured in CUDA graphs.  What DeepCompile actually *decides* — when every shard is gathered, released and reduced, what
stays resident, what is offloaded — is independent of FX, so it is expressed here on the ZeRO unit trace the engine
already records:

* :mod:`.plan`           — the IR: an ordered list of gather / compute / release / reduce nodes with byte sizes and
  measured times, plus a memory simulator;
* :mod:`.profilers`      — per-unit compute time (CUDA events) and collective bandwidth tables;
* :mod:`.list_schedule`  — memory-bounded list scheduling of the gather nodes (``fast_free_schedule``,
  ``simple_prefetch``);
* :mod:`.passes`         — prefetch, selective (persistent) gather, optimizer-state / parameter / activation offload;
* :mod:`.backend`        — pass registry + ``launch_compile_passes`` stepping through a schedule of passes;
* :mod:`.cuda_graph`     — static-shape capture/replay of a callable (the "compiled function").
"""
from .config import CompileConfig
from .backend import (init_schedule, launch_compile_passes, make_backend, register_compile_pass, run_opt_passes)
from .plan import ExecutionPlan, PlanNode, build_plan_from_engine
from .cuda_graph import GraphedCallable
from .util import deepcompile_backward_prologue, get_deepcompile_handle, is_deepcompile_supported
