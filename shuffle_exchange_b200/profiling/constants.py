"""Key names and defaults of the ``"flops_profiler"`` config section (reference: ``profiling/constants.py``).

``FLOPS_PROFILER_<KEY>`` is the JSON key, ``FLOPS_PROFILER_<KEY>_DEFAULT`` its default; both are generated from the
one table below so the two can never drift apart."""
FLOPS_PROFILER = "flops_profiler"

_FIELDS = {
    "enabled": False,              # run the profiler at `profile_step`
    "recompute_fwd_factor": 0.0,   # extra forward FLOPs counted for activation recomputation
    "profile_step": 1,             # global step that is profiled
    "module_depth": -1,            # depth of the per-module table (-1: innermost)
    "top_modules": 1,              # how many modules per depth in the aggregated table
    "detailed": True,              # print the per-module tree
    "output_file": None,           # None: stdout
}

for _key, _default in _FIELDS.items():
    globals()[f"FLOPS_PROFILER_{_key.upper()}"] = _key
    globals()[f"FLOPS_PROFILER_{_key.upper()}_DEFAULT"] = _default
del _key, _default
