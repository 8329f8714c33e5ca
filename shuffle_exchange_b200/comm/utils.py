"""Launcher-environment probes and call-introspection helpers for the comm logger (reference: ``comm/utils.py``)."""
from __future__ import annotations

import inspect
import os


def _first_env(names, default=None):
    for n in names:
        v = os.environ.get(n)
        if v is not None:
            return v
    return default


def get_local_rank_from_launcher() -> int:
    """torchrun / the sxb launcher set LOCAL_RANK; Open MPI sets OMPI_COMM_WORLD_LOCAL_RANK; a bare ``python`` run is rank 0."""
    return int(_first_env(("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "MV2_COMM_WORLD_LOCAL_RANK", "SLURM_LOCALID"), 0))


def get_world_rank_from_launcher() -> int:
    return int(_first_env(("RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", "SLURM_PROCID"), 0))


def get_world_size_from_launcher() -> int:
    size = int(_first_env(("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "SLURM_NTASKS"), 1))
    if get_world_rank_from_launcher() == 0:
        print(f"set world size to {size}")
    return size


def get_default_args(func) -> dict:
    return {k: v.default for k, v in inspect.signature(func).parameters.items() if v.default is not inspect.Parameter.empty}


def get_tensor_position(func) -> int:
    """Index of the positional parameter that carries the payload of a collective."""
    params = list(inspect.signature(func).parameters)
    for name in ("tensor", "tensors", "input_list", "input_tensor_list", "input_tensor", "input", "output_tensor"):
        if name in params:
            return params.index(name)
    return -1


def get_tensor_kwarg(func, kwargs):
    merged = {**get_default_args(func), **kwargs}
    for name in ("tensor", "tensors", "input_list", "input_tensor_list", "input_tensor", "input"):
        if merged.get(name) is not None:
            return merged[name]
    return None


def get_msg_size_from_args(func, *args, **kwargs) -> int:
    """Bytes moved by a collective call, from whichever argument holds the tensor(s)."""
    pos = get_tensor_position(func)
    payload = args[pos] if 0 <= pos < len(args) else get_tensor_kwarg(func, kwargs)
    if payload is None:
        return 0
    if isinstance(payload, (list, tuple)):
        return sum(t.element_size() * t.nelement() for t in payload)
    return payload.element_size() * payload.nelement()


def get_debug_log_name(func_args: dict, debug) -> str:
    return f"{func_args['log_name']} | [Caller Func: {debug}]" if debug else func_args["log_name"]
