"""Backward-compatible access to a Megatron-style ``mpu`` object whose method names changed over time
(reference: deepspeed/utils/bwc.py:7-104)."""


def _first(mpu, *names):
    for n in names:
        if hasattr(mpu, n):
            return getattr(mpu, n)()
    raise AttributeError(f"mpu object has none of {names}")


def bwc_tensor_model_parallel_rank(mpu=None):
    return 0 if mpu is None else _first(mpu, "get_tensor_model_parallel_rank", "get_slice_parallel_rank",
                                        "get_model_parallel_rank")


def bwc_tensor_model_parallel_world_size(mpu=None):
    return 1 if mpu is None else _first(mpu, "get_tensor_model_parallel_world_size", "get_slice_parallel_world_size",
                                        "get_model_parallel_world_size")


def bwc_tensor_model_parallel_group(mpu=None):
    return None if mpu is None else _first(mpu, "get_tensor_model_parallel_group", "get_slice_parallel_group",
                                           "get_model_parallel_group")


def bwc_pipeline_parallel_world_size(mpu=None):
    if mpu is None:
        return 1
    for n in ("get_pipeline_model_parallel_world_size", "get_pipe_parallel_world_size"):
        if hasattr(mpu, n):
            return getattr(mpu, n)()
    return 1


def bwc_pipeline_parallel_group(mpu=None):
    if mpu is None:
        return None
    for n in ("get_pipeline_model_parallel_group", "get_pipe_parallel_group"):
        if hasattr(mpu, n):
            return getattr(mpu, n)()
    raise AttributeError("mpu has no pipeline parallel group accessor")
