"""``zero`` namespace (reference: deepspeed/runtime/zero/__init__.py)."""
from . import partition_parameters  # noqa: F401
from .partition_parameters import (GatheredParameters, Init, ZeroParamStatus, ZeroParamType,  # noqa: F401
                                   register_external_parameter, unregister_external_parameter)
from .config import DeepSpeedZeroConfig, ZeroStageEnum  # noqa: F401


def __getattr__(name):
    import importlib
    table = {
        "TiledLinear": (".tiling", "TiledLinear"),
        "TiledLinearReturnBias": (".tiling", "TiledLinearReturnBias"),
        "MiCS_Init": (".mics", "MiCS_Init"),
        "estimate_zero2_model_states_mem_needs": (".stage_1_and_2", "estimate_zero2_model_states_mem_needs"),
        "estimate_zero2_model_states_mem_needs_all_live": (".stage_1_and_2", "estimate_zero2_model_states_mem_needs_all_live"),
        "estimate_zero2_model_states_mem_needs_all_cold": (".stage_1_and_2", "estimate_zero2_model_states_mem_needs_all_cold"),
        "estimate_zero3_model_states_mem_needs": (".stage3", "estimate_zero3_model_states_mem_needs"),
        "estimate_zero3_model_states_mem_needs_all_live": (".stage3", "estimate_zero3_model_states_mem_needs_all_live"),
        "estimate_zero3_model_states_mem_needs_all_cold": (".stage3", "estimate_zero3_model_states_mem_needs_all_cold"),
    }
    if name in table:
        mod, attr = table[name]
        return getattr(importlib.import_module(mod, __name__), attr)
    raise AttributeError(name)
