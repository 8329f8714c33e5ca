"""Public utils surface (reference: deepspeed/utils/__init__.py)."""
from .logging import logger, log_dist  # noqa: F401
from .comms_logging import get_caller_func  # noqa: F401
from .init_on_device import OnDevice  # noqa: F401
from .groups import *  # noqa: F401,F403
from .nvtx import instrument_w_nvtx  # noqa: F401
from .tensor_fragment import (safe_get_full_fp32_param, safe_get_full_grad, safe_get_full_optimizer_state,  # noqa: F401
                              safe_set_full_fp32_param, safe_set_full_grad, safe_set_full_optimizer_state,
                              safe_get_local_fp32_param, safe_get_local_grad, safe_get_local_optimizer_state,
                              safe_set_local_fp32_param, safe_set_local_grad, safe_set_local_optimizer_state,
                              safe_update_full_grad_vectorized, get_full_hp_param, set_full_hp_param)
from .z3_leaf_module import set_z3_leaf_modules, unset_z3_leaf_modules, get_z3_leaf_modules, z3_leaf_module, z3_leaf_parameter  # noqa: F401
from .mixed_precision_linkage import link_hp_params, lazy_init_hp_params_optimizer_state  # noqa: F401
from .numa import get_numactl_cmd  # noqa: F401
from .timer import SynchronizedWallClockTimer, ThroughputTimer, NoopTimer  # noqa: F401


def __getattr__(name):
    if name == "RepeatingLoader":
        from ..runtime.dataloader import RepeatingLoader
        return RepeatingLoader
    raise AttributeError(name)
