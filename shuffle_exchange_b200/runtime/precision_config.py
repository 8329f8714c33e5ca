"""fp16 / bf16 config sections under the reference's names (reference: ``runtime/precision_config.py``)."""
from .config import BF16Config as DeepSpeedBF16Config  # noqa: F401
from .config import FP16Config as DeepSpeedFP16Config  # noqa: F401

# "bf16": {"enabled": true, "immediate_grad_update": true, "check_grad_overflow": false}
BFLOAT16 = "bf16"
BFLOAT16_OLD = "bfloat16"          # kept for configs written before the key was shortened
# "fp16": {"enabled": true, "auto_cast": false, "loss_scale": 0, "initial_scale_power": 16, "loss_scale_window": 1000,
#          "hysteresis": 2, "consecutive_hysteresis": false, "min_loss_scale": 1}
FP16 = "fp16"


def get_bfloat16_config(param_dict) -> DeepSpeedBF16Config:
    d = param_dict.get(BFLOAT16, param_dict.get(BFLOAT16_OLD, {}))
    return DeepSpeedBF16Config(**d)


def get_float16_config(param_dict) -> DeepSpeedFP16Config:
    return DeepSpeedFP16Config(**param_dict.get(FP16, {}))
