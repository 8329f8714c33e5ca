from .fused_adam import FusedAdam
from .cpu_adam import DeepSpeedCPUAdam
from .multi_tensor_apply import MultiTensorApply
