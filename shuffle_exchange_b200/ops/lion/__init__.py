from .fused_lion import FusedLion
from .cpu_lion import DeepSpeedCPULion
