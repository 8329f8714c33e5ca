from .autotuner import Autotuner
from .config import DeepSpeedAutotuningConfig
