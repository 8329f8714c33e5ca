from .config import AUTOTP_MODE, TPConfig, TPTrainingConfig, get_tensor_parallel_config
from .tp_manager import TpTrainingManager
