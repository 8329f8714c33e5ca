"""``DeepSpeedNebulaConfig`` (reference: ``deepspeed/nebula/config.py:11-43``)."""
from ..runtime.config_utils import DeepSpeedConfigObject, get_scalar_param
from .constants import *  # noqa: F401,F403
from . import constants as C


class DeepSpeedNebulaConfig(DeepSpeedConfigObject):

    def __init__(self, param_dict):
        super().__init__()
        self.enabled = None
        self.persistent_storage_path = None
        self.persistent_time_interval = None
        self.num_of_version_in_retention = None
        self.enable_nebula_load = None
        self.load_path = None
        self._initialize(param_dict.get(C.NEBULA, {}) if C.NEBULA in param_dict else {})

    def _initialize(self, d):
        self.enabled = get_scalar_param(d, C.NEBULA_ENABLED, C.NEBULA_ENABLED_DEFAULT)
        self.load_path = get_scalar_param(d, C.NEBULA_LOAD_PATH, C.NEBULA_LOAD_PATH_DEFAULT)
        self.enable_nebula_load = get_scalar_param(d, C.NEBULA_ENABLE_NEBULA_LOAD, C.NEBULA_ENABLE_NEBULA_LOAD_DEFAULT)
        self.persistent_storage_path = get_scalar_param(d, C.NEBULA_PERSISTENT_STORAGE_PATH,
                                                        C.NEBULA_PERSISTENT_STORAGE_PATH_DEFAULT)
        self.persistent_time_interval = get_scalar_param(d, C.NEBULA_PERSISTENT_TIME_INTERVAL,
                                                         C.NEBULA_PERSISTENT_TIME_INTERVAL_DEFAULT)
        self.num_of_version_in_retention = get_scalar_param(d, C.NEBULA_NUM_OF_VERSION_IN_RETENTION,
                                                            C.NEBULA_NUM_OF_VERSION_IN_RETENTION_DEFAULT)
