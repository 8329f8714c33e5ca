"""Pluggable inference modules (reference ``deepspeed/inference/v2/modules``)."""
from . import heuristics
from .configs import *  # noqa: F401,F403
from .ds_module import DSModuleBase, DSModuleConfig
from .interfaces import *  # noqa: F401,F403
from .module_registry import ConfigBundle, DSModuleRegistryBase
