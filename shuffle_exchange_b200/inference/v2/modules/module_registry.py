"""Per-interface registries of named implementations (reference ``inference/v2/modules/module_registry.py``)."""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Any, Dict, Type

from ....runtime.config_utils import DeepSpeedConfigModel
from .ds_module import DSModuleBase


class ConfigBundle(DeepSpeedConfigModel):
    """A module implementation name, its module config and optional implementation-specific settings."""
    name: str
    config: Any
    implementation_config: Dict = {}


class DSModuleRegistryBase(ABC):
    """Subclasses define ``registry: Dict[str, Type[DSModuleBase]]`` and ``associated_class()``."""
    registry: Dict[str, Type[DSModuleBase]] = {}

    @classmethod
    def instantiate_config(cls, config_bundle: ConfigBundle) -> DSModuleBase:
        if config_bundle.name not in cls.registry:
            raise KeyError(f"Unknown DSModule: {config_bundle.name}, cls.registry={sorted(cls.registry)}")
        target = cls.registry[config_bundle.name]
        if not target.supports_config(config_bundle.config):
            raise ValueError(f"Config {config_bundle.config} is not supported by {target}")
        return target(config_bundle.config, config_bundle.implementation_config)

    @staticmethod
    @abstractmethod
    def associated_class() -> Type[DSModuleBase]:
        ...

    @classmethod
    def register_module(cls, child_class: Type[DSModuleBase]) -> Type[DSModuleBase]:
        if not issubclass(child_class, cls.associated_class()):
            raise TypeError(f"Can only register subclasses of {cls.associated_class()}, got {child_class}")
        cls.registry[child_class.name()] = child_class
        return child_class
