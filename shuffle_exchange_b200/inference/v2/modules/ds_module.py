"""Base classes of pluggable inference modules (reference ``inference/v2/modules/ds_module.py``): every module kind
(attention, linear, MoE, norms, embedding, unembedding) has a pydantic config, an abstract interface and one or more
named implementations; ``supports_config`` lets a registry pick an implementation for a concrete config."""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Any, Dict, Optional

import torch

from ....runtime.config_utils import DeepSpeedConfigModel


class DSModuleConfig(DeepSpeedConfigModel):
    max_tokens: int


class DSModuleBase(torch.nn.Module, ABC):
    """``name()`` identifies the implementation, ``config_class()`` its config type and ``supports_config(config)``
    whether this implementation can run that config on this device."""

    @staticmethod
    @abstractmethod
    def config_class():
        ...

    @staticmethod
    @abstractmethod
    def name() -> str:
        ...

    @staticmethod
    @abstractmethod
    def supports_config(config: DSModuleConfig) -> bool:
        ...

    def __init__(self, config: DSModuleConfig, implementation_config: Optional[Dict[str, Any]] = None) -> None:
        super().__init__()
        self._config = config
        self._implementation_config = implementation_config or {}
