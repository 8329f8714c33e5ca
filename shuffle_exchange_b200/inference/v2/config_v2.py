"""Configuration of the ragged inference engine (reference: ``inference/v2/config_v2.py:15-43``)."""
from __future__ import annotations

from dataclasses import dataclass, field

from .ragged.ragged_manager import DSStateManagerConfig


@dataclass
class DeepSpeedTPConfig:
    tp_size: int = 1


@dataclass
class QuantizationConfig:
    quantization_mode: str = None      # None | "wf6af16" | "fp8"


@dataclass
class RaggedInferenceEngineConfig:
    tensor_parallel: DeepSpeedTPConfig = field(default_factory=DeepSpeedTPConfig)
    state_manager: DSStateManagerConfig = field(default_factory=DSStateManagerConfig)
    quantization: QuantizationConfig = field(default_factory=QuantizationConfig)
    kv_block_size: int = 128
