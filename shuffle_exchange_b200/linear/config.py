"""Configuration of the optimized / LoRA / quantized linear layers (reference: ``deepspeed/linear/config.py:12-55``)."""
from dataclasses import dataclass, field
from typing import List

import torch


@dataclass
class LoRAConfig:
    """``lora_r`` rank, ``lora_alpha`` scaling, ``base_weight_sharding`` = number of ranks the frozen base weight is
    sharded over (set it to the dp world size), ``offload`` frozen weights to host when idle, ``delay_lora_init`` to
    create the adapters later with :func:`init_lora`, ``target_mods`` = module-name suffixes that receive adapters."""
    lora_r: int = 64
    lora_alpha: float = 16.0
    base_weight_sharding: int = 1
    offload: bool = False
    offload_ratio: float = 0.0
    delay_lora_init: bool = False
    target_mods: List[str] = field(default_factory=lambda: ["q_proj", "k_proj", "v_proj", "o_proj", "gate_proj",
                                                           "up_proj", "down_proj"])


@dataclass
class QuantizationConfig:
    """Group-scaled miniature-float weight format: ``q_bits`` total bits of which ``mantissa_bits`` mantissa, one fp32
    scale per ``group_size`` elements; payload stored as ``q_dtype`` bytes."""
    q_bits: int = 8
    mantissa_bits: int = 3
    group_size: int = 512
    q_dtype: torch.dtype = torch.uint8
