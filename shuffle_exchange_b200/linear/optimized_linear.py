"""``OptimizedLinear``: factory for plain / LoRA / quantised-base linear layers.

Reference: ``deepspeed/linear/optimized_linear.py:18-222``.  ``LoRAOptimizedLinear`` keeps the frozen base weight
sharded over ``base_weight_sharding`` ranks (``1/n`` of the rows... the reference shards the INPUT dimension) and
all-gathers it for every forward; adapters ``lora_weight_1: [in -> r]`` and ``lora_weight_2: [r -> out]`` are the
only trainable tensors; output = ``x W_base^T + (alpha / r) * (x A^T) B^T``.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as TF
from torch.nn.modules.linear import Linear as _TorchLinear   # immune to linear.Init's nn.Linear patch

from .. import comm as dist
from .config import LoRAConfig, QuantizationConfig
from .quantization import QuantizedLinear, QuantizedParameter


class OptimizedLinear(nn.Module):
    """``OptimizedLinear(in, out, lora_config=..., quantization_config=...)`` returns the cheapest layer that provides
    the requested features: ``nn.Linear``, :class:`QuantizedLinear` or :class:`LoRAOptimizedLinear`."""

    def __new__(cls, input_dim: int, output_dim: int, lora_config: LoRAConfig = None,
                quantization_config: QuantizationConfig = None, dtype=torch.bfloat16, linear_cls=nn.Linear):
        if quantization_config is not None and not torch.cuda.is_available():
            raise RuntimeError("quantized linear layers need a CUDA device")
        if lora_config is None and quantization_config is None:
            return _TorchLinear(input_dim, output_dim, bias=False, dtype=dtype)
        if lora_config is not None:
            return LoRAOptimizedLinear(input_dim=input_dim, output_dim=output_dim, lora_config=lora_config,
                                       quantization_config=quantization_config, dtype=dtype)
        return QuantizedLinear(input_dim=input_dim, output_dim=output_dim, quantization_config=quantization_config,
                               dtype=dtype)


class LoRAOptimizedLinear(nn.Module):
    def __init__(self, input_dim: int, output_dim: int, lora_config: LoRAConfig = None,
                 quantization_config: QuantizationConfig = None, device=None, dtype=torch.bfloat16):
        super().__init__()
        self.input_dim, self.output_dim = input_dim, output_dim
        self.lora_config, self.quantization_config = lora_config, quantization_config
        self.device = device if device is not None else ("cuda" if torch.cuda.is_available() else "cpu")
        self.dtype = dtype
        assert lora_config is not None, "DSOptimizedLinear requires a LoRA config"
        assert lora_config.lora_r > 0, "lora_r must be positive"
        self.zero_shards = lora_config.base_weight_sharding
        assert input_dim % self.zero_shards == 0, "base weight sharding must divide the input dimension"
        self.sharded_weight_size = input_dim // self.zero_shards
        shard = torch.empty(output_dim, self.sharded_weight_size, dtype=dtype)
        if quantization_config is not None:
            self.weight = QuantizedParameter(shard, quantization_config=quantization_config)
        else:
            self.weight = nn.Parameter(shard, requires_grad=False)
        self.disabled = False
        self.lora_weight_1 = self.lora_weight_2 = None
        if not lora_config.delay_lora_init:
            self.init_lora()

    def disable(self):
        """Turn the layer into its frozen base (adapters dropped)."""
        self.disabled = True
        self.weight = nn.Parameter(self.weight.data, requires_grad=False)
        self.lora_weight_1 = self.lora_weight_2 = None

    def init_lora(self):
        if self.disabled:
            return
        if self.quantization_config is not None and not isinstance(self.weight, QuantizedParameter):
            self.weight = QuantizedParameter(self.weight.data, quantization_config=self.quantization_config)
        r = self.lora_config.lora_r
        dev = self.weight.device if self.weight.device.type != "meta" else self.device
        self.lora_scaling_factor = self.lora_config.lora_alpha / r
        self.lora_weight_1 = _TorchLinear(self.input_dim, r, bias=False, device=dev, dtype=self.dtype)
        self.lora_weight_2 = _TorchLinear(r, self.output_dim, bias=False, device=dev, dtype=self.dtype)
        nn.init.kaiming_uniform_(self.lora_weight_1.weight, a=math.sqrt(5))
        nn.init.zeros_(self.lora_weight_2.weight)           # adapters start as the identity update
        self.weight.requires_grad = False

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
        """Accept a FULL base weight ``[out, in]`` and keep only this rank's input-dimension shard."""
        key = prefix + "weight"
        if key in state_dict:
            w = state_dict[key]
            if w.dim() == 2 and w.shape[1] == self.input_dim and self.zero_shards > 1:
                rank = dist.get_rank() % self.zero_shards if dist.is_initialized() else 0
                state_dict[key] = w[:, rank * self.sharded_weight_size: (rank + 1) * self.sharded_weight_size].contiguous()
        if self.lora_weight_1 is None:
            for k in [k for k in state_dict if k.startswith(prefix + "lora_weight_")]:
                state_dict.pop(k)
        super()._load_from_state_dict(state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs)

    def full_weight(self) -> torch.Tensor:
        base = self.weight.dequantized().view(self.output_dim, self.sharded_weight_size).to(self.dtype) \
            if isinstance(self.weight, QuantizedParameter) else self.weight.data
        if self.zero_shards <= 1 or not dist.is_initialized():
            return base
        parts = [torch.empty_like(base) for _ in range(self.zero_shards)]
        dist.all_gather(parts, base.contiguous())
        return torch.cat(parts, dim=1)

    def forward(self, input_tensor):
        if self.disabled:
            return TF.linear(input_tensor, self.weight)
        assert self.lora_weight_1 is not None and self.lora_weight_2 is not None, "init_lora was never called, please initialize lora weights first"
        with torch.no_grad():
            base = TF.linear(input_tensor, self.full_weight())
        lora = self.lora_weight_2(self.lora_weight_1(input_tensor))
        return base + self.lora_scaling_factor * lora
