"""Group-wise asymmetric quantiser / dequantiser + small model-surgery helpers.

Reference: ``deepspeed/inference/quantization/utils.py:43-288`` — config keys ``num_bits`` (4|8), ``group_size``,
``group_dim``, ``symmetric`` (must be False).  ``q = round((x - min) * (2^bits - 1) / (max - min))`` per group, INT4
packs two codes per byte (even index in the high nibble).  Dequantisation is ``q / scale + min`` in the compute dtype."""
from __future__ import annotations

from typing import Callable, Dict, Tuple

import torch
from torch import Tensor, nn


def _grouped_shape(shape, group_dim: int, group_size: int):
    n = shape[group_dim] // group_size
    return tuple(shape[:group_dim]) + (n, group_size) + tuple(shape[group_dim + 1:])


def _check(config: Dict) -> None:
    assert config["num_bits"] in (4, 8), "Only INT4 and INT8 quantization is supported."
    assert config["symmetric"] is False, "Only asymmetric quantization is supported at this moment."


class Quantizer:

    def __init__(self, config: Dict) -> None:
        _check(config)
        self.config = config

    def quantize(self, tensor: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
        """-> ``(codes uint8 [packed for 4 bits], scale, min)``; ``scale``/``min`` keep a singleton group-element dim."""
        c = self.config
        gd, gs = c["group_dim"], c["group_size"]
        assert tensor.shape[gd] % gs == 0, f"Tensor shape: {tensor.shape} quantization config {c}"
        shape = tensor.shape
        grouped = tensor.detach().float().reshape(_grouped_shape(shape, gd, gs))
        codes, scale, mn = self._quantize_int8(grouped)
        codes = codes.reshape(shape)
        if c["num_bits"] == 4:
            codes = self._compress_uint8_to_uint4(codes)
        return codes, scale.to(tensor.dtype), mn.to(tensor.dtype)

    def _quantize_int8(self, tensor: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
        q_range = 2 ** self.config["num_bits"] - 1
        dim = self.config["group_dim"] + 1
        mn = tensor.amin(dim=dim, keepdim=True)
        mx = tensor.amax(dim=dim, keepdim=True)
        scale = q_range / (mx - mn).clamp_min(1e-8)
        codes = ((tensor - mn) * scale).clamp_(0, q_range).round_().to(torch.uint8)
        return codes, scale, mn

    def _compress_uint8_to_uint4(self, tensor: Tensor) -> Tensor:
        assert tensor.shape[-1] % 2 == 0
        return (tensor[..., 0::2] << 4) | tensor[..., 1::2]


class DeQuantizer:

    def __init__(self, config: Dict, dtype: torch.dtype) -> None:
        _check(config)
        self.config, self.dtype = config, dtype

    def dequantize(self, tensor: Tensor, quant_scale: Tensor, quant_min: Tensor) -> Tensor:
        c = self.config
        if c["num_bits"] == 4:
            tensor = self._decompress_uint4_to_uint8(tensor)
        shape = tensor.shape
        grouped = tensor.reshape(_grouped_shape(shape, c["group_dim"], c["group_size"]))
        return self._dequantize_int8(grouped, quant_scale, quant_min).reshape(shape)

    def _dequantize_int8(self, tensor: Tensor, quant_scale: Tensor, quant_min: Tensor) -> Tensor:
        assert tensor.dtype == torch.uint8
        return (tensor.to(torch.float32) / quant_scale.float() + quant_min.float()).to(self.dtype)

    def _decompress_uint4_to_uint8(self, tensor: Tensor) -> Tensor:
        out = torch.empty(tensor.shape[:-1] + (tensor.shape[-1] * 2,), dtype=torch.uint8, device=tensor.device)
        out[..., 0::2] = tensor >> 4
        out[..., 1::2] = tensor & 0xF
        return out


def tensor_clamp(tensor: Tensor, min, max) -> Tensor:
    return tensor.clamp_(min, max)


def tensor_round(tensor: Tensor) -> Tensor:
    return tensor.round_()


def get_AsyncPartitionedParameterSwapper(model: nn.Module):
    for p in model.parameters():
        sw = getattr(p, "nvme_swapper", None)
        if sw is not None:
            return sw
    return None


def recursive_setattr(model: nn.Module, module_name: str, module: nn.Module) -> None:
    """``recursive_setattr(m, "a.b.c", x)`` == ``m.a.b.c = x`` (integer path components index containers)."""
    parts = module_name.split(".")
    parent = model
    for p in parts[:-1]:
        parent = getattr(parent, p)
    setattr(parent, parts[-1], module)


def concat_to_compat_param(quantized_weight: Tensor, quant_scale: Tensor, quant_min: Tensor,
                           return_param: bool = True):
    """Pack codes, scales and mins into ONE flat byte tensor so the triple can live in a single (ZeRO-partitionable)
    parameter; returns it with a closure that recovers the three views."""
    shape_w, shape_s, shape_m = quantized_weight.shape, quant_scale.shape, quant_min.shape
    dt = quant_scale.dtype
    flat = torch.cat([quantized_weight.reshape(-1), quant_scale.contiguous().view(torch.uint8).reshape(-1),
                      quant_min.contiguous().view(torch.uint8).reshape(-1)])

    def deconcat(compat: Tensor) -> Tuple[Tensor, Tensor, Tensor]:
        nw = shape_w.numel()
        ns = shape_s.numel() * dt.itemsize
        w = compat[:nw].reshape(shape_w)
        s = compat[nw:nw + ns].view(dt).reshape(shape_s)
        m = compat[nw + ns:].view(dt).reshape(shape_m)
        return w, s, m

    if return_param:
        flat = nn.Parameter(flat, requires_grad=False)
    return flat, deconcat


def _quantize_param(param: nn.Parameter, quant_config: Dict):
    assert not hasattr(param, "weight_quantized"), "Parameter has already been quantized."
    q = Quantizer(quant_config)
    dq = DeQuantizer(quant_config, param.dtype)
    codes, scale, mn = q.quantize(param.data)
    param.data = codes.view(param.dtype) if codes.element_size() == param.element_size() else codes
    param.quant_scale, param.quant_min = scale, mn
    param.dequantizer = dq
    param.weight_quantized = True


def wrap_quantized_functional(f: Callable) -> Callable:
    """Make ``F.linear``-like functionals accept parameters quantised by :func:`_quantize_param`."""

    def wrapper(input: Tensor, weight: nn.Parameter, *args, **kwargs) -> Tensor:
        if getattr(weight, "weight_quantized", False):
            w = weight.dequantizer.dequantize(weight.data.view(torch.uint8), weight.quant_scale, weight.quant_min)
            return f(input, w, *args, **kwargs)
        return f(input, weight, *args, **kwargs)

    return wrapper


def wrap_load_from_state_dict(f: Callable) -> Callable:
    """Quantise matching parameters right after a state dict is loaded into a module."""

    def wrapper(model, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
        replaced = {}
        for name, p in model.named_parameters(recurse=False):
            if getattr(p, "weight_quantized", False):
                replaced[name] = p
                setattr(model, name, nn.Parameter(torch.empty_like(state_dict[prefix + name]), requires_grad=False))
        f(model, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs)
        for name, old in replaced.items():
            new = getattr(model, name)
            _quantize_param(new, old.dequantizer.config)

    return wrapper
