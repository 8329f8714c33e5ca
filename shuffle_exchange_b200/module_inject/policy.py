"""Policy base classes and checkpoint-copy helpers (reference: ``module_inject/policy.py``).

``TransformerPolicy`` lives in ``containers/base.py``; this module keeps the reference's import path and the small
helpers the containers' checkpoint loaders share: they copy a tensor (optionally a fused group of tensors) from a state
dict into a parameter of the fused layer, transposing into its ``[in, out]`` layout, slicing for tensor parallelism and
quantising when a weight quantizer is supplied."""
from __future__ import annotations

from typing import Optional, Sequence

import torch

from .containers.base import DSPolicy, TransformerPolicy  # noqa: F401


def transpose(data: torch.Tensor) -> torch.Tensor:
    """Materialised 2-D transpose (the fused inference layer computes ``x @ W``)."""
    return data.t().contiguous() if data.dim() == 2 else data


def _transpose(x: torch.Tensor, heads: int = 1, mp_replace=None) -> torch.Tensor:
    """Megatron-v2 fused QKV rows are ``[heads, 3, d]``; regroup to ``[3, heads, d]`` (q|k|v blocks)."""
    heads = heads // (mp_replace.mp_size if mp_replace is not None else 1)
    d = x.shape[-1] // (3 * heads) if x.dim() == 1 else x.shape[0] // (3 * heads)
    if x.dim() == 1:
        return x.view(heads, 3, d).permute(1, 0, 2).reshape(-1).contiguous()
    return x.view(heads, 3, d, x.shape[-1]).permute(1, 0, 2, 3).reshape(x.shape[0], x.shape[-1]).contiguous()


def _assign(module, dst_name: str, value: torch.Tensor, weight_quantizer=None, mp_replace=None, transpose_2d=True,
            qkv=False, allocate_tensor=False):
    dst = getattr(module, dst_name)
    if value.dim() == 2 and transpose_2d:
        value = transpose(value)
    if weight_quantizer is not None and value.dim() == 2:
        value = weight_quantizer.quantize(value)
    if mp_replace is not None:
        new = (mp_replace.strided_copy(dst, value, num_splits=3, int8=getattr(weight_quantizer, "q_int8", False),
                                       allocate_tensor=allocate_tensor) if qkv else
               mp_replace.copy(dst, value, int8=getattr(weight_quantizer, "q_int8", False),
                               allocate_tensor=allocate_tensor))
    else:
        new = value
    if isinstance(dst, torch.nn.Parameter) and dst.shape == new.shape:
        dst.data.copy_(new.to(dst.dtype))
    else:
        setattr(module, dst_name, torch.nn.Parameter(new.detach(), requires_grad=False))


def maybe_copy(module, sd, weight_quantizer, mp_replace, dst_name: str, src_name: str, qkv: bool = False,
               megatron_v2: bool = False, split_qkv: bool = False, heads: int = 1) -> None:
    """Copy ``sd[src_name]`` into ``module.<dst_name>`` if the key exists."""
    if src_name not in sd:
        return
    value = sd[src_name]
    if qkv and megatron_v2:
        value = _transpose(value, heads=heads, mp_replace=mp_replace)
    _assign(module, dst_name, value, weight_quantizer, mp_replace, qkv=qkv and split_qkv)


def maybe_copy_qkv(module, sd, weight_quantizer, mp_replace, dst_name: str, src_names: Sequence[str],
                   split_qkv: bool = False) -> None:
    """Fuse separate q, k, v tensors (``src_names``) along the output dimension and copy."""
    if src_names[0] not in sd:
        return
    fused = torch.cat([sd[n] for n in src_names], dim=0)
    _assign(module, dst_name, fused, weight_quantizer, mp_replace, qkv=split_qkv)


def maybe_copy_geglu(module, sd, weight_quantizer, mp_replace, dst_name: str, src_names: Sequence[str]) -> None:
    """Fuse the two halves of a gated MLP input projection (``src_names = (up, gate)``) and copy."""
    if src_names[0] not in sd:
        return
    fused = torch.cat([sd[n] for n in src_names], dim=0)
    _assign(module, dst_name, fused, weight_quantizer, mp_replace)


def pack_lora_weights(p) -> list:
    """``[lora_right, lora_left, scaling]`` of a LoRA-wrapped parameter holder, or ``[]``."""
    if hasattr(p, "lora_right_weight"):
        return [p.lora_right_weight, p.lora_left_weight, p.lora_scaling]
    return []


def maybe_get_lora(p) -> list:
    return pack_lora_weights(p)
