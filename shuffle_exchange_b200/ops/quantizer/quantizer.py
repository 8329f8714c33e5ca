"""Grouped int8 / int4 quantisation (symmetric or asymmetric, optional stochastic rounding).

Reference API: ``deepspeed/ops/quantizer/quantizer.py:15-29`` (``ds_quantizer``: fake-quantise in place),
the ``quantize / dequantize / swizzle_quant / quantized_reduction`` functions of the native module used by
ZeRO++ (``runtime/zero/partition_parameters.py:824-862`` ``CUDAQuantizer`` for qwZ,
``runtime/comm/coalesced_collectives.py:30-79`` for qgZ).  Kernels: ``csrc/quantize.cu``; the pure-PyTorch
branch below is the CPU oracle with identical rounding.
"""
from __future__ import annotations

from typing import Tuple

import torch

from ...native import lib as _lib

_lib.register({
    "sxb_quantize_int": "ppp" "l" "iiii" "i" "ii" "i" "p",
    "sxb_dequantize_int": "ppp" "l" "iii" "i" "p",
    "sxb_quantized_reduction": "pppp" "l" "iiii" "p",
})


def _check(t: torch.Tensor, groups: int) -> int:
    n = t.numel()
    if n % groups:
        raise ValueError(f"{n} elements cannot be split into {groups} groups")
    gs = n // groups
    if gs % 8 or gs > 8192:
        raise ValueError(f"group size {gs} must be a multiple of 8 and <= 8192")
    return gs


def _swizzle_perm(groups: int, nodes: int, devs: int, device) -> torch.Tensor:
    """output group index of every input group for the (node, dev, r) -> (dev, node, r) reorder."""
    g = torch.arange(groups, device=device)
    if nodes <= 1 or devs <= 1:
        return g
    per = groups // (nodes * devs)
    part, r = g // per, g % per
    node, dev = part // devs, part % devs
    return (dev * nodes + node) * per + r


def quantize(t: torch.Tensor, groups: int, bits: int = 8, symmetric: bool = True, stochastic: bool = False,
             seed: int = 0, nodes: int = 1, devs: int = 1) -> Tuple[torch.Tensor, torch.Tensor]:
    """Returns ``(q, scales)``: ``q`` int8 ``[groups, group_size * bits / 8]`` (int4 packs two values per byte,
    low nibble first), ``scales`` fp32 ``[groups]`` or ``[groups, 2]`` = (scale, min) when asymmetric."""
    assert bits in (4, 8)
    gs = _check(t, groups)
    t = t.contiguous()
    q = torch.empty(groups, gs * bits // 8, dtype=torch.int8, device=t.device)
    scales = torch.empty((groups,) if symmetric else (groups, 2), dtype=torch.float32, device=t.device)
    if t.is_cuda:
        _lib.call("sxb_quantize_int", t.data_ptr(), q.data_ptr(), scales.data_ptr(), groups, gs, bits, int(symmetric),
                  int(stochastic), seed & 0xFFFFFFFF, nodes, devs, _lib.dtype_code(t.dtype), _lib.stream_ptr())
        return q, scales
    x = t.reshape(groups, gs).float()
    if symmetric:
        qmax = (1 << (bits - 1)) - 1
        amax = x.abs().amax(dim=1, keepdim=True)
        sc = torch.where(amax > 0, amax / qmax, torch.ones_like(amax))
        y = x / sc
        r = torch.floor(y + torch.rand_like(y)) if stochastic else torch.round(y)
        r = r.clamp_(-qmax - 1, qmax)
        sc_out = sc.squeeze(1)
    else:
        levels = (1 << bits) - 1
        mn, mx = x.amin(dim=1, keepdim=True), x.amax(dim=1, keepdim=True)
        sc = torch.where(mx > mn, (mx - mn) / levels, torch.ones_like(mx))
        y = (x - mn) / sc
        r = torch.floor(y + torch.rand_like(y)) if stochastic else torch.round(y)
        r = r.clamp_(0, levels)
        sc_out = torch.cat([sc, mn], dim=1)
    ri = r.to(torch.int32)
    if bits == 8:
        packed = (ri & 0xFF).to(torch.uint8).view(torch.int8)
    else:
        ri = ri & 0xF
        packed = (ri[:, 0::2] | (ri[:, 1::2] << 4)).to(torch.uint8).view(torch.int8)
    perm = _swizzle_perm(groups, nodes, devs, t.device)
    q[perm] = packed
    scales[perm] = sc_out
    return q, scales


def dequantize(q: torch.Tensor, scales: torch.Tensor, groups: int, bits: int = 8, symmetric: bool = True,
               dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    assert bits in (4, 8)
    gs = q.numel() * 8 // bits // groups
    out = torch.empty(groups, gs, dtype=dtype, device=q.device)
    if q.is_cuda:
        _lib.call("sxb_dequantize_int", q.data_ptr(), scales.data_ptr(), out.data_ptr(), groups, gs, bits,
                  int(symmetric), _lib.dtype_code(dtype), _lib.stream_ptr())
        return out
    raw = q.reshape(groups, -1).view(torch.uint8).to(torch.int32)
    if bits == 8:
        vals = raw
    else:
        vals = torch.stack([raw & 0xF, raw >> 4], dim=2).reshape(groups, gs)
    if symmetric:
        sign_bit = 1 << (bits - 1)
        vals = torch.where(vals >= sign_bit, vals - (1 << bits), vals).float()
        out.copy_(vals * scales.reshape(groups, 1))
    else:
        out.copy_(vals.float() * scales[:, :1] + scales[:, 1:2])
    return out


def swizzle_quant(t: torch.Tensor, groups: int, bits: int, symmetric: bool, pipeline: int, nodes: int,
                  devices_per_node: int):
    """qgZ first step: quantise and lay the groups out (dev, node, r)-major (``pipeline`` is accepted for call
    compatibility; the kernel needs no software pipelining at this size)."""
    return quantize(t, groups, bits, symmetric, nodes=nodes, devs=devices_per_node)


def quantized_reduction(q: torch.Tensor, scales: torch.Tensor, in_groups: int, out_groups: int, bits: int,
                        symmetric: bool, devices_per_node: int):
    """qgZ middle step: ``q`` holds ``devices_per_node`` chunks of ``out_groups`` groups; returns the re-quantised
    sum over chunks."""
    assert in_groups == out_groups * devices_per_node
    gs = q.numel() * 8 // bits // in_groups
    qo = torch.empty(out_groups, gs * bits // 8, dtype=torch.int8, device=q.device)
    so = torch.empty((out_groups,) if symmetric else (out_groups, 2), dtype=torch.float32, device=q.device)
    if q.is_cuda:
        _lib.call("sxb_quantized_reduction", q.data_ptr(), scales.data_ptr(), qo.data_ptr(), so.data_ptr(), out_groups,
                  gs, bits, int(symmetric), devices_per_node, _lib.stream_ptr())
        return qo, so
    full = dequantize(q, scales, in_groups, bits, symmetric, torch.float32)
    summed = full.reshape(devices_per_node, out_groups, gs).sum(0)
    return quantize(summed, out_groups, bits, symmetric)


def ds_quantizer(input: torch.Tensor, groups: int = 1, bit_num: int = 8, sr: bool = False, asym: bool = False):
    """In-place fake quantisation (quantise then dequantise) as used by MoQ / compression."""
    q, s = quantize(input, groups, bit_num, symmetric=not asym, stochastic=sr)
    input.copy_(dequantize(q, s, groups, bit_num, not asym, input.dtype).view_as(input))
    return input


class Quantizer:
    """Object flavour with the reference's method names."""

    def __init__(self, q_bits: int = 8, symmetric: bool = True, group_size: int = 2048):
        self.q_bits, self.symmetric, self.group_size = q_bits, symmetric, group_size

    def _groups(self, t):
        gs = self.group_size
        while t.numel() % gs:
            gs //= 2
        return t.numel() // gs

    def quantize(self, t):
        return quantize(t, self._groups(t), self.q_bits, self.symmetric)

    def dequantize(self, q, scales, dtype=torch.bfloat16):
        return dequantize(q, scales, scales.shape[0], self.q_bits, self.symmetric, dtype)


class CUDAQuantizer:
    """int8 symmetric weight quantiser of ZeRO++ qwZ (reference partition_parameters.py:824-862): groups of
    at most ``group_size`` elements, chosen so that the group count divides the tensor."""
    group_size = 2048

    def quantize(self, param: torch.Tensor, groups=None):
        if groups is None:
            n = param.numel()
            gs = self.group_size
            while n % gs or gs % 8:
                gs -= 8
                if gs <= 0:
                    raise ValueError(f"cannot group {n} elements")
            groups = n // gs
        return quantize(param, groups, 8, True)

    def dequantize(self, quantized: torch.Tensor, scale: torch.Tensor, dtype=torch.bfloat16):
        return dequantize(quantized, scale, scale.shape[0], 8, True, dtype)
