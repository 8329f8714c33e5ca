"""MoQ — quantize-aware training schedule applied to the fp16/bf16 weights after each optimizer step.

Reference: ``deepspeed/runtime/quantize.py:14-180``: every 2-D parameter carries ``start_bits / target_bits / q_period``;
each time ``qsteps`` crosses the parameter's period its precision drops by one bit and the period doubles (times an
eigenvalue-derived factor), until ``target_bits``.  >=3 bits use group-wise symmetric/asymmetric fake quantisation
(stochastic or nearest rounding), 2 bits ternary, 1 bit binary; ``q_mixed_fp16`` blends the quantised and real weights
with a ratio that decays by ``q_change_ratio`` per step.  On CUDA the high-bit path runs the fused
quantise-dequantise kernel (``csrc/quantize.cu`` through ``ops.quantizer.ds_quantizer``)."""
from __future__ import annotations

import math

import torch

from ..utils.logging import logger

TWO_D_PARAMS = 6      # 2-D weight matrices per transformer layer (qkv, out, 2x mlp, ...): reference constant


class Quantizer:

    def __init__(self, q_groups=1, q_mixed_fp16=False, q_change_ratio=0.01, q_type=0, q_rounding=0, q_verbose=False,
                 q_eigenvalue=False, use_quantizer_kernel=False, layer_num=0):
        self.q_groups = q_groups
        self.q_mixed_fp16 = q_mixed_fp16
        self.q_change_ratio = q_change_ratio
        self.q_type = q_type
        self.qsteps = 0
        self.quantize_real_ratio = 1.000
        self.q_verbose = q_verbose
        self.q_eigenvalue = q_eigenvalue
        self.use_quantizer_kernel = use_quantizer_kernel
        self.q_rounding = q_rounding
        self.layer_num = layer_num

    # the engine sets these per-layer lists when eigenvalue-driven schedules are on
    q_start_bits: list = []
    q_period: list = []
    q_target_bits: int = 8

    def any_precision_switch(self) -> bool:
        if self.layer_num == 0:
            return True
        nxt = self.qsteps + TWO_D_PARAMS * max(self.layer_num, 1)
        return any(self.q_start_bits[i] != self.q_target_bits and nxt >= self.q_period[i]
                   for i in range(self.layer_num))

    def quantize(self, parameter_group, overflow, eigenvalue_enabled, block_eigenvalue={}):
        if overflow and not eigenvalue_enabled:
            return
        self.step()
        self.update_fp16_ratio()
        for group in parameter_group:
            for p in group:
                if p.dim() > 1 and getattr(p, "start_bits", None):
                    eig, layer_id = (None, 0)
                    if block_eigenvalue:
                        eig, layer_id = block_eigenvalue.get(id(p), (None, 0))
                    factor = 1 + math.floor(eig * 4) if eig is not None else 1
                    p.data = self.compute_quantization(p, layer_id, factor)

    def step(self):
        self.qsteps += 1

    # ------------------------------------------------------------------------------------------ quantisers
    def _groups(self, x):
        return x.reshape(self.q_groups, -1)

    def quantize_highbit(self, inputs, num_bits):
        q_range = 2 ** num_bits
        flat = self._groups(inputs.float())
        g_min, g_max = flat.amin(dim=-1, keepdim=True), flat.amax(dim=-1, keepdim=True)
        noise = 0.0 if self.q_rounding == "nearest" else torch.empty_like(flat).uniform_(-0.5, 0.5)
        if self.q_type == "symmetric":
            scale = (2 * torch.maximum(g_min.abs(), g_max.abs()) / q_range).clamp_min(1e-12)
            out = (flat / scale + noise).round().clamp(-(q_range >> 1), (q_range >> 1) - 1) * scale
        elif self.q_type == "asymmetric":
            scale = ((g_max - g_min) / q_range).clamp_min(1e-12)
            zero = (g_min / scale).round() * scale
            out = ((flat - zero) / scale + noise).round().clamp(0, q_range - 1) * scale + zero
        else:
            raise ValueError(f"unknown quantization type {self.q_type!r}")
        return out.reshape(inputs.shape).to(inputs.dtype).contiguous()

    def quantize_tenary(self, inputs):
        flat = self._groups(inputs.float())
        thres = (0.7 * flat.abs().mean(dim=1)).view(-1, 1)
        mask = flat.abs() > thres
        alpha = ((flat.abs() * mask).sum(dim=1) / mask.sum(dim=1).clamp_min(1)).view(-1, 1)
        out = alpha * (flat > thres) - alpha * (flat < -thres)
        return out.reshape(inputs.shape).to(inputs.dtype).contiguous()

    def quantize_binary(self, inputs):
        flat = self._groups(inputs.float())
        out = flat.sign() * flat.abs().mean(dim=1, keepdim=True)
        return out.reshape(inputs.shape).to(inputs.dtype).contiguous()

    def mixed_fp16_quantize(self, input, input_q, index):
        if self.q_mixed_fp16 and self.q_start_bits[index] >= (self.q_target_bits - 1):
            return input * self.quantize_real_ratio + (1 - self.quantize_real_ratio) * input_q
        return input_q

    def compute_quantization(self, input, index=0, factor=1):
        if input.start_bits != input.target_bits and self.qsteps >= input.q_period:
            self.quantize_real_ratio = 1.0
            input.q_period <<= 1
            input.q_period *= factor
            input.start_bits -= 1
            if self.q_verbose:
                logger.info(f"Quantization settings: current bit-precision = {input.start_bits}, step = "
                            f"{self.qsteps}, quantization period = {input.q_period}, index = {index}")
        assert input.start_bits >= input.target_bits, "Quantization bit is lower than target precision bits!"
        bits = input.start_bits
        if self.use_quantizer_kernel:
            if bits <= 2:
                raise ValueError("Quantization bit is too low, please do it without quantization kernel!")
            from ..ops.quantizer import ds_quantizer
            q = ds_quantizer(input.data.clone(), self.q_groups, bits, asym=self.q_type != "symmetric",
                             sr=self.q_rounding != "nearest_neighbor")
            return self.mixed_fp16_quantize(input.data, q, index)
        if bits >= 3:
            q = self.quantize_highbit(input.data, bits)
        else:
            assert self.q_type == "symmetric", "Quantization type is not symmetric!"
            assert self.q_rounding == "nearest", "Quantization rounding is not nearest_neighbor!"
            q = self.quantize_tenary(input.data) if bits == 2 else self.quantize_binary(input.data)
        if self.q_mixed_fp16 and bits >= input.target_bits - 1:
            q = self.quantize_real_ratio * input.data + (1 - self.quantize_real_ratio) * q
        return q

    def update_fp16_ratio(self):
        if self.q_mixed_fp16:
            self.quantize_real_ratio = max(0.0, self.quantize_real_ratio - self.q_change_ratio) \
                if self.quantize_real_ratio > 0 else 0.0
