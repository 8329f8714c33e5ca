"""Group-scaled miniature-float quantiser (FP8 e4m3 / FP6 e3m2 / FP4 e2m1 / FP12 e7m4).

Reference: ``deepspeed/ops/fp_quantizer/quantize.py:16-166`` (``FP_Quantize.quantize / dequantize /
selective_dequantize / get_scales``; payload ``[num_groups, group_size*q_bits/8 + 4]`` uint8 with the fp32 scale in
the last four bytes of every group).  Kernels in ``csrc/quantize.cu``; the PyTorch branch is the CPU oracle.
"""
from __future__ import annotations

import abc

import torch

from ...native import lib as _lib

_lib.register({
    "sxb_fp_quantize": "pp" "l" "iii" "i" "i" "p",
    "sxb_fp_gemv": "ppp" "iiiii" "ll" "p",
    "sxb_fp_dequantize": "pp" "l" "ii" "p" "l" "i" "p",
})

_FORMATS = {8: (4, 3), 6: (3, 2), 4: (2, 1), 12: (7, 4)}   # q_bits -> (exponent bits, mantissa bits)


def _fmt_max(e: int, m: int) -> float:
    bias = (1 << (e - 1)) - 1
    return (2.0 - 2.0 ** (-m)) * 2.0 ** ((1 << e) - 1 - bias)


def _encode_cpu(x: torch.Tensor, e: int, m: int, stochastic: bool) -> torch.Tensor:
    """x already divided by the group scale; returns integer codes (int64)."""
    bias = (1 << (e - 1)) - 1
    sign = (x < 0) | ((x == 0) & (torch.signbit(x)))
    a = x.abs().clamp(max=_fmt_max(e, m)).double()
    mant, ex = torch.frexp(a)                      # a = mant * 2^ex, mant in [0.5, 1)
    ef = ex - 1 + bias
    sub = ef <= 0
    quantum = torch.where(sub, torch.full_like(a, 2.0 ** (1 - bias - m)), torch.pow(2.0, (ex - 1 - m).double()))
    q = a / quantum
    r = torch.floor(q + torch.rand_like(q)) if stochastic else torch.round(q)   # torch.round = half-to-even
    r = r.to(torch.int64)
    ef = torch.where(sub, torch.zeros_like(ef), ef).to(torch.int64)
    carry = (~sub) & (r >= (2 << m))
    r = torch.where(carry, r >> 1, r)
    ef = torch.where(carry, ef + 1, ef)
    over = ef > ((1 << e) - 1)
    ef = torch.where(over, torch.full_like(ef, (1 << e) - 1), ef)
    r = torch.where(over, torch.full_like(r, (2 << m) - 1), r)
    code = torch.where(sub, r, (ef << m) | (r & ((1 << m) - 1)))
    code = torch.where(a == 0, torch.zeros_like(code), code)
    return code | (sign.to(torch.int64) << (e + m))


def _decode_cpu(code: torch.Tensor, e: int, m: int) -> torch.Tensor:
    bias = (1 << (e - 1)) - 1
    sign = (code >> (e + m)) & 1
    ef = (code >> m) & ((1 << e) - 1)
    mant = code & ((1 << m) - 1)
    sub = ef == 0
    v = torch.where(sub, mant.double() * 2.0 ** (1 - bias - m),
                    (1.0 + mant.double() * 2.0 ** (-m)) * torch.pow(2.0, (ef - bias).double()))
    return torch.where(sign == 1, -v, v).float()


def _pack_cpu(code: torch.Tensor, bits: int) -> torch.Tensor:
    """[groups, gs] codes -> [groups, gs*bits/8] bytes; 8 codes form a little-endian bit stream of `bits` bytes."""
    g, n = code.shape
    c = code.reshape(g, n // 8, 8)
    # byte-wise assembly (exact, avoids int64 sign issues)
    out = torch.zeros(g, n // 8, bits, dtype=torch.uint8)
    for b in range(bits):
        byte = torch.zeros(g, n // 8, dtype=torch.int64)
        for i in range(8):
            lo, hi_ = i * bits, (i + 1) * bits
            s, e_ = max(lo, 8 * b), min(hi_, 8 * b + 8)
            if s < e_:
                byte |= ((c[:, :, i] >> (s - lo)) & ((1 << (e_ - s)) - 1)) << (s - 8 * b)
        out[:, :, b] = byte.to(torch.uint8)
    return out.reshape(g, n * bits // 8)


def _unpack_cpu(raw: torch.Tensor, bits: int, gs: int) -> torch.Tensor:
    g = raw.shape[0]
    r = raw.reshape(g, gs // 8, bits).to(torch.int64)
    code = torch.zeros(g, gs // 8, 8, dtype=torch.int64)
    for i in range(8):
        lo, hi_ = i * bits, (i + 1) * bits
        for b in range(bits):
            s, e_ = max(lo, 8 * b), min(hi_, 8 * b + 8)
            if s < e_:
                code[:, :, i] |= ((r[:, :, b] >> (s - 8 * b)) & ((1 << (e_ - s)) - 1)) << (s - lo)
    return code.reshape(g, gs)


GEMV_MAX_ROWS = 4


def fp_gemv_ok(x: torch.Tensor, out_features: int, in_features: int, group_size: int, q_bits: int) -> bool:
    """Whether :func:`fp_gemv` takes this problem (decode-sized batch, bf16 CUDA activations, aligned K)."""
    rows = x.numel() // max(1, x.shape[-1])
    return (x.is_cuda and x.dtype == torch.bfloat16 and rows <= GEMV_MAX_ROWS and q_bits in (4, 6, 8)
            and in_features % 512 == 0 and in_features % group_size == 0 and group_size % 16 == 0
            and x.shape[-1] == in_features)


def fp_gemv(x: torch.Tensor, wq: torch.Tensor, out_features: int, in_features: int, q_bits: int = 6,
            group_size: int = 512) -> torch.Tensor:
    """``x @ W^T`` for a weight kept in :func:`fp_quantize`'s packed layout (``W`` is ``[out, in]``, quantised flat):
    the packed codes are read once and decoded in registers (``csrc/quantize.cu: fp_gemv_kernel``).  For decode-sized
    ``x`` (at most :data:`GEMV_MAX_ROWS` rows) this is bound by the packed weight bytes — 0.75 B per weight at 6 bits
    instead of the 4.75 B of dequantise-to-scratch + GEMM."""
    x2 = x.reshape(-1, in_features)
    if not x2.is_contiguous():
        x2 = x2.contiguous()
    y = torch.empty(x2.shape[0], out_features, dtype=torch.bfloat16, device=x.device)
    _lib.call("sxb_fp_gemv", x2.data_ptr(), wq.data_ptr(), y.data_ptr(), x2.shape[0], out_features, in_features,
              int(q_bits), int(group_size), x2.stride(0), y.stride(0), _lib.stream_ptr())
    return y.view(*x.shape[:-1], out_features)


def fp_quantize(x: torch.Tensor, group_size: int = 512, q_bits: int = 8, stochastic: bool = False, seed: int = 0):
    e, m = _FORMATS[q_bits]
    n = x.numel()
    assert n % group_size == 0 and group_size % 8 == 0, "tensor size must be a multiple of the group size"
    groups = n // group_size
    out = torch.empty(groups, group_size * q_bits // 8 + 4, dtype=torch.uint8, device=x.device)
    x = x.contiguous()
    if x.is_cuda:
        _lib.call("sxb_fp_quantize", x.data_ptr(), out.data_ptr(), groups, group_size, q_bits, int(stochastic),
                  seed & 0xFFFFFFFF, _lib.dtype_code(x.dtype), _lib.stream_ptr())
        return out
    v = x.reshape(groups, group_size).float()
    amax = v.abs().amax(dim=1, keepdim=True)
    scale = torch.where(amax > 0, amax / _fmt_max(e, m), torch.ones_like(amax))
    inv = 1.0 / scale                       # the kernel multiplies by the reciprocal
    code = _encode_cpu(v * inv, e, m, stochastic)
    out[:, :-4] = _pack_cpu(code, q_bits)
    out[:, -4:] = scale.contiguous().view(torch.uint8).reshape(groups, 4)
    return out


def fp_dequantize(q: torch.Tensor, group_size: int = 512, q_bits: int = 8, dtype=torch.bfloat16, rows=None,
                  groups_per_row: int = 1, out: torch.Tensor = None):
    e, m = _FORMATS[q_bits]
    gbytes = group_size * q_bits // 8 + 4
    q = q.reshape(-1, gbytes)
    if rows is not None:
        rows = rows.to(torch.int32).contiguous()
        out_groups = rows.numel() * groups_per_row
    else:
        out_groups = q.shape[0]
    if out is None:
        out = torch.empty(out_groups, group_size, dtype=dtype, device=q.device)
    if q.is_cuda:
        _lib.call("sxb_fp_dequantize", q.data_ptr(), out.data_ptr(), out_groups, group_size, q_bits,
                  None if rows is None else rows.data_ptr(), groups_per_row, _lib.dtype_code(out.dtype),
                  _lib.stream_ptr())
        return out
    if rows is not None:
        sel = (rows.long()[:, None] * groups_per_row + torch.arange(groups_per_row)[None, :]).reshape(-1)
        q = q[sel]
    scale = q[:, -4:].contiguous().view(torch.float32)
    code = _unpack_cpu(q[:, :-4], q_bits, group_size)
    out.view(out_groups, group_size).copy_(_decode_cpu(code, e, m) * scale)
    return out


class Quantizer(abc.ABC):
    def __init__(self, group_size=512) -> None:
        self.group_size = group_size

    @abc.abstractmethod
    def quantize(self, input, q_bits=8, q_mantisa_bits=3, stochastic_mode=False, return_meta_tensor=False): ...

    @abc.abstractmethod
    def dequantize(self, input_q, fp_out=None, q_bits=8, q_mantisa_bits=3, scale=None): ...


class FP_Quantize(Quantizer):
    def __init__(self, quantization_config=None, group_size: int = None) -> None:
        gs = group_size if group_size is not None else getattr(quantization_config, "group_size", 512)
        super().__init__(group_size=gs)
        self.q_config = quantization_config
        self.orig_dtype = self.orig_shape = self.num_groups = self.input_q = self.scale = None

    def quantize(self, input, q_bits=8, q_mantisa_bits=3, stochastic_mode=False, return_meta_tensor=False):
        assert input.dtype in (torch.bfloat16, torch.float16, torch.float32)
        if return_meta_tensor:
            assert q_bits == 8, "meta tensor is only supported with q_bit=8"
        if q_bits not in _FORMATS:
            raise AssertionError(f"Missing {q_bits}-quantization")
        self.orig_dtype, self.orig_shape = input.dtype, input.shape
        self.num_groups = input.numel() // self.group_size
        out = fp_quantize(input, self.group_size, q_bits, stochastic_mode)
        self.input_q = out
        if return_meta_tensor:
            data, self.scale = out[:, :-4], out[:, -4:]
            return data.contiguous().reshape(input.shape), self.scale
        return out

    def to(self, *args, **kwargs):
        for name in ("input_q", "scale"):
            t = getattr(self, name)
            if t is not None:
                setattr(self, name, t.to(*args, **kwargs))

    def get_scales(self):
        return self.input_q[:, -4:].contiguous().view(torch.float32)

    def _assemble(self, input_q, scale, q_bits):
        if scale is None:
            return input_q
        payload = input_q.reshape(-1, self.group_size * q_bits // 8)
        return torch.cat([payload, scale.reshape(-1, 4).to(payload.device)], dim=1).contiguous()

    def dequantize(self, input_q, fp_out=None, q_bits=8, q_mantisa_bits=3, scale=None):
        if q_bits not in _FORMATS:
            raise AssertionError(f"Missing {q_bits}-dequantization")
        q = self._assemble(input_q, scale, q_bits)
        dtype = self.orig_dtype or torch.bfloat16
        out = fp_dequantize(q, self.group_size, q_bits, dtype)
        out = out.reshape(self.orig_shape) if self.orig_shape is not None and out.numel() == torch.Size(self.orig_shape).numel() else out
        if fp_out is not None:
            fp_out.copy_(out.view_as(fp_out))
            return fp_out
        return out

    def selective_dequantize(self, input_q, indexes, fp_out=None, q_bits=8, q_mantisa_bits=3, scale=None):
        assert self.orig_shape is not None and len(self.orig_shape) == 3, "Selective-Dequantization works on 3d tensor only!"
        q = self._assemble(input_q, scale, q_bits)
        per_row = (self.orig_shape[1] * self.orig_shape[2]) // self.group_size
        out = fp_dequantize(q, self.group_size, q_bits, self.orig_dtype, rows=indexes, groups_per_row=per_row)
        out = out.reshape(indexes.numel(), self.orig_shape[1], self.orig_shape[2])
        if fp_out is not None:
            fp_out.copy_(out)
            return fp_out
        return out
