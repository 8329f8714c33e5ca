"""Fused bandwidth-bound building blocks shared by the training transformer layer, Evoformer attention and the
block-sparse attention (kernels: ``csrc/softmax_dropout.cu``).

* :func:`bias_dropout_residual` — ``residual + dropout(x + bias)`` in one pass, 1-byte keep mask for the backward
  (the fused epilogues of the reference's native BERT layer, ``ops/transformer/transformer.py:149-293``).
* :func:`biased_softmax` — ``softmax(scale * s + bias1 + bias2)`` over ``[B, H, Sq, Sk]`` with broadcastable biases
  and bias gradients (Evoformer mask + pair bias, ``ops/deepspeed4science/evoformer_attn.py:29-106``; BERT additive
  mask).
* :func:`bsr_softmax` — softmax over the non-zero blocks of a block-sparse score tensor with optional relative
  position bias, key-padding mask and attention mask (``ops/sparse_attention/softmax.py:123-296``).

Each op has a PyTorch branch with identical semantics for CPU tensors (also the oracle in the tests).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from ..native import lib as _lib

_lib.register({
    "sxb_bias_dropout_residual": "ppppp" "l" "i" "f" "ll" "i" "p",
    "sxb_dropout_bwd": "ppp" "l" "f" "i" "p",
    "sxb_biased_softmax_fwd": "pp" "pllll" "pllll" "l" "iii" "f" "i" "p",
    "sxb_biased_softmax_bwd": "pppp" "l" "i" "f" "i" "p",
    "sxb_bsr_softmax_fwd": "pppp" "l" "iiiii" "pl" "pi" "pi" "f" "i" "p",
    "sxb_bsr_softmax_bwd": "ppppp" "l" "iiiii" "f" "i" "p",
})

_FLOAT = (torch.float32, torch.float16, torch.bfloat16)


def _philox_args(device) -> Tuple[int, int]:
    """(seed, offset) drawn from torch's generator so ``torch.manual_seed`` makes dropout reproducible."""
    gen = torch.cuda.default_generators[device.index if device.index is not None else torch.cuda.current_device()]
    seed = int(gen.initial_seed()) & 0x7FFFFFFFFFFFFFFF
    off = int(torch.randint(0, 2 ** 62, (1,), device="cpu").item())
    return seed, off


# ------------------------------------------------------------------------------------- bias + dropout + residual
class _BiasDropoutResidual(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bias, residual, p, training):
        p_eff = float(p) if training else 0.0
        x = x.contiguous()
        out = torch.empty_like(x)
        keep = torch.empty(x.shape, dtype=torch.uint8, device=x.device) if p_eff > 0 else None
        seed, off = _philox_args(x.device) if p_eff > 0 else (0, 0)
        res = residual.contiguous() if residual is not None else None
        _lib.call("sxb_bias_dropout_residual", x.data_ptr(), _lib.ptr(bias), _lib.ptr(res), out.data_ptr(),
                  _lib.ptr(keep), x.numel(), x.shape[-1], p_eff, seed, off, _lib.dtype_code(x.dtype),
                  _lib.stream_ptr())
        ctx.p = p_eff
        ctx.has_bias = bias is not None
        ctx.has_res = residual is not None
        ctx.save_for_backward(keep) if keep is not None else ctx.save_for_backward()
        return out

    @staticmethod
    def backward(ctx, dy):
        dy = dy.contiguous()
        if ctx.p > 0:
            (keep,) = ctx.saved_tensors
            dx = torch.empty_like(dy)
            _lib.call("sxb_dropout_bwd", dy.data_ptr(), keep.data_ptr(), dx.data_ptr(), dy.numel(), ctx.p,
                      _lib.dtype_code(dy.dtype), _lib.stream_ptr())
        else:
            dx = dy
        dbias = dx.reshape(-1, dx.shape[-1]).sum(0) if ctx.has_bias else None
        return dx, dbias, (dy if ctx.has_res else None), None, None


def bias_dropout_residual(x: torch.Tensor, bias: Optional[torch.Tensor] = None,
                          residual: Optional[torch.Tensor] = None, p: float = 0.0,
                          training: bool = True) -> torch.Tensor:
    """``residual + dropout(x + bias, p)``; ``bias`` broadcasts over the last dim."""
    if x.is_cuda and x.dtype in _FLOAT and x.shape[-1] % 4 == 0:
        if bias is not None:
            bias = bias.to(x.dtype).contiguous()
        if residual is not None:
            residual = residual.to(x.dtype)
        return _BiasDropoutResidual.apply(x, bias, residual, p, training)
    y = x if bias is None else x + bias
    y = torch.nn.functional.dropout(y, p, training)
    return y if residual is None else y + residual


def dropout(x: torch.Tensor, p: float, training: bool = True) -> torch.Tensor:
    if p <= 0 or not training:
        return x
    return bias_dropout_residual(x, None, None, p, training)


# ------------------------------------------------------------------------------------- biased softmax
def _bias_strides(bias: Optional[torch.Tensor], shape) -> Tuple[Optional[torch.Tensor], Tuple[int, int, int, int]]:
    if bias is None:
        return None, (0, 0, 0, 0)
    while bias.dim() < 4:
        bias = bias.unsqueeze(0)
    st = []
    for d in range(4):
        if bias.shape[d] == 1:
            st.append(0)
        elif bias.shape[d] == shape[d]:
            st.append(bias.stride(d))
        else:
            raise ValueError(f"bias shape {tuple(bias.shape)} does not broadcast to scores {tuple(shape)}")
    return bias, tuple(st)


def _reduce_to(grad: torch.Tensor, bias_shape) -> torch.Tensor:
    """Sum a full-shape gradient down to a broadcast operand's shape."""
    shape = list(bias_shape)
    lead = grad.dim() - len(shape)
    if lead:
        grad = grad.sum(dim=tuple(range(lead)))
    dims = tuple(i for i, s in enumerate(shape) if s == 1 and grad.shape[i] != 1)
    if dims:
        grad = grad.sum(dim=dims, keepdim=True)
    return grad


class _BiasedSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scores, bias1, bias2, scale):
        B, H, Sq, Sk = scores.shape
        scores = scores.contiguous()
        out = torch.empty_like(scores)
        b1, s1 = _bias_strides(bias1.to(scores.dtype) if bias1 is not None else None, scores.shape)
        b2, s2 = _bias_strides(bias2.to(scores.dtype) if bias2 is not None else None, scores.shape)
        _lib.call("sxb_biased_softmax_fwd", scores.data_ptr(), out.data_ptr(), _lib.ptr(b1), *s1, _lib.ptr(b2), *s2,
                  B, H, Sq, Sk, float(scale), _lib.dtype_code(scores.dtype), _lib.stream_ptr())
        ctx.scale = float(scale)
        ctx.b1_shape = tuple(bias1.shape) if bias1 is not None else None
        ctx.b2_shape = tuple(bias2.shape) if bias2 is not None else None
        ctx.need_b = (bias1 is not None and bias1.requires_grad, bias2 is not None and bias2.requires_grad)
        ctx.save_for_backward(out)
        return out

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        dy = dy.contiguous()
        ds = torch.empty_like(y)
        need_b = any(ctx.need_b)
        db = torch.empty_like(y) if need_b else None
        rows = y.numel() // y.shape[-1]
        _lib.call("sxb_biased_softmax_bwd", y.data_ptr(), dy.data_ptr(), ds.data_ptr(), _lib.ptr(db), rows,
                  y.shape[-1], ctx.scale, _lib.dtype_code(y.dtype), _lib.stream_ptr())
        g1 = _reduce_to(db.float(), ctx.b1_shape).to(y.dtype) if ctx.need_b[0] else None
        g2 = _reduce_to(db.float(), ctx.b2_shape).to(y.dtype) if ctx.need_b[1] else None
        return ds, g1, g2, None


def biased_softmax(scores: torch.Tensor, bias1: Optional[torch.Tensor] = None,
                   bias2: Optional[torch.Tensor] = None, scale: float = 1.0) -> torch.Tensor:
    """``softmax(scale * scores + bias1 + bias2, dim=-1)`` for ``scores [B, H, Sq, Sk]``; biases broadcast."""
    if scores.is_cuda and scores.dtype in _FLOAT and scores.dim() == 4:
        return _BiasedSoftmax.apply(scores, bias1, bias2, scale)
    s = scores.float() * scale
    if bias1 is not None:
        s = s + bias1.float()
    if bias2 is not None:
        s = s + bias2.float()
    return torch.softmax(s, dim=-1).to(scores.dtype)


# ------------------------------------------------------------------------------------- block-sparse softmax
class BsrLayout:
    """CSR view of a ``[H, nrb, ncb]`` 0/1 block layout; block order = ``layout.nonzero()`` order (head, row, col)."""

    def __init__(self, layout: torch.Tensor, block: int):
        layout = layout.to(torch.int64).cpu()
        self.layout = layout
        self.block = int(block)
        self.H, self.nrb, self.ncb = layout.shape
        nz = layout.nonzero()
        self.nnz = int(nz.shape[0])
        self.head_idx = nz[:, 0].contiguous()
        self.row_idx = nz[:, 1].contiguous()
        self.col_idx = nz[:, 2].contiguous()
        counts = layout.reshape(self.H * self.nrb, self.ncb).sum(-1)
        self.row_ptr = torch.zeros(self.H * self.nrb + 1, dtype=torch.int32)
        self.row_ptr[1:] = torch.cumsum(counts, 0).to(torch.int32)
        self._dev = {}

    def on(self, device):
        key = str(device)
        if key not in self._dev:
            self._dev[key] = (self.row_ptr.to(device), self.col_idx.to(torch.int32).to(device),
                              self.head_idx.to(device), self.row_idx.to(device), self.col_idx.to(device))
        return self._dev[key]


def _bsr_softmax_ref(x, lay: BsrLayout, scale, rpe, kpm, attn_mask, kpm_mode, mask_mode):
    """Dense-expand oracle: scatter the blocks into [B,H,S,S], mask the zero blocks, softmax, gather back."""
    B = x.shape[0]
    blk, H = lay.block, lay.H
    S, Sk = lay.nrb * blk, lay.ncb * blk
    _, _, hi, ri, ci = lay.on(x.device)
    dense = torch.full((B, H, lay.nrb, lay.ncb, blk, blk), float("-inf"), dtype=torch.float32, device=x.device)
    dense[:, hi, ri, ci] = x.float() * scale
    dense = dense.permute(0, 1, 2, 4, 3, 5).reshape(B, H, S, Sk)
    if rpe is not None:
        dense = dense + rpe.float()
    if kpm is not None:
        m = kpm.float()[:, None, None, :]
        dense = dense.masked_fill(m == 0, float("-inf")) if kpm_mode == "mul" else dense + m
    if attn_mask is not None:
        m = attn_mask.float()[None, None]
        dense = dense.masked_fill(m == 0, float("-inf")) if mask_mode == "mul" else dense + m
    prob = torch.softmax(dense, dim=-1)
    prob = torch.nan_to_num(prob, nan=0.0)
    prob = prob.reshape(B, H, lay.nrb, blk, lay.ncb, blk).permute(0, 1, 2, 4, 3, 5)
    return prob[:, hi, ri, ci].to(x.dtype)


class _BsrSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, lay, scale, rpe, kpm, attn_mask, kpm_mode, mask_mode):
        x = x.contiguous()
        y = torch.empty_like(x)
        row_ptr, col_idx = lay.on(x.device)[:2]
        B = x.shape[0]
        rpe_sb = 0
        if rpe is not None:
            rpe = rpe.to(x.dtype).contiguous()
            if rpe.dim() == 4:
                rpe_sb = rpe.stride(0) if rpe.shape[0] > 1 else 0
        kpm_c = kpm.to(x.dtype).contiguous() if kpm is not None else None
        am_c = attn_mask.to(x.dtype).contiguous() if attn_mask is not None else None
        _lib.call("sxb_bsr_softmax_fwd", x.data_ptr(), y.data_ptr(), row_ptr.data_ptr(), col_idx.data_ptr(), B, lay.H,
                  lay.nrb, lay.ncb, lay.block, lay.nnz, _lib.ptr(rpe), rpe_sb, _lib.ptr(kpm_c), 2 if kpm_mode == "mul" else 1,
                  _lib.ptr(am_c), 2 if mask_mode == "mul" else 1, float(scale), _lib.dtype_code(x.dtype),
                  _lib.stream_ptr())
        ctx.lay, ctx.scale = lay, float(scale)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        lay = ctx.lay
        dy = dy.contiguous()
        dx = torch.empty_like(y)
        row_ptr, col_idx = lay.on(y.device)[:2]
        _lib.call("sxb_bsr_softmax_bwd", y.data_ptr(), dy.data_ptr(), dx.data_ptr(), row_ptr.data_ptr(),
                  col_idx.data_ptr(), y.shape[0], lay.H, lay.nrb, lay.ncb, lay.block, lay.nnz, ctx.scale,
                  _lib.dtype_code(y.dtype), _lib.stream_ptr())
        return dx, None, None, None, None, None, None, None


def bsr_softmax(x: torch.Tensor, lay: BsrLayout, scale: float = 1.0, rpe: Optional[torch.Tensor] = None,
                key_padding_mask: Optional[torch.Tensor] = None, attn_mask: Optional[torch.Tensor] = None,
                key_padding_mask_mode: str = "add", attn_mask_mode: str = "add") -> torch.Tensor:
    """``x`` is ``[B, nnz, block, block]`` (blocks in ``layout.nonzero()`` order); softmax runs over every score of
    one query row across the non-zero blocks of its block-row."""
    if x.is_cuda and x.dtype in _FLOAT:
        return _BsrSoftmax.apply(x, lay, scale, rpe, key_padding_mask, attn_mask, key_padding_mask_mode,
                                 attn_mask_mode)
    return _bsr_softmax_ref(x, lay, scale, rpe, key_padding_mask, attn_mask, key_padding_mask_mode, attn_mask_mode)
