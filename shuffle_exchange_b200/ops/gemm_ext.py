"""Gated, grouped and fp8 GEMMs on the extended CTA-pair tcgen05 kernel (``csrc/gemm_ext.cu``).

* :class:`Gate` + ``gemm(..., gate=)`` — K2: the weight operand is read from a buffer a ZeRO-3 gather is still
  filling; the kernel's TMA producer waits on the gather's arrival flags tile by tile
  (reference: ``all_gather_into_tensor`` then ``F.linear``, runtime/zero/partition_parameters.py:1221-1261).
* :func:`grouped_linear` — K4/K5: every local expert's forward / dgrad / wgrad GEMM in ONE launch over device-side
  row offsets, no host synchronisation (reference: python loop over experts, moe/experts.py:28-38).
* :func:`fp8_rowwise_quant`, :func:`fp8_gemm_nt`, ``grouped_linear(..., fp8=True)`` — K14: e4m3 operands on
  ``tcgen05.mma kind::f8f6f4`` with per-row / per-output-channel scales applied in the epilogue
  (reference: ops/fp_quantizer/fp8_gemm_triton.py:17-165 dequantises before the dot).

CPU tensors take a plain PyTorch branch (test oracle).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from ..native import lib as _n

GEMM_NT, GEMM_NN, GEMM_TN = 0, 1, 2
PLAIN, RAGGED_M, RAGGED_K = 0, 1, 2
GROUP_ALIGN_M = 256          # row alignment of every group in ragged-M problems (one CTA-pair tile)
GROUP_ALIGN_K = 64

_n.register({
    "sxb_gemm_ext": "ppp" "iiii" "lll" "iii" "f" "pilll" "iipll" "pp" "i" "p",
    "sxb_gemmx_hang_info": "p",
    "sxb_stream_write_u32": "pip",
    "sxb_fp8_rowwise_quant": "ppp" "li" "ll" "p",
})


@dataclass
class Gate:
    """Arrival flags of the buffer a gated operand lives in: ``flags[c] >= value`` (as wrapping 32-bit sequence
    numbers) <=> bytes ``[c*chunk_bytes, (c+1)*chunk_bytes)`` of the buffer have landed."""
    flags: torch.Tensor          # int32 / uint32, device
    value: int
    chunk_bytes: int
    buffer_ptr: int              # data_ptr() of the first byte the flags cover

    def base_of(self, t: torch.Tensor) -> int:
        off = t.data_ptr() - self.buffer_ptr
        if off < 0 or off >= self.flags.numel() * self.chunk_bytes:
            raise _n.NativeError("gated operand does not live in the flagged buffer")
        return off


def _call(a, b, c, M, N, K, layout, lda, ldb, ldc, ab_dtype, c_dtype, accumulate, alpha, gate: Optional[Gate],
          b_tensor, group_mode, num_groups, offsets, total_rows, c_group_stride, scale_a, scale_b, max_ctas=0):
    gf, gv, gb, gc, gn = None, 0, 0, 0, 0
    if gate is not None:
        gf, gv, gb, gc, gn = gate.flags.data_ptr(), int(gate.value) & 0xFFFFFFFF, gate.base_of(b_tensor), \
            int(gate.chunk_bytes), gate.flags.numel()
    _n.call("sxb_gemm_ext", a, b, c, int(M), int(N), int(K), int(layout), int(lda), int(ldb), int(ldc),
            int(ab_dtype), int(c_dtype), int(bool(accumulate)), float(alpha), gf, gv, int(gb), int(gc), int(gn),
            int(group_mode), int(num_groups), _n.ptr(offsets), int(total_rows), int(c_group_stride),
            _n.ptr(scale_a), _n.ptr(scale_b), int(max_ctas), _n.stream_ptr())


def gemm_gated(a: torch.Tensor, b: torch.Tensor, layout: int, gate: Optional[Gate], out: Optional[torch.Tensor] = None,
               out_dtype=None, accumulate: bool = False, alpha: float = 1.0) -> torch.Tensor:
    """Same contract as ``ops.functional.gemm`` (bf16 operands) with ``b`` possibly still arriving."""
    if layout == GEMM_NT:
        M, K = a.shape
        N = b.shape[0]
    elif layout == GEMM_NN:
        M, K = a.shape
        N = b.shape[1]
    else:
        K, M = a.shape
        N = b.shape[1]
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype or a.dtype, device=a.device)
        accumulate = False
    if not a.is_cuda:
        fa, fb = a.float(), b.float()
        r = alpha * (fa @ fb.t() if layout == GEMM_NT else (fa @ fb if layout == GEMM_NN else fa.t() @ fb))
        out.copy_((r + out.float() if accumulate else r).to(out.dtype))
        return out
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16:
        raise _n.NativeError("gated gemm expects bf16 operands")
    _call(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, layout, a.stride(0), b.stride(0), out.stride(0),
          _n.BF16, _n.dtype_code(out.dtype), accumulate, alpha, gate, b, PLAIN, 0, None, 0, 0, None, None)
    return out


def stream_write_u32(flags: torch.Tensor, index: int, value: int) -> None:
    """Stream-ordered write of ``flags[index] = value`` by the stream front end (no kernel, no SM)."""
    _n.call("sxb_stream_write_u32", flags.data_ptr() + 4 * int(index), int(value) & 0xFFFFFFFF, _n.stream_ptr())


# --------------------------------------------------------------------------------------- fp8
def fp8_rowwise_quant(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """bf16 ``[rows, cols]`` -> (e4m3 bytes ``[rows, cols]`` as uint8, fp32 scale per row); value = q * scale."""
    rows, cols = x.shape
    if not x.is_cuda:
        amax = x.float().abs().amax(dim=1)
        scale = torch.where(amax > 0, amax / 448.0, torch.ones_like(amax))
        q = (x.float() / scale[:, None]).to(torch.float8_e4m3fn)
        return q.view(torch.uint8), scale
    if x.dtype != torch.bfloat16 or x.stride(1) != 1:
        raise _n.NativeError("fp8_rowwise_quant expects a bf16 matrix with unit inner stride")
    q = torch.empty((rows, cols), dtype=torch.uint8, device=x.device)
    s = torch.empty(rows, dtype=torch.float32, device=x.device)
    _n.call("sxb_fp8_rowwise_quant", x.data_ptr(), q.data_ptr(), s.data_ptr(), rows, cols, x.stride(0), q.stride(0),
            _n.stream_ptr())
    return q, s


def _deq(q: torch.Tensor, s: torch.Tensor) -> torch.Tensor:
    return q.view(torch.float8_e4m3fn).float() * s[:, None]


def fp8_gemm_nt(qa: torch.Tensor, sa: torch.Tensor, qb: torch.Tensor, sb: torch.Tensor,
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``(qa*sa) @ (qb*sb)^T`` -> bf16 ``[M, N]``; qa ``[M, K]``, qb ``[N, K]`` e4m3 bytes."""
    M, K = qa.shape
    N = qb.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=qa.device)
    if not qa.is_cuda:
        out.copy_((_deq(qa, sa) @ _deq(qb, sb).t()).to(out.dtype))
        return out
    _call(qa.data_ptr(), qb.data_ptr(), out.data_ptr(), M, N, K, GEMM_NT, qa.stride(0), qb.stride(0), out.stride(0),
          _n.FP8E4M3, _n.dtype_code(out.dtype), False, 1.0, None, None, PLAIN, 0, None, 0, 0, sa, sb)
    return out


# ------------------------------------------------------------------------------------ grouped
def aligned_offsets(counts: torch.Tensor, align: int = GROUP_ALIGN_M) -> torch.Tensor:
    """int32 ``[G+1]`` row offsets with every group padded to ``align`` rows — computed on the device, no sync."""
    padded = ((counts.to(torch.int64) + align - 1) // align) * align
    off = torch.zeros(counts.numel() + 1, dtype=torch.int32, device=counts.device)
    off[1:] = torch.cumsum(padded, 0).to(torch.int32)
    return off


def _groups_cpu(offsets):
    o = offsets.tolist()
    return [(g, o[g], o[g + 1]) for g in range(len(o) - 1)]


def grouped_gemm_nt(x, w, offsets, out=None, scale_x=None, qw=None, scale_w=None):
    """``out[rows_g] = x[rows_g] @ w[g]^T``; x ``[R, K]``, w ``[G, N, K]``, offsets int32 ``[G+1]`` (multiples of 256).
    With ``qw``/``scale_w`` (and ``x`` already e4m3 + ``scale_x``) the product runs in fp8."""
    G, N, K = w.shape if qw is None else qw.shape
    R = x.shape[0]
    if out is None:
        out = torch.empty((R, N), dtype=torch.bfloat16, device=x.device)
    if not x.is_cuda:
        for g, lo, hi in _groups_cpu(offsets):
            if hi > lo:
                if qw is None:
                    out[lo:hi] = (x[lo:hi].float() @ w[g].float().t()).to(out.dtype)
                else:
                    out[lo:hi] = (_deq(x[lo:hi], scale_x[lo:hi]) @ _deq(qw[g], scale_w[g]).t()).to(out.dtype)
        return out
    if qw is None:
        _call(x.data_ptr(), w.data_ptr(), out.data_ptr(), 0, N, K, GEMM_NT, x.stride(0), w.stride(1), out.stride(0),
              _n.BF16, _n.dtype_code(out.dtype), False, 1.0, None, None, RAGGED_M, G, offsets, R, 0, None, None)
    else:
        _call(x.data_ptr(), qw.data_ptr(), out.data_ptr(), 0, N, K, GEMM_NT, x.stride(0), qw.stride(1), out.stride(0),
              _n.FP8E4M3, _n.dtype_code(out.dtype), False, 1.0, None, None, RAGGED_M, G, offsets, R, 0,
              scale_x, scale_w.reshape(-1))
    return out


def grouped_gemm_nn(dy, w, offsets, out=None):
    """``out[rows_g] = dy[rows_g] @ w[g]``; dy ``[R, N]``, w ``[G, N, K]`` -> ``[R, K]`` (dgrad)."""
    G, N, K = w.shape
    R = dy.shape[0]
    if out is None:
        out = torch.empty((R, K), dtype=torch.bfloat16, device=dy.device)
    if not dy.is_cuda:
        for g, lo, hi in _groups_cpu(offsets):
            if hi > lo:
                out[lo:hi] = (dy[lo:hi].float() @ w[g].float()).to(out.dtype)
        return out
    # GEMM view: A = dy [R, N] (K_gemm = N), B_g = w[g] stored [N, K] = [K_gemm, N_gemm]
    _call(dy.data_ptr(), w.data_ptr(), out.data_ptr(), 0, K, N, GEMM_NN, dy.stride(0), w.stride(1), out.stride(0),
          _n.BF16, _n.dtype_code(out.dtype), False, 1.0, None, None, RAGGED_M, G, offsets, R, 0, None, None)
    return out


def grouped_gemm_tn(dy, x, offsets, out, accumulate=False):
    """``out[g] (+)= dy[rows_g]^T @ x[rows_g]``; dy ``[R, N]``, x ``[R, K]``, out ``[G, N, K]`` (wgrad)."""
    G, N, K = out.shape
    R = dy.shape[0]
    if not dy.is_cuda:
        for g, lo, hi in _groups_cpu(offsets):
            r = dy[lo:hi].float().t() @ x[lo:hi].float()
            out[g] = ((out[g].float() + r) if accumulate else r).to(out.dtype)
        return out
    _call(dy.data_ptr(), x.data_ptr(), out.data_ptr(), N, K, 0, GEMM_TN, dy.stride(0), x.stride(0), out.stride(1),
          _n.BF16, _n.dtype_code(out.dtype), accumulate, 1.0, None, None, RAGGED_K, G, offsets, R, out.stride(0),
          None, None)
    return out


def _as3d(w):
    return w if w.dim() == 3 else w.unsqueeze(0)


class _GroupedLinear(torch.autograd.Function):
    """y[rows_g] = x[rows_g] @ w[g]^T for all groups; backward = one grouped dgrad + one grouped wgrad launch.
    ``w`` is ``[G, N, K]`` or, for a single group, the ``[N, K]`` parameter itself; it may carry the main-grad protocol
    (``_sxb_main_grad`` view of the same shape)."""

    @staticmethod
    def forward(ctx, x, w, offsets, fp8, out):
        from . import functional as F
        w3 = _as3d(w)
        if fp8 and x.is_cuda:
            qx, sx = fp8_rowwise_quant(x)
            G, N, K = w3.shape
            qw, sw = fp8_rowwise_quant(w3.reshape(G * N, K))
            y = grouped_gemm_nt(qx, None, offsets, out=out, scale_x=sx, qw=qw.view(G, N, K), scale_w=sw.view(G, N))
        else:
            y = grouped_gemm_nt(x, w3, offsets, out=out)
        F._note_use(w)
        ctx.save_for_backward(x, w, offsets)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, offsets = ctx.saved_tensors
        if not dy.is_contiguous():
            dy = dy.contiguous()
        w3 = _as3d(w)
        dx = grouped_gemm_nn(dy, w3, offsets) if ctx.needs_input_grad[0] else None
        dw = None
        if ctx.needs_input_grad[1]:
            mg = getattr(w, "_sxb_main_grad", None)
            if mg is not None:
                first = not getattr(w, "_sxb_grad_written", False)
                grouped_gemm_tn(dy, x, offsets, _as3d(mg), accumulate=not first)
                w._sxb_grad_written = True
                hook = getattr(w, "_sxb_grad_ready", None)
                if hook is not None:
                    hook(w)
            else:
                dw = grouped_gemm_tn(dy, x, offsets, torch.empty_like(w3)).view(w.shape)
        return dx, dw, None, None, None


def grouped_linear(x: torch.Tensor, w: torch.Tensor, offsets: torch.Tensor, fp8: bool = False,
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x ``[R, K]`` (rows sorted by group, groups padded to 256 rows with ZERO rows), w ``[G, N, K]`` (or ``[N, K]``
    for one group).  ``out``: optional preallocated ``[R, N]`` destination (e.g. a symmetric-memory buffer)."""
    return _GroupedLinear.apply(x, w, offsets, fp8, out)
