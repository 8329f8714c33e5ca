"""Token routing primitives for mixture-of-experts layers (training and inference).

Kernels: ``csrc/moe_ragged.cu`` — ``sxb_top_k_gating`` (softmax over all experts + top-k + per-expert arrival order),
``sxb_moe_scatter`` (dropless grouping of token rows by expert), ``sxb_moe_gather`` (score-weighted combine) and its
backward.  They replace the dense one-hot einsum dispatch/combine of the reference (``moe/sharded_moe.py:609-610,670``:
O(S·E·C·M) FLOPs and memory) and are the same operations as its inference kernels
(``inference/v2/kernels/ragged_ops/{top_k_gating,moe_scatter,moe_gather}``).  Each op has a PyTorch branch with the
same semantics, used on CPU and as the oracle in tests.
"""
from __future__ import annotations

from typing import Tuple

import torch

from ..native import lib as _lib

_lib.register({
    "sxb_top_k_gating": "ppppp" "iiii" "i" "p",
    "sxb_moe_scatter": "ppppppp" "iiiii" "p",
    "sxb_moe_gather": "pppp" "iii" "p",
    "sxb_moe_gather_bwd": "pppppp" "iii" "p",
})


def top_k_gating(logits: torch.Tensor, top_k: int, renormalize: bool = True):
    """``logits`` [T, E] -> ``(scores fp32 [T,K], assignments int32 [T,K], offsets int32 [T,K], counts int32 [E])``.

    ``offsets[t,k]`` is the arrival index of token ``t`` inside expert ``assignments[t,k]`` (a permutation of
    ``range(counts[e])`` per expert; on CUDA the order is nondeterministic, on CPU it is token order)."""
    T, E = logits.shape
    dev = logits.device
    scores = torch.empty(T, top_k, dtype=torch.float32, device=dev)
    assign = torch.empty(T, top_k, dtype=torch.int32, device=dev)
    offsets = torch.empty(T, top_k, dtype=torch.int32, device=dev)
    counts = torch.empty(E, dtype=torch.int32, device=dev)
    if logits.is_cuda:
        logits = logits.contiguous()
        _lib.call("sxb_top_k_gating", logits.data_ptr(), scores.data_ptr(), assign.data_ptr(), offsets.data_ptr(),
                  counts.data_ptr(), T, E, top_k, int(renormalize), _lib.dtype_code(logits.dtype), _lib.stream_ptr())
        return scores, assign, offsets, counts
    probs = torch.softmax(logits.float(), dim=-1)
    s, a = torch.topk(probs, top_k, dim=-1)
    if renormalize:
        s = s / s.sum(dim=-1, keepdim=True)
    flat = a.reshape(-1)
    onehot = torch.nn.functional.one_hot(flat, E)
    order = (torch.cumsum(onehot, dim=0) - 1).gather(1, flat[:, None]).squeeze(1)
    scores.copy_(s)
    assign.copy_(a.to(torch.int32))
    offsets.copy_(order.view(T, top_k).to(torch.int32))
    counts.copy_(onehot.sum(0).to(torch.int32))
    return scores, assign, offsets, counts


def _scatter_fwd(x, assign, offsets, counts):
    T, H = x.shape
    K = assign.shape[1]
    E = counts.shape[0]
    moe_in = torch.empty(T * K, H, dtype=x.dtype, device=x.device)
    cumsum = torch.empty(E, dtype=torch.int64, device=x.device)
    slots = torch.empty(T, K, dtype=torch.int32, device=x.device)
    if x.is_cuda:
        x = x.contiguous()
        _lib.call("sxb_moe_scatter", x.data_ptr(), moe_in.data_ptr(), assign.data_ptr(), offsets.data_ptr(),
                  counts.data_ptr(), cumsum.data_ptr(), slots.data_ptr(), T, E, K, H, x.element_size(), _lib.stream_ptr())
        return moe_in, cumsum, slots
    cumsum.copy_(torch.cumsum(counts.long(), 0))
    base = torch.cat([cumsum.new_zeros(1), cumsum[:-1]])
    sl = base[assign.long()] + offsets.long()
    slots.copy_(sl.to(torch.int32))
    moe_in[sl.reshape(-1)] = x.repeat_interleave(K, dim=0)
    return moe_in, cumsum, slots


def _gather_fwd(moe_out, scores, slots):
    T, K = slots.shape
    H = moe_out.shape[1]
    if moe_out.is_cuda and moe_out.dtype == torch.bfloat16:
        out = torch.empty(T, H, dtype=moe_out.dtype, device=moe_out.device)
        _lib.call("sxb_moe_gather", moe_out.data_ptr(), out.data_ptr(), None if scores is None else scores.data_ptr(),
                  slots.data_ptr(), T, K, H, _lib.stream_ptr())
        return out
    rows = moe_out[slots.long().reshape(-1)].view(T, K, H).float()
    if scores is not None:
        rows = rows * scores[:, :, None]
    return rows.sum(1).to(moe_out.dtype)


class _MoEScatter(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, assign, offsets, counts):
        moe_in, cumsum, slots = _scatter_fwd(x, assign, offsets, counts)
        ctx.save_for_backward(slots)
        ctx.mark_non_differentiable(cumsum, slots)
        return moe_in, cumsum, slots

    @staticmethod
    def backward(ctx, d_moe_in, _c, _s):
        slots, = ctx.saved_tensors
        return _gather_fwd(d_moe_in.contiguous(), None, slots), None, None, None


class _MoEGather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, moe_out, scores, slots):
        ctx.save_for_backward(moe_out, scores, slots)
        return _gather_fwd(moe_out, scores, slots)

    @staticmethod
    def backward(ctx, d_out):
        moe_out, scores, slots = ctx.saved_tensors
        T, K = slots.shape
        H = moe_out.shape[1]
        d_out = d_out.contiguous()
        if moe_out.is_cuda and moe_out.dtype == torch.bfloat16:
            d_moe = torch.empty_like(moe_out)
            d_scores = torch.empty(T, K, dtype=torch.float32, device=moe_out.device)
            _lib.call("sxb_moe_gather_bwd", d_out.data_ptr(), moe_out.data_ptr(), scores.data_ptr(), slots.data_ptr(),
                      d_moe.data_ptr(), d_scores.data_ptr(), T, K, H, _lib.stream_ptr())
            return d_moe, d_scores, None
        sl = slots.long().reshape(-1)
        d_moe = torch.zeros_like(moe_out)
        d_moe[sl] = (d_out.float()[:, None, :] * scores[:, :, None]).reshape(T * K, H).to(moe_out.dtype)
        d_scores = (moe_out[sl].view(T, K, H).float() * d_out.float()[:, None, :]).sum(-1)
        return d_moe, d_scores, None


def moe_scatter(x: torch.Tensor, assign: torch.Tensor, offsets: torch.Tensor, counts: torch.Tensor):
    """``x`` [T, H] -> ``(moe_input [T*K, H] grouped by expert, expert_cumsum int64 [E], mapped_slots int32 [T, K])``."""
    return _MoEScatter.apply(x, assign, offsets, counts)


def moe_gather(moe_output: torch.Tensor, scores: torch.Tensor, mapped_slots: torch.Tensor) -> torch.Tensor:
    """``out[t] = sum_k scores[t,k] * moe_output[mapped_slots[t,k]]`` (differentiable in both tensor arguments)."""
    return _MoEGather.apply(moe_output, scores, mapped_slots)
