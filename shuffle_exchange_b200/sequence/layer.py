"""Ulysses sequence parallelism: ``DistributedAttention``.

Inputs arrive sequence-sharded; before attention q/k/v go through an all-to-all that scatters heads
and gathers the sequence so every rank runs ordinary full-sequence attention on ``h/P`` heads, and a
second all-to-all restores the sequence sharding (reference: deepspeed/sequence/layer.py —
``single_all_to_all`` :221-254, uneven heads :111-218, ``_SeqAllToAll`` :277-328, ``DistributedAttention``
:331-440).  Differences: q, k and v travel in ONE all-to-all when their head counts allow it (the
reference issues three and carries a TODO to merge them, :388), layouts are computed by one helper
for any (scatter_idx, gather_idx), and on CUDA the transport can be the symmetric-memory row all-to-all
(``comm/symm.py``) instead of NCCL.
"""
from __future__ import annotations

from typing import Any, Optional, Tuple

import torch
from torch import Tensor
from torch.nn import Module

from ..comm import comm as dist


def pre_all2all_fun(P: int, scatter_idx: int):
    """Layout step in front of the equal-split all-to-all: split dim ``scatter_idx`` into ``[P, s/P]`` and bring the
    ``P`` axis to the front, contiguous (one block per destination rank)."""
    def pre(x: Tensor) -> Tensor:
        assert x.shape[scatter_idx] % P == 0, f"dim {scatter_idx} ({x.shape[scatter_idx]}) not divisible by sp={P}"
        shp = list(x.shape)
        shp[scatter_idx: scatter_idx + 1] = [P, shp[scatter_idx] // P]
        return x.reshape(shp).movedim(scatter_idx, 0).contiguous()
    return pre


def post_all2all(gather_idx: int):
    """Layout step behind it: ``out[p]`` is the block received from rank ``p``; concatenate the blocks along
    ``gather_idx``."""
    def post(out: Tensor) -> Tensor:
        out = out.movedim(0, gather_idx)               # [.., P, gather, ..]
        shp = list(out.shape)
        shp[gather_idx: gather_idx + 2] = [shp[gather_idx] * shp[gather_idx + 1]]
        return out.reshape(shp)
    return post


def _a2a_equal(x: Tensor, scatter_idx: int, gather_idx: int, group) -> Tensor:
    """All-to-all that splits dim ``scatter_idx`` into P parts and concatenates the received parts
    along ``gather_idx`` (P = group size).  Works for any tensor rank."""
    P = dist.get_world_size(group)
    if P == 1:
        return x
    t = pre_all2all_fun(P, scatter_idx)(x)
    out = torch.empty_like(t)
    dist.all_to_all_single(out, t, group=group)
    return post_all2all(gather_idx)(out)


def _a2a_uneven_heads(x: Tensor, scatter_idx: int, gather_idx: int, group, num_heads: int) -> Tensor:
    """Head counts not divisible by P (GQA/MQA): rank r receives ``ceil``/``floor`` shares of the heads."""
    P = dist.get_world_size(group)
    r = dist.get_rank(group)
    base, extra = divmod(num_heads, P)
    split = [base + (1 if i < extra else 0) for i in range(P)]
    if scatter_idx != gather_idx and x.shape[scatter_idx] == num_heads:
        # forward direction: scatter heads, gather sequence
        parts = [p.contiguous() for p in x.split(split, dim=scatter_idx)]
        recv = [torch.empty_like(parts[r]) for _ in range(P)]
        dist.all_to_all(recv, parts, group=group)
        return torch.cat(recv, dim=gather_idx)
    # reverse: scatter sequence, gather heads
    parts = [p.contiguous() for p in x.chunk(P, dim=scatter_idx)]
    recv = []
    for i in range(P):
        shp = list(parts[0].shape)
        shp[gather_idx] = split[i]
        recv.append(torch.empty(shp, dtype=x.dtype, device=x.device))
    dist.all_to_all(recv, parts, group=group)
    return torch.cat(recv, dim=gather_idx)


def uneven_heads_all2all(input: Tensor, scatter_idx: int, gather_idx: int, batch_dim_idx: int, group,
                         num_heads: Optional[int] = None) -> Tensor:
    """Ulysses all-to-all for head counts that do not divide the sequence-parallel size (reference
    ``sequence/layer.py:110-215``): the first ``num_heads % P`` ranks take one head more.  Forward direction (heads
    scattered, sequence gathered) infers ``num_heads`` from the tensor; the reverse direction needs it."""
    if num_heads is None:
        if scatter_idx >= 2:
            num_heads = input.shape[scatter_idx]
        else:
            from ..module_inject.tp_shard import get_num_attention_heads
            num_heads = get_num_attention_heads()
            if not num_heads:
                raise ValueError("reverse uneven all-to-all needs num_heads")
    return _a2a_uneven_heads(input, scatter_idx, gather_idx, group, num_heads)


def single_all_to_all(input: Tensor, scatter_idx: int, gather_idx: int, batch_dim_idx: int, group,
                      async_op: bool = False, handle=None, type=None) -> Tensor:
    return _a2a_equal(input, scatter_idx, gather_idx, group)


class _SeqAllToAll(torch.autograd.Function):
    @staticmethod
    def forward(ctx: Any, group, input: Tensor, scatter_idx: int, gather_idx: int, batch_dim_idx: int = 0,
                stream=None, handle=None, type=None, is_fwd: bool = True, num_heads: Optional[int] = None) -> Tensor:
        ctx.group, ctx.scatter_idx, ctx.gather_idx = group, scatter_idx, gather_idx
        ctx.batch_dim_idx, ctx.num_heads = batch_dim_idx, num_heads
        P = dist.get_world_size(group)
        if num_heads is not None and num_heads % P != 0:
            return _a2a_uneven_heads(input, scatter_idx, gather_idx, group, num_heads)
        return _a2a_equal(input, scatter_idx, gather_idx, group)

    @staticmethod
    def backward(ctx: Any, *grad_output: Tensor):
        g = _SeqAllToAll.apply(ctx.group, grad_output[0], ctx.gather_idx, ctx.scatter_idx, ctx.batch_dim_idx,
                               None, None, None, False, ctx.num_heads)
        return (None, g, None, None, None, None, None, None, None, None)


class DistributedAttention(Module):
    """Wrap any local attention ``local_attn(q, k, v, *args)`` so it runs under Ulysses SP.

    Tensors are ``[s/P, b, h, d]`` (``scatter_idx=2, gather_idx=0``, Megatron layout) or ``[b, s/P, h, d]``
    (``gather_idx=1``).  After the first all-to-all every rank holds ``[s, b, h/P, d]``.
    """

    def __init__(self, local_attention: Module, sequence_process_group, scatter_idx: int = 2, gather_idx: int = 0,
                 sp_stream=None) -> None:
        super().__init__()
        self.local_attn = local_attention
        self.spg = sequence_process_group
        self.scatter_idx = scatter_idx
        self.gather_idx = gather_idx
        self.sp_overlap_comm = sp_stream is not None
        self.sp_stream = sp_stream
        self.overlap_handles = None

    def layer_sync(self, layer):
        if self.sp_overlap_comm and hasattr(layer, "done_event"):
            torch.cuda.current_stream().wait_event(layer.done_event)

    def forward(self, query: Tensor, key: Tensor, value: Tensor, batch_dim_idx: int = None, rotary_pos_emb=None,
                *args: Any, **kwargs) -> Tensor:
        if batch_dim_idx is None:
            batch_dim_idx = 1 if self.gather_idx == 0 else 0
        hq, hk = query.shape[self.scatter_idx], key.shape[self.scatter_idx]
        P = dist.get_world_size(self.spg)
        if hq == hk and hq % P == 0 and query.shape == key.shape == value.shape:
            # one collective for q|k|v: stack on a new leading axis that the layout helper ignores
            qkv = torch.stack([query, key, value], dim=0)
            out = _SeqAllToAll.apply(self.spg, qkv, self.scatter_idx + 1, self.gather_idx + 1, batch_dim_idx + 1)
            q, k, v = out[0], out[1], out[2]
        else:
            q = _SeqAllToAll.apply(self.spg, query, self.scatter_idx, self.gather_idx, batch_dim_idx, None, None,
                                   None, True, hq)
            k = _SeqAllToAll.apply(self.spg, key, self.scatter_idx, self.gather_idx, batch_dim_idx, None, None,
                                   None, True, hk)
            v = _SeqAllToAll.apply(self.spg, value, self.scatter_idx, self.gather_idx, batch_dim_idx, None, None,
                                   None, True, hk)
        if rotary_pos_emb is not None:
            pos_cos, pos_sin = rotary_pos_emb
            q = apply_rotary_pos_emb(q, pos_cos, pos_sin)
            k = apply_rotary_pos_emb(k, pos_cos, pos_sin)
        context = self.local_attn(q, k, v, *args, **kwargs)
        return _SeqAllToAll.apply(self.spg, context, self.gather_idx, self.scatter_idx, batch_dim_idx, None, None,
                                  None, False, hq)


def _rotate_half(x):
    x1, x2 = x[..., : x.shape[-1] // 2], x[..., x.shape[-1] // 2:]
    return torch.cat((-x2, x1), dim=-1)


def apply_rotary_pos_emb(t, freqs_cos, freqs_sin):
    """Rotate the leading ``rot_dim`` features of every head (reference: sequence/layer.py:84-103)."""
    rot_dim = freqs_cos.shape[-1]
    t_rot, t_pass = t[..., :rot_dim], t[..., rot_dim:]
    t_rot = (t_rot * freqs_cos) + (_rotate_half(t_rot) * freqs_sin)
    return t_rot if t_pass.shape[-1] == 0 else torch.cat((t_rot, t_pass), dim=-1)
