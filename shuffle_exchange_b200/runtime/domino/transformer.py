"""Domino: tensor parallelism that hides its all-reduces by splitting every batch into two micro-batches and
interleaving them — while micro-batch 0's row-parallel output is being all-reduced, micro-batch 1 computes
(reference: ``runtime/domino/transformer.py:250-409``, https://arxiv.org/abs/2409.15241).

``DominoTransformerLayer`` is a pre-LN decoder layer with column-parallel QKV / up projections and row-parallel
output / down projections over ``tp_group``; collectives are issued asynchronously and waited for just before their
result is needed.  On one rank (or without a group) it is an ordinary transformer layer.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as TF

from ... import comm as dist
from ...ops import functional as F
from .async_linear import DominoAsyncColumnParallelLinear


class DominoUtil:
    """Shared dictionary of in-flight all-reduce handles keyed by a batch id."""
    HANDLE_DIC = {"0": None, "1": None, "2": None, "3": None}


def _tp(group):
    return dist.get_world_size(group=group) if (group is not None and dist.is_initialized()) else 1


class _AsyncReduce:
    """In-place sum of a row-parallel partial result over the group, waited for by the consumer.  The reduction runs on
    the detached storage: for autograd it is the identity (every rank back-propagates the same replicated loss)."""

    def __init__(self, t, group):
        self.t = t
        self.h = dist.all_reduce(t.detach(), group=group, async_op=True) if _tp(group) > 1 else None

    def wait(self):
        if self.h is not None:
            self.h.wait()
        return self.t


class _WaitInputGrad(torch.autograd.Function):
    """Identity placed in front of a column-parallel linear.  Its backward runs when autograd reaches the producer of
    the linear's input: only there is the asynchronous all-reduce of the input gradient (issued by the linear's backward
    and parked in ``handle_dic[key]``) waited for — the other micro-batch's backward computes in between."""

    @staticmethod
    def forward(ctx, x, handle_dic, key):
        ctx.handle_dic, ctx.key = handle_dic, key
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        h = ctx.handle_dic.get(ctx.key)
        if h is not None:
            h.wait()
            ctx.handle_dic[ctx.key] = None
        return g, None, None


class _RowParallel(torch.autograd.Function):
    """y = x W^T summed over the group, asynchronously; backward is identity on the reduction."""

    @staticmethod
    def forward(ctx, x, weight):
        ctx.save_for_backward(x, weight)
        return F.matmul_nt(x, weight)

    @staticmethod
    def backward(ctx, g):
        x, w = ctx.saved_tensors
        return F.matmul_nn(g, w), F.matmul_tn(g, x)


class DominoTransformerLayer(nn.Module):
    def __init__(self, hidden_size: int, num_heads: int, ffn_hidden_size: Optional[int] = None, tp_group=None,
                 layernorm_epsilon: float = 1e-5, attention_dropout: float = 0.0, hidden_dropout: float = 0.0):
        super().__init__()
        self.tp_group = tp_group
        tp = _tp(tp_group)
        assert num_heads % tp == 0
        self.h, self.heads_local, self.hd = hidden_size, num_heads // tp, hidden_size // num_heads
        ffn = ffn_hidden_size or 4 * hidden_size
        self.input_layernorm = nn.LayerNorm(hidden_size, eps=layernorm_epsilon)
        self.post_attention_layernorm = nn.LayerNorm(hidden_size, eps=layernorm_epsilon)
        self.query_key_value = DominoAsyncColumnParallelLinear(hidden_size, 3 * hidden_size, tp_group)
        self.dense = nn.Parameter(torch.empty(hidden_size, hidden_size // tp))
        self.dense_bias = nn.Parameter(torch.zeros(hidden_size))
        self.dense_h_to_4h = DominoAsyncColumnParallelLinear(hidden_size, ffn, tp_group)
        self.dense_4h_to_h = nn.Parameter(torch.empty(hidden_size, ffn // tp))
        self.mlp_bias = nn.Parameter(torch.zeros(hidden_size))
        nn.init.xavier_normal_(self.dense)
        nn.init.xavier_normal_(self.dense_4h_to_h)
        self.attn_p, self.hidden_p = attention_dropout, hidden_dropout

    def _attn_partial(self, x, key):
        B, S, _ = x.shape
        qkv = self.query_key_value(x, DominoUtil.HANDLE_DIC, key).view(B, S, 3, self.heads_local, self.hd)
        q, k, v = qkv.unbind(dim=2)                       # [B, S, heads, hd] each
        if self.attn_p > 0.0 and self.training:
            # attention dropout has no fused kernel here: PyTorch attention for that configuration only
            ctx = TF.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), is_causal=True,
                                                  dropout_p=self.attn_p).transpose(1, 2)
        else:
            ctx = F.flash_attention(q.contiguous(), k.contiguous(), v.contiguous(), causal=True)
        return _RowParallel.apply(ctx.reshape(B, S, self.heads_local * self.hd), self.dense)

    def _mlp_partial(self, x, key):
        return _RowParallel.apply(TF.gelu(self.dense_h_to_4h(x, DominoUtil.HANDLE_DIC, key)), self.dense_4h_to_h)

    def forward(self, hidden_states, attention_mask=None, rotary_pos_emb=None):
        """``hidden_states`` [B, S, H]; the batch is cut into two halves that ping-pong between compute and
        communication."""
        B = hidden_states.shape[0]
        halves = hidden_states.chunk(2, dim=0) if B >= 2 else (hidden_states,)
        # ---- attention: issue half 0's reduce, compute half 1 meanwhile.  Both halves' inputs are wrapped BEFORE
        # either half computes, so that in backward half 1's input-gradient all-reduce overlaps half 0's backward.
        D = DominoUtil.HANDLE_DIC
        ln = [_WaitInputGrad.apply(self.input_layernorm(h), D, str(i)) for i, h in enumerate(halves)]
        parts = []
        for i, x in enumerate(ln):
            parts.append(_AsyncReduce(self._attn_partial(x, str(i)), self.tp_group))
        attn_out = []
        for h, p in zip(halves, parts):
            o = p.wait() + self.dense_bias
            attn_out.append(h + TF.dropout(o, self.hidden_p, self.training))
        # ---- MLP, same interleave
        ln = [_WaitInputGrad.apply(self.post_attention_layernorm(h), D, str(2 + i)) for i, h in enumerate(attn_out)]
        parts = []
        for i, x in enumerate(ln):
            parts.append(_AsyncReduce(self._mlp_partial(x, str(2 + i)), self.tp_group))
        outs = []
        for h, p in zip(attn_out, parts):
            o = p.wait() + self.mlp_bias
            outs.append(h + TF.dropout(o, self.hidden_p, self.training))
        return torch.cat(outs, dim=0) if len(outs) > 1 else outs[0]


class DominoTransformer(nn.Module):
    """Stack of Domino layers with the reference's two schedules (``transformer.py:467-605``):

    * ``intra_layer_overlap=True`` — every layer ping-pongs its two half-batches and joins them at its end;
    * inter-layer overlap (the default, as in the reference) — the half-batches are never joined between layers, so the
      last all-reduce of layer ``i`` for one half overlaps the first GEMMs of layer ``i + 1`` for the other half.

    The inter-layer schedule is written as one generator per half-batch that runs through the WHOLE stack and yields
    each time it has issued an asynchronous all-reduce; :meth:`_interleave` resumes the halves round-robin, so a half
    waits for its reduction only after the other half has been given the device.  The order of operations this
    produces is the reference's hand-unrolled one; the layer count and the number of halves are free."""

    def __init__(self, num_layers: int, hidden_size: int, num_heads: int, ffn_hidden_size: Optional[int] = None,
                 tp_group=None, post_layer_norm: bool = True, intra_layer_overlap: bool = False, **kw):
        super().__init__()
        self.layers = nn.ModuleList([DominoTransformerLayer(hidden_size, num_heads, ffn_hidden_size, tp_group, **kw)
                                     for _ in range(num_layers)])
        self.final_layernorm = nn.LayerNorm(hidden_size) if post_layer_norm else None
        self.tp_group = tp_group
        self.intra_layer_overlap = intra_layer_overlap

    def forward(self, hidden_states, attention_mask=None, rotary_pos_emb=None):
        if self.intra_layer_overlap or hidden_states.shape[0] < 2:
            return self.intra_layer_overlap_forward(hidden_states, attention_mask, rotary_pos_emb)
        return self.inter_layer_overlap_forward(hidden_states, attention_mask, rotary_pos_emb)

    def intra_layer_overlap_forward(self, hidden_states, attention_mask=None, rotary_pos_emb=None):
        for layer in self.layers:
            hidden_states = layer(hidden_states, attention_mask, rotary_pos_emb)
        return self.final_layernorm(hidden_states) if self.final_layernorm is not None else hidden_states

    def _half_batch_stream(self, h, mb: int):
        """One half-batch through every layer; yields after each issued all-reduce, returns the final hidden state."""
        D = DominoUtil.HANDLE_DIC
        k_attn, k_mlp = str(mb), str(2 + mb)
        for layer in self.layers:
            x = _WaitInputGrad.apply(layer.input_layernorm(h), D, k_attn)
            pending = _AsyncReduce(layer._attn_partial(x, k_attn), layer.tp_group)
            yield
            h = h + TF.dropout(pending.wait() + layer.dense_bias, layer.hidden_p, layer.training)
            x = _WaitInputGrad.apply(layer.post_attention_layernorm(h), D, k_mlp)
            pending = _AsyncReduce(layer._mlp_partial(x, k_mlp), layer.tp_group)
            yield
            h = h + TF.dropout(pending.wait() + layer.mlp_bias, layer.hidden_p, layer.training)
        return self.final_layernorm(h) if self.final_layernorm is not None else h

    @staticmethod
    def _interleave(streams):
        """Resume the generators in turn until all have returned; -> their return values in order."""
        results = [None] * len(streams)
        live = list(enumerate(streams))
        while live:
            still = []
            for i, g in live:
                try:
                    next(g)
                    still.append((i, g))
                except StopIteration as done:
                    results[i] = done.value
            live = still
        return results

    def inter_layer_overlap_forward(self, hidden_states, attention_mask=None, rotary_pos_emb=None):
        halves = hidden_states.chunk(2, dim=0)
        outs = self._interleave([self._half_batch_stream(h, mb) for mb, h in enumerate(halves)])
        return torch.cat(outs, dim=0)
