"""Random layer-token-drop primitives: token sampling, sorted gather and scatter.

Reference: ``deepspeed/ops/random_ltd/dropping_utils.py:17-134`` — ``gpt_sample_tokens`` / ``bert_sample_tokens`` draw
``reserved_length`` distinct positions per sequence with ``torch.multinomial``, sort them (native ``token_sort_``) and
build the matching attention mask; ``GatherTokens`` / ``ScatterTokens`` are autograd functions over the native
``token_gather`` / ``token_scatter_``.  Kernels: ``csrc/quantize.cu`` (``sxb_token_sort`` bitonic sort in shared
memory, ``sxb_token_move`` 16-byte row copies).
"""
from __future__ import annotations

import torch

from ...native import lib as _lib

_lib.register({"sxb_token_sort": "piip", "sxb_token_move": "ppp" "iiiiiii" "p"})


def token_sort_(idx: torch.Tensor) -> torch.Tensor:
    """Sort every row of int32 ``idx`` [..., k] ascending, in place."""
    k = idx.shape[-1]
    if idx.is_cuda and k <= 4096:
        assert idx.dtype == torch.int32 and idx.is_contiguous()
        _lib.call("sxb_token_sort", idx.data_ptr(), idx.numel() // k, k, _lib.stream_ptr())
        return idx
    idx.copy_(torch.sort(idx, dim=-1).values)
    return idx


def _move(full: torch.Tensor, part: torch.Tensor, idx: torch.Tensor, batch_first: bool, scatter: bool):
    if batch_first:
        batch, seq, hidden = full.shape
    else:
        seq, batch, hidden = full.shape
    k = idx.shape[-1]
    if full.is_cuda and (hidden * full.element_size()) % 16 == 0:
        assert full.is_contiguous() and part.is_contiguous()
        idx = idx.to(torch.int32).contiguous()
        _lib.call("sxb_token_move", full.data_ptr(), part.data_ptr(), idx.data_ptr(), batch, seq, k, hidden,
                  full.element_size(), int(batch_first), int(scatter), _lib.stream_ptr())
        return
    ii = idx.long()
    if batch_first:
        g = ii[:, :, None].expand(batch, k, hidden)
        if scatter:
            full.scatter_(1, g, part)
        else:
            part.copy_(torch.gather(full, 1, g))
    else:
        g = ii.t()[:, :, None].expand(k, batch, hidden)
        if scatter:
            full.scatter_(0, g, part)
        else:
            part.copy_(torch.gather(full, 0, g))


def token_gather(activations: torch.Tensor, sorted_indices: torch.Tensor, batch_first: bool) -> torch.Tensor:
    k = sorted_indices.shape[-1]
    if batch_first:
        out = activations.new_empty(activations.shape[0], k, activations.shape[2])
    else:
        out = activations.new_empty(k, activations.shape[1], activations.shape[2])
    _move(activations.contiguous(), out, sorted_indices, batch_first, scatter=False)
    return out


def token_scatter_(all_activations: torch.Tensor, layer_activations: torch.Tensor, sorted_indices: torch.Tensor,
                   batch_first: bool) -> torch.Tensor:
    _move(all_activations, layer_activations.contiguous(), sorted_indices, batch_first, scatter=True)
    return all_activations


def gpt_sample_tokens(reserved_length: int, seq_length: int, batch_size: int, layers: int = 1,
                      device: str = "cpu", attn_mask: torch.Tensor = None):
    prob = torch.ones((layers * batch_size, seq_length), device=device)
    sampled = torch.multinomial(prob, reserved_length).reshape(layers, batch_size, reserved_length).to(torch.int32)
    sampled = token_sort_(sampled.contiguous())
    new_mask = attn_mask[:, :, :reserved_length, :reserved_length] if attn_mask is not None else None
    return sampled, new_mask


def bert_sample_tokens(reserved_length: int, seq_length: int, batch_size: int, layers: int = 1,
                       device: str = "cpu", attn_mask: torch.Tensor = None):
    assert attn_mask is not None
    prob = torch.ones((layers * batch_size, seq_length), device=device)
    sampled = torch.multinomial(prob, reserved_length).reshape(layers, batch_size, reserved_length).to(torch.int32)
    sampled = token_sort_(sampled.contiguous())
    dtype = sampled.dtype
    sampled = sampled.to(torch.long)
    new_mask = []
    for l in range(layers):
        rows = []
        for b in range(batch_size):
            sel = sampled[l][b]
            rows.append(attn_mask[b:b + 1, :, sel, :][:, :, :, sel])
        new_mask.append(torch.cat(rows, dim=0))
    return sampled.to(dtype), new_mask


class GatherTokens(torch.autograd.Function):
    @staticmethod
    def forward(ctx, activations: torch.Tensor, sorted_indices: torch.Tensor, batch_first: bool):
        ctx.save_for_backward(activations, sorted_indices)
        ctx.batch_first = batch_first
        return activations, token_gather(activations, sorted_indices, batch_first)

    @staticmethod
    def backward(ctx, a_gradients: torch.Tensor, g_gradients: torch.Tensor):
        g_gradients = g_gradients.contiguous()
        a_gradients = a_gradients.contiguous().clone()
        activations, sorted_indices = ctx.saved_tensors
        # the dropped-path gradient is ADDED at the sampled positions
        tmp = token_gather(a_gradients, sorted_indices, ctx.batch_first)
        token_scatter_(a_gradients, tmp + g_gradients, sorted_indices, ctx.batch_first)
        return a_gradients, None, None


class ScatterTokens(torch.autograd.Function):
    @staticmethod
    def forward(ctx, all_activations: torch.Tensor, layer_activations: torch.Tensor, sorted_indices: torch.Tensor,
                batch_first: bool):
        out = token_scatter_(all_activations.clone().contiguous(), layer_activations, sorted_indices, batch_first)
        ctx.save_for_backward(sorted_indices)
        ctx.batch_first = batch_first
        return out

    @staticmethod
    def backward(ctx, out_gradients: torch.Tensor):
        out_gradients = out_gradients.contiguous()
        sorted_indices, = ctx.saved_tensors
        ret_val = token_gather(out_gradients, sorted_indices, ctx.batch_first)
        # positions overwritten by the layer output receive no gradient through the bypass
        a_grad = out_gradients.clone()
        zeros = torch.zeros_like(ret_val)
        token_scatter_(a_grad, zeros, sorted_indices, ctx.batch_first)
        return a_grad, ret_val, None, None
