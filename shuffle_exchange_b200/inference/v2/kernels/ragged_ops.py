"""Python entry points of the ragged-batch kernels (``csrc/moe_ragged.cu``).

Reference modules: ``inference/v2/kernels/ragged_ops/{embed,logits_gather,linear_blocked_kv_rotary,blocked_flash}``.
Every wrapper has a PyTorch branch with identical semantics (CPU execution and test oracle).
"""
from __future__ import annotations

import math

import torch

from ....native import lib as _lib

_lib.register({
    "sxb_ragged_embed": "ppppp" "iii" "p",
    "sxb_logits_gather": "ppp" "ii" "p",
    "sxb_kv_rotary_blocked": "ppppp" "iiiiiii" "f" "p",
    "sxb_paged_attention": "pppppp" "iiiiii" "l" "f" "p",
})


def ragged_embed(ids: torch.Tensor, table: torch.Tensor, positions: torch.Tensor = None,
                 pos_table: torch.Tensor = None, pos_offset: int = 0) -> torch.Tensor:
    """``out[t] = table[ids[t]] (+ pos_table[positions[t] + pos_offset])``; ids/positions int32 [T]."""
    T, H = ids.numel(), table.shape[1]
    if table.is_cuda and table.dtype == torch.bfloat16 and (H * 2) % 16 == 0:
        out = torch.empty(T, H, dtype=table.dtype, device=table.device)
        _lib.call("sxb_ragged_embed", ids.data_ptr(), None if positions is None else positions.data_ptr(),
                  table.data_ptr(), None if pos_table is None else pos_table.data_ptr(), out.data_ptr(), T, H,
                  pos_offset, _lib.stream_ptr())
        return out
    out = table[ids.long()]
    if pos_table is not None:
        out = out + pos_table[positions.long() + pos_offset]
    return out


def logits_gather(hidden: torch.Tensor, last_token: torch.Tensor) -> torch.Tensor:
    """Rows of ``hidden`` [T, H] at the last token of every sequence (``last_token`` int32 [n_seqs])."""
    if hidden.is_cuda and hidden.dtype == torch.bfloat16 and (hidden.shape[1] * 2) % 16 == 0:
        out = torch.empty(last_token.numel(), hidden.shape[1], dtype=hidden.dtype, device=hidden.device)
        _lib.call("sxb_logits_gather", hidden.data_ptr(), last_token.data_ptr(), out.data_ptr(), last_token.numel(),
                  hidden.shape[1], _lib.stream_ptr())
        return out
    return hidden[last_token.long()]


def _rope_ref(x: torch.Tensor, pos: torch.Tensor, rot_dim: int, theta: float) -> torch.Tensor:
    """HF rotate_half convention on the first ``rot_dim`` channels of x [T, H, D]."""
    half = rot_dim // 2
    inv = theta ** (-torch.arange(0, half, dtype=torch.float32, device=x.device) * 2.0 / rot_dim)
    ang = pos.float()[:, None] * inv[None, :]
    cos, sin = ang.cos()[:, None, :], ang.sin()[:, None, :]
    x0, x1 = x[..., :half].float(), x[..., half:rot_dim].float()
    out = x.clone()
    out[..., :half] = (x0 * cos - x1 * sin).to(x.dtype)
    out[..., half:rot_dim] = (x1 * cos + x0 * sin).to(x.dtype)
    return out


def kv_rotary_blocked(qkv: torch.Tensor, cache: torch.Tensor, tok_seq: torch.Tensor, tok_pos: torch.Tensor,
                      block_tables: torch.Tensor, n_q: int, n_kv: int, head_dim: int, rot_dim: int = None,
                      theta: float = 10000.0) -> None:
    """In place: rotate q and k inside ``qkv`` [T, (n_q + 2 n_kv) * D]; append rotated k and v to the paged ``cache``
    [n_blocks, block_size, 2, n_kv, D] at each token's (sequence, position)."""
    rot_dim = head_dim if rot_dim is None else rot_dim
    T = qkv.shape[0]
    block_size = cache.shape[1]
    if qkv.is_cuda and qkv.dtype == torch.bfloat16:
        _lib.call("sxb_kv_rotary_blocked", qkv.data_ptr(), cache.data_ptr(), tok_seq.data_ptr(), tok_pos.data_ptr(),
                  block_tables.data_ptr(), T, block_tables.shape[1], block_size, n_q, n_kv, head_dim, rot_dim,
                  float(theta), _lib.stream_ptr())
        return
    D = head_dim
    v3 = qkv.view(T, n_q + 2 * n_kv, D)
    v3[:, : n_q + n_kv] = _rope_ref(v3[:, : n_q + n_kv], tok_pos, rot_dim, theta)
    blk = block_tables[tok_seq.long(), (tok_pos // block_size).long()].long()
    slot = tok_pos.long() % block_size
    cache[blk, slot, 0] = v3[:, n_q: n_q + n_kv]
    cache[blk, slot, 1] = v3[:, n_q + n_kv:]


def paged_attention(q: torch.Tensor, cache: torch.Tensor, tok_seq: torch.Tensor, tok_pos: torch.Tensor,
                    block_tables: torch.Tensor, scale: float = None) -> torch.Tensor:
    """Causal attention of every token against its sequence's cached keys/values (positions 0..tok_pos).
    ``q`` [T, n_q, D] (may be a strided view of the fused qkv buffer)."""
    T, n_q, D = q.shape
    n_kv, block_size = cache.shape[3], cache.shape[1]
    scale = 1.0 / math.sqrt(D) if scale is None else scale
    if q.is_cuda and q.dtype == torch.bfloat16 and D in (64, 128) and q.stride(2) == 1 and q.stride(1) == D:
        out = torch.empty(T, n_q, D, dtype=q.dtype, device=q.device)
        _lib.call("sxb_paged_attention", q.data_ptr(), out.data_ptr(), cache.data_ptr(), tok_seq.data_ptr(),
                  tok_pos.data_ptr(), block_tables.data_ptr(), T, block_tables.shape[1], block_size, n_q, n_kv, D,
                  q.stride(0), float(scale), _lib.stream_ptr())
        return out
    out = torch.empty(T, n_q, D, dtype=q.dtype, device=q.device)
    group = n_q // n_kv
    for t in range(T):
        n = int(tok_pos[t]) + 1
        pos = torch.arange(n, device=q.device)
        blk = block_tables[int(tok_seq[t]), (pos // block_size)].long()
        k = cache[blk, pos % block_size, 0].float()          # [n, n_kv, D]
        v = cache[blk, pos % block_size, 1].float()
        k = k.repeat_interleave(group, dim=1)
        v = v.repeat_interleave(group, dim=1)
        s = torch.einsum("hd,nhd->hn", q[t].float(), k) * scale
        p = torch.softmax(s, dim=-1)
        out[t] = torch.einsum("hn,nhd->hd", p, v).to(q.dtype)
    return out
