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
    "sxb_attn_fwd_atoms": "ppp" "pii" "piii" "iiii" "ll" "f" "p",
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


def atoms_supported(q: torch.Tensor, cache: torch.Tensor) -> bool:
    """Whether the tensor-core atom kernel takes this query / cache layout."""
    bs = cache.shape[1]
    return (q.is_cuda and q.dtype == torch.bfloat16 and cache.dtype == torch.bfloat16 and q.shape[-1] in (64, 128)
            and q.stride(2) == 1 and q.stride(1) == q.shape[-1] and q.stride(0) % 8 == 0 and cache.is_contiguous()
            and bs % 8 == 0 and (64 % bs == 0 if bs < 64 else bs % 64 == 0))


def paged_attention_atoms(q: torch.Tensor, cache: torch.Tensor, atoms: torch.Tensor, block_tables: torch.Tensor,
                          scale: float = None, n_packed: int = 0) -> torch.Tensor:
    """The same attention as :func:`paged_attention`, scheduled by attention atoms on the tcgen05 flash kernel: one CTA
    per (atom, q head) runs 128 queries of one sequence against that sequence's pages, K/V tiles fetched by TMA through
    the block table (``csrc/attention.cu``: ``sxb_attn_fwd_atoms``).  Prefill re-reads K/V once per 128 queries instead
    of once per query.  ``n_packed``: the last that many atoms have at most ``128 // (n_q // n_kv)`` queries and run
    with the GQA group packed into the tile (K/V read once per KV head: decode)."""
    T, n_q, D = q.shape
    n_kv = cache.shape[3]
    scale = 1.0 / math.sqrt(D) if scale is None else scale
    out = torch.empty(T, n_q, D, dtype=q.dtype, device=q.device)
    _lib.call("sxb_attn_fwd_atoms", q.data_ptr(), cache.data_ptr(), out.data_ptr(), atoms.data_ptr(), atoms.shape[0],
              int(n_packed), block_tables.data_ptr(), block_tables.shape[1], cache.shape[0], cache.shape[1], T, n_q, n_kv, D,
              q.stride(0), out.stride(0), float(scale), _lib.stream_ptr())
    return out


def ragged_attention(q: torch.Tensor, cache: torch.Tensor, batch, scale: float = None) -> torch.Tensor:
    """Attention of a finalized ragged batch: atoms on the tensor cores where the layout allows, else per token."""
    atoms = batch.atoms() if hasattr(batch, "atoms") else None
    if atoms is not None and atoms.shape[0] > 0 and atoms_supported(q, cache):
        group = q.shape[1] // cache.shape[3]
        n_packed = 0
        if group > 1 and 128 % group == 0 and hasattr(batch, "atoms_longer_than"):
            n_packed = atoms.shape[0] - batch.atoms_longer_than(128 // group)
        return paged_attention_atoms(q, cache, atoms, batch.block_tables(), scale, n_packed)
    return paged_attention(q, cache, batch.token_to_seq(), batch.token_positions(), batch.block_tables(), scale)


# ================================================================================================ class interface
# The reference exposes every ragged kernel as a small callable object constructed once per model with the static
# shape information (``inference/v2/kernels/ragged_ops/*``); these are those objects over the functions above.  All of
# them take the :class:`RaggedBatchWrapper` where the reference does.

def get_q_block_size(head_size: int) -> int:
    """Query rows per attention atom (reference ``ragged_ops/blocked_flash``: larger tiles for smaller heads)."""
    return 128 if head_size <= 64 else 64


def get_kv_block_size(head_size: int) -> int:
    return 128 if head_size <= 128 else 64


class RaggedEmbeddingKernel:
    """``ragged_ops/embed``: token (+ learned position) embedding of a ragged batch."""
    supported_dtypes = (torch.float16, torch.bfloat16, torch.float32)

    def __init__(self, embed_dtype: torch.dtype, token_dtype: torch.dtype, embed_dim: int):
        if embed_dtype not in self.supported_dtypes:
            raise ValueError(f"Unsupported embedding data type: {embed_dtype}")
        if token_dtype not in (torch.int32, torch.int64):
            raise ValueError(f"Unsupported token data type: {token_dtype}")
        if embed_dim % (16 // torch.tensor([], dtype=embed_dtype).element_size()) != 0:
            raise ValueError(f"Embedding dimension {embed_dim} must be a multiple of 16 bytes")
        self.embed_dim = embed_dim

    def __call__(self, embedded_tokens: torch.Tensor, ragged_wrapper, embedding_weight: torch.Tensor,
                 position_embed_weight: torch.Tensor = None, position_embed_offset: int = 0) -> torch.Tensor:
        n = ragged_wrapper.current_tokens
        out = ragged_embed(ragged_wrapper.input_ids()[:n], embedding_weight,
                           ragged_wrapper.token_positions()[:n] if position_embed_weight is not None else None,
                           position_embed_weight, position_embed_offset)
        embedded_tokens[:n].copy_(out)
        return embedded_tokens


class RaggedLogitsGather:
    """``ragged_ops/logits_gather``: hidden state of the last token of every sequence."""

    def __init__(self, model_dim: int, fp_dtype: torch.dtype):
        self.model_dim = model_dim

    def __call__(self, final_token_activations: torch.Tensor, all_activations: torch.Tensor, ragged_wrapper):
        n = ragged_wrapper.current_sequences
        final_token_activations[:n].copy_(logits_gather(all_activations, ragged_wrapper.last_token_index()[:n]))
        return final_token_activations


class BlockedRotaryEmbeddings:
    """``ragged_ops/linear_blocked_kv_rotary``: rotate q/k in place (frequencies from ``theta_base``) and append k/v to
    the paged cache."""

    def __init__(self, head_size: int, n_q_heads: int, n_kv_heads: int, dtype: torch.dtype, rotary_dim: int,
                 theta_base: float):
        if head_size not in (64, 80, 96, 128):
            raise ValueError(f"Unsupported head size: {head_size}")
        if n_q_heads % n_kv_heads:
            raise ValueError("query heads must be a multiple of key/value heads")
        self.head_size, self.n_q, self.n_kv = head_size, n_q_heads, n_kv_heads
        self.rotary_dim, self.theta_base = rotary_dim, theta_base

    def __call__(self, kv_cache: torch.Tensor, qkv: torch.Tensor, ragged_batch) -> None:
        n = ragged_batch.current_tokens
        kv_rotary_blocked(qkv[:n], kv_cache, ragged_batch.token_to_seq()[:n], ragged_batch.token_positions()[:n],
                          ragged_batch.block_tables(), self.n_q, self.n_kv, self.head_size, self.rotary_dim,
                          self.theta_base)


class BlockedTrainedRotaryEmbeddings(BlockedRotaryEmbeddings):
    """Same with an explicit inverse-frequency table: accepted when it equals ``theta ** (-2i/d)`` (the kernels
    regenerate frequencies from theta), otherwise refused loudly."""

    def __init__(self, head_size: int, n_q_heads: int, n_kv_heads: int, dtype: torch.dtype):
        super().__init__(head_size, n_q_heads, n_kv_heads, dtype, head_size, 10000.0)

    def __call__(self, kv_cache, qkv, ragged_batch, inverse_freqs: torch.Tensor) -> None:
        d = inverse_freqs.numel() * 2
        theta = float(inverse_freqs[1].item() ** (-d / 2.0)) if inverse_freqs.numel() > 1 else self.theta_base
        expect = theta ** (-torch.arange(0, d, 2, dtype=torch.float32, device=inverse_freqs.device) / d)
        if not torch.allclose(expect, inverse_freqs.float(), rtol=1e-3, atol=1e-6):
            raise ValueError("inverse frequency table is not a plain theta ** (-2i/d) schedule")
        self.rotary_dim, self.theta_base = d, theta
        super().__call__(kv_cache, qkv, ragged_batch)


class LinearBlockedKVCopy(BlockedRotaryEmbeddings):
    """No positional rotation (learned / no position embeddings): only the cache append."""

    def __init__(self, head_size: int, n_q_heads: int, n_kv_heads: int, dtype: torch.dtype):
        super().__init__(head_size, n_q_heads, n_kv_heads, dtype, 0, 10000.0)


class BlockedFlashAttn:
    """``ragged_ops/blocked_flash``: causal attention of the batch's new tokens over their sequences' paged KV."""

    def __init__(self, head_size: int, dtype: torch.dtype):
        if head_size not in (64, 128):
            raise ValueError(f"Unsupported head size: {head_size} (64 and 128 run on the sm_100a kernel)")
        self.head_size = head_size

    def __call__(self, out: torch.Tensor, q: torch.Tensor, kv_cache: torch.Tensor, atoms_or_batch,
                 softmax_scale: float) -> torch.Tensor:
        b = atoms_or_batch
        n = b.current_tokens
        T = q.shape[0]
        q3 = q[:n].view(n, -1, self.head_size)
        res = ragged_attention(q3, kv_cache, b, softmax_scale)
        out[:n].copy_(res.reshape(n, -1))
        return out


class RaggedTopKGating:
    """``ragged_ops/top_k_gating``: softmax top-k with per-expert counts / slots."""

    def __init__(self, logit_dtype: torch.dtype = torch.float32):
        self.logit_dtype = logit_dtype

    def __call__(self, logits: torch.Tensor, top_k: int, renormalize: bool = True):
        from ....ops import moe_ops
        return moe_ops.top_k_gating(logits, top_k, renormalize=renormalize)


class MoEScatter:
    """``ragged_ops/moe_scatter``: tokens -> expert-contiguous rows."""

    def __init__(self, dtype: torch.dtype, channels: int):
        self.channels = channels

    def __call__(self, activations, assignments, offsets, expert_counts):
        from ....ops import moe_ops
        return moe_ops.moe_scatter(activations, assignments, offsets, expert_counts)


class MoEGather:
    """``ragged_ops/moe_gather``: score-weighted sum of each token's expert outputs."""

    def __init__(self, dtype: torch.dtype, channels: int, normalize_scores: bool = False):
        self.normalize_scores = normalize_scores

    def __call__(self, moe_output, scores, mapped_slots):
        from ....ops import moe_ops
        if self.normalize_scores:
            scores = scores / scores.sum(dim=-1, keepdim=True)
        return moe_ops.moe_gather(moe_output, scores, mapped_slots)
