"""Position-bias builders that must be tensor-parallel aware (reference: ``module_inject/auto_tp_model_utils.py``).

ALiBi slopes are a function of the *global* head index.  After AutoTP a rank only holds ``heads / tp`` heads, so the
client's builder (which enumerates ``num_heads`` slopes from zero) would hand every rank the slopes of rank 0; these
replacements build all slopes and return the rank's slice."""
from __future__ import annotations

import math

import torch

from .. import comm as dist
from .tp_shard import get_shard_size, get_shard_size_list


def _alibi_slopes(num_heads: int, device=None) -> torch.Tensor:
    closest = 2 ** math.floor(math.log2(num_heads))
    base = 2.0 ** (-(2.0 ** -(math.log2(closest) - 3)))
    slopes = torch.pow(torch.tensor(base, device=device), torch.arange(1, closest + 1, device=device))
    if closest != num_heads:
        extra_base = 2.0 ** (-(2.0 ** -(math.log2(2 * closest) - 3)))
        n_extra = min(closest, num_heads - closest)
        extra = torch.pow(torch.tensor(extra_base, device=device), torch.arange(1, 1 + 2 * n_extra, 2, device=device))
        slopes = torch.cat([slopes, extra])
    return slopes.float()


def _tp_rank_world():
    if dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def _rank_head_slice(num_heads: int, rank=None, world=None) -> slice:
    r, w = _tp_rank_world()
    rank, world = (r if rank is None else rank), (w if world is None else world)
    sizes = get_shard_size_list(num_heads, world)
    lo = sum(sizes[:rank])
    return slice(lo, lo + sizes[rank])


def build_bloom_alibi_tensor(attention_mask: torch.Tensor, num_heads: int, dtype: torch.dtype, rank=None,
                             world=None) -> torch.Tensor:
    """BLOOM ALiBi ``[batch * local_heads, 1, seq]``: slope_h * (position of each unmasked token).  ``num_heads`` is
    the GLOBAL head count; the result covers this rank's heads only."""
    B, S = attention_mask.shape
    slopes = _alibi_slopes(num_heads, attention_mask.device)
    pos = ((attention_mask.cumsum(dim=-1) - 1) * attention_mask)[:, None, :].float()
    alibi = slopes[None, :, None] * pos                                   # [B, H, S]
    hs = _rank_head_slice(num_heads, rank, world)
    alibi = alibi[:, hs]
    return alibi.reshape(B * alibi.shape[1], 1, S).to(dtype)


def get_alibi_mask(self, tensor: torch.Tensor, seq_length_with_past: int, rank=None, world=None) -> torch.Tensor:
    """Baichuan: additive ``[local_heads, S, S]`` mask = causal -inf triangle + slope_h * key position."""
    n_head = getattr(self, "n_head", None) or getattr(self, "num_heads")
    slopes = _alibi_slopes(n_head, tensor.device)
    pos = torch.arange(seq_length_with_past, device=tensor.device).float()
    bias = slopes[:, None, None] * pos[None, None, :]
    tri = torch.full((seq_length_with_past, seq_length_with_past), float("-inf"), device=tensor.device).triu(1)
    mask = (bias + tri[None]).to(tensor.dtype)
    return mask[_rank_head_slice(n_head, rank, world)]


def build_mpt_alibi_tensor(self, num_heads: int, sequence_length: int, alibi_bias_max: int = 8, device=None,
                           rank=None, world=None) -> torch.Tensor:
    """MPT ALiBi ``[local_heads, 1, seq]``: slope_h * (key position - (seq - 1)) with MPT's slope schedule."""
    pos = torch.arange(1 - sequence_length, 1, dtype=torch.int32, device=device).view(1, 1, sequence_length).float()
    padded = 2 ** math.ceil(math.log2(num_heads))
    base = torch.arange(1, padded + 1, device=device).float() * (alibi_bias_max / padded)
    slopes = 1.0 / torch.pow(2, base)
    if padded != num_heads:
        slopes = torch.cat([slopes[1::2], slopes[::2]])[:num_heads]
    alibi = pos * slopes.view(num_heads, 1, 1)
    return alibi[_rank_head_slice(num_heads, rank, world)]


def build_mpt_atten_bias_tensor(self, device, dtype, attention_mask=None, prefix_mask=None, sequence_id=None,
                                rank=None, world=None):
    """MPT ``_attn_bias`` replacement: let the client build the full-head bias, then keep this rank's heads."""
    attn_bias, attention_mask = self._attn_bias_orig(device=device, dtype=dtype, attention_mask=attention_mask,
                                                     prefix_mask=prefix_mask, sequence_id=sequence_id)
    if attn_bias is not None and attn_bias.dim() == 4 and attn_bias.shape[1] > 1:
        attn_bias = attn_bias[:, _rank_head_slice(attn_bias.shape[1], rank, world)]
    return attn_bias, attention_mask
