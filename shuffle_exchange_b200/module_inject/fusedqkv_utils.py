"""Tensor-parallel slicing of fused QKV (and fused gate|up) weights whose row order is architecture specific.

Reference: ``module_inject/fusedqkv_utils.py:1-205``.  A fused weight ``[rows, in]`` must be cut so that every rank
receives *its heads'* q, k and v rows and still sees them in the order the client attention module expects.  The row
orders in the wild:

=============  ==========================================  =============================
type           row order                                    examples
=============  ==========================================  =============================
``bloomtype``  ``[heads, (q|k|v), d]``                       BLOOM, GPT-NeoX, Falcon-7B MHA
``glmtype``    ``[(q|k|v), heads, d]``                       ChatGLM, GPT-2 ``c_attn``, MPT ``Wqkv``
``qwentype``   same as ``glmtype``                           Qwen-1 ``c_attn``
``bigcodetype`` ``[q (heads*d) | k (d) | v (d)]`` (MQA)      GPTBigCode: q cut, the single k/v head replicated
``codegentype`` ``[mp_blocks(4), (q|v|k), heads/4, d]``      CodeGen
``phi3type``   ``[q (hq*d) | k (hkv*d) | v (hkv*d)]`` (GQA)  Phi-3 ``qkv_proj`` and its ``gate_up_proj``
=============  ==========================================  =============================
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from ..utils.logging import warning_once
from .tp_shard import get_num_attention_heads, get_num_kv_heads, get_shard_size, get_shard_size_list

FUSED_QKV_TYPES = {
    "CodeGenBlock": "codegentype", "BloomBlock": "bloomtype", "GLMBlock": "glmtype", "MPTBlock": "glmtype",
    "MptBlock": "glmtype", "BaichuanLayer": "glmtype", "QWenBlock": "qwentype", "FalconDecoderLayer": "bloomtype",
    "GPTBigCodeBlock": "bigcodetype", "DecoderLayer": "glmtype", "Phi3DecoderLayer": "phi3type",
    "GPTNeoXLayer": "bloomtype", "GPT2Block": "glmtype",
}
_FUSED_NAMES = ("query_key_value", "attn.Wqkv", "self_attn.W_pack", "c_attn", "qkv_proj")


def require_tp_fused_qkvw(name: str, mp_size: int) -> bool:
    """Does the parameter / module path ``name`` denote a fused QKV projection that needs head-aware slicing?"""
    return mp_size > 1 and any(f in name for f in _FUSED_NAMES)


def split_by_qkvlist_and_refuse(qkv_list: Sequence[torch.Tensor], split_size, split_dim: int = 0,
                                cat_dim: int = 0) -> List[torch.Tensor]:
    """Cut each of q, k, v into per-rank pieces and re-fuse rank by rank: ``out[r] = cat(q_r, k_r, v_r)``.
    ``split_size`` is one size for all ranks or a per-rank list (uneven heads)."""
    pieces = [torch.split(t, split_size, dim=split_dim) for t in qkv_list]
    return [torch.cat([p[r] for p in pieces], dim=cat_dim) for r in range(len(pieces[0]))]


def _rank_rows(total: int, mp_size: int, rank: int) -> slice:
    sizes = get_shard_size_list(total, mp_size)
    lo = sum(sizes[:rank])
    return slice(lo, lo + sizes[rank])


def prepare_tp_fused_qkvw(module, src: torch.Tensor, mp_size: int, gpu_index: int,
                          fused_type: Optional[str] = None) -> torch.Tensor:
    """Rows of ``src`` (a fused QKV weight ``[rows, in]`` or bias ``[rows]``) that belong to rank ``gpu_index``."""
    if src is None or mp_size == 1:
        return src
    ftype = fused_type or FUSED_QKV_TYPES.get(type(module).__name__ if not isinstance(module, str) else module)
    if ftype is None:
        for cls_name, t in FUSED_QKV_TYPES.items():
            if cls_name in str(module):
                ftype = t
                break
    if ftype is None:
        warning_once(f"fused QKV layout of {type(module).__name__} unknown; assuming [(q|k|v), heads, d]")
        ftype = "glmtype"
    rows = src.shape[0]
    if ftype == "bloomtype":
        # heads outermost: a contiguous row range IS a set of whole heads with their q, k, v together
        heads = get_num_attention_heads() or mp_size
        per_head = rows // heads
        hs = _rank_rows(heads, mp_size, gpu_index)
        return src[hs.start * per_head: hs.stop * per_head].contiguous()
    if ftype in ("glmtype", "qwentype"):
        q, k, v = src.chunk(3, dim=0)
        n = q.shape[0]
        sizes = get_shard_size_list(n, mp_size, "qkv")
        return split_by_qkvlist_and_refuse([q, k, v], sizes)[gpu_index].contiguous()
    if ftype == "bigcodetype":
        heads = get_num_attention_heads()
        d = rows // (heads + 2)
        q, kv = src[: heads * d], src[heads * d:]
        return torch.cat([q[_rank_rows(heads, mp_size, gpu_index).start * d:
                            _rank_rows(heads, mp_size, gpu_index).stop * d], kv], dim=0).contiguous()
    if ftype == "codegentype":
        blocks = 4                           # CodeGen's checkpoint layout is [4 blocks][q|v|k][heads/4 * d]
        if mp_size > blocks or blocks % mp_size:
            raise ValueError("CodeGen fused QKV can be split over 1, 2 or 4 ranks")
        view = src.reshape(blocks, 3, rows // (3 * blocks), *src.shape[1:])
        per = blocks // mp_size
        return view[gpu_index * per:(gpu_index + 1) * per].reshape(-1, *src.shape[1:]).contiguous()
    if ftype == "phi3type":
        hq, hkv = get_num_attention_heads(), get_num_kv_heads() or get_num_attention_heads()
        d = rows // (hq + 2 * hkv)
        q, k, v = src.split([hq * d, hkv * d, hkv * d], dim=0)
        qs, ks = _rank_rows(hq, mp_size, gpu_index), _rank_rows(hkv, mp_size, gpu_index)
        return torch.cat([q[qs.start * d: qs.stop * d], k[ks.start * d: ks.stop * d], v[ks.start * d: ks.stop * d]],
                         dim=0).contiguous()
    raise ValueError(f"unknown fused QKV type {ftype}")


def shard_value_with_share_qk(weight: torch.Tensor, bias: Optional[torch.Tensor], rank: int, world_size: int,
                              shard_value: bool = True) -> torch.Tensor:
    """Layers that fuse a *shared* query/key projection with a value projection (``[qk | v]``): the q/k half is cut per
    rank, and so is the value half when ``shard_value`` (otherwise it is replicated)."""
    t = weight if bias is None else bias
    half = t.shape[0] // 2
    qk, v = t[:half], t[half:]
    qk_r = qk[_rank_rows(half, world_size, rank)]
    v_r = v[_rank_rows(half, world_size, rank)] if shard_value else v
    return torch.cat([qk_r, v_r], dim=0).contiguous()


def shard_chunk_mlp(weight: torch.Tensor, bias: Optional[torch.Tensor], rank: int, world_size: int) -> torch.Tensor:
    """Fused gate|up projection (Phi-3 ``gate_up_proj``): each half is cut for ``rank`` and the halves re-fused."""
    t = weight if bias is None else bias
    gate, up = t.chunk(2, dim=0)
    rows = _rank_rows(gate.shape[0], world_size, rank)
    return torch.cat([gate[rows], up[rows]], dim=0).contiguous()
