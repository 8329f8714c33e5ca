"""Cut a whole v2 weight dictionary for one tensor-parallel rank.

Reference: every ``*Parameter`` class under ``model_implementations/common_parameters`` calls
``inference_model.transform_*_param`` which in turn calls the matching ``shard_*_param`` helper
(``inference_transformer_base.py:199-560``).  Here the fused per-layer dictionary (see
``inference_transformer_base.DSTransformerModelBase``) is mapped key by key.
"""
from __future__ import annotations

from typing import Dict

from .attn_out import shard_attn_out_param
from .mlp import shard_mlp_1_param, shard_mlp_2_param
from .qkv import shard_qkv_param
from .unembed import shard_unembed_param


def _is_gated(act) -> bool:
    return getattr(act, "name", str(act)).upper() in ("GEGLU", "REGLU", "SIGLU")


def shard_layer(lw, spec, rank: int, size: int):
    out = type(lw)()
    hq, hkv, d = spec.n_heads, spec.n_heads_kv, spec.head_size
    kv = None if hkv == hq else hkv
    gated = _is_gated(spec.mlp_activation)
    for k, v in lw.items():
        if v is None:
            continue
        if k in ("wqkv", "bqkv"):
            v = shard_qkv_param(v, rank, size, d, hq, kv)
        elif k in ("wo", "bo"):
            v = shard_attn_out_param(v, rank, size, d, hq, kv)
        elif k in ("w_in", "b_in", "shared_w_in"):
            v = shard_mlp_1_param(v, rank, size, gated=gated)
        elif k in ("w_out", "b_out", "shared_w_out"):
            v = shard_mlp_2_param(v, rank, size)
        elif k in ("moe_w1", "moe_b1"):
            v = shard_mlp_1_param(v, rank, size, gated=gated, is_moe=True)
        elif k in ("moe_w2", "moe_b2"):
            v = shard_mlp_2_param(v, rank, size, is_moe=True)
        # norms, router gate, shared-expert gate: replicated
        if v is not None:
            out[k] = v
    return out


def shard_weights(weights: Dict, spec, rank: int, size: int) -> Dict:
    """Local view of ``weights`` for ``rank`` of ``size``.  Embeddings, positions and norms stay replicated (the
    residual stream is replicated between layers); the unembedding is cut along the vocabulary."""
    if size == 1:
        return weights
    local = {k: v for k, v in weights.items() if k not in ("layers", "unembed", "unembed_b", "_transformed")}
    local["unembed"] = shard_unembed_param(weights["unembed"], rank, size)
    if weights.get("unembed_b") is not None:
        local["unembed_b"] = shard_unembed_param(weights["unembed_b"], rank, size)
    local["layers"] = [shard_layer(lw, spec, rank, size) for lw in weights["layers"]]
    return local
