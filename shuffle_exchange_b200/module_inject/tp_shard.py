"""Shard-size arithmetic for tensor parallelism with uneven head counts (reference: ``module_inject/tp_shard.py``):
sizes are multiples of the head dimension when kv-head information is known, so that no head is split."""
from __future__ import annotations

_NUM_KV_HEADS = None
_N_EMBD = None
_VOCAB = None
_TP_GRAIN = 1


def set_num_kv_heads(num):
    global _NUM_KV_HEADS
    _NUM_KV_HEADS = num


def set_num_attention_heads(num):
    global _NUM_ATT
    _NUM_ATT = num


def set_n_embd(num):
    global _N_EMBD
    _N_EMBD = num


def set_tp_grain_size(num):
    global _TP_GRAIN
    _TP_GRAIN = max(1, int(num or 1))


def set_vocab_size(num):
    global _VOCAB
    _VOCAB = num


def get_num_kv_heads():
    return _NUM_KV_HEADS


def get_num_attention_heads():
    return globals().get("_NUM_ATT")


def get_n_embd():
    return _N_EMBD


def get_shard_size(total_size, mp_size, name=None, rank=None):
    """Size of rank ``rank``'s shard (default: this rank)."""
    from .. import comm as dist
    if rank is None:
        rank = dist.get_rank() % mp_size if dist.is_initialized() else 0
    return get_shard_size_list(total_size, mp_size, name)[rank]


def get_shard_size_list(total_size, mp_size, name=None):
    """Per-rank sizes: whole kv heads per rank when head info is known and the tensor is a head-structured one
    (not lm_head / embed_out), else multiples of the TP grain; the first ranks get the remainder."""
    last = (name or "").split(".")[-1]
    if _NUM_KV_HEADS is not None and _N_EMBD and total_size % _NUM_KV_HEADS == 0 and "mlp" not in (name or "") \
            and last not in ("lm_head", "embed_out"):
        per_head = total_size // _NUM_KV_HEADS
        base, rem = divmod(_NUM_KV_HEADS, mp_size)
        return [(base + (1 if r < rem else 0)) * per_head for r in range(mp_size)]
    grain = _TP_GRAIN
    if total_size >= grain and total_size % grain == 0:
        units = total_size // grain
        base, rem = divmod(units, mp_size)
        return [(base + (1 if r < rem else 0)) * grain for r in range(mp_size)]
    base, rem = divmod(total_size, mp_size)
    return [base + (1 if r < rem else 0) for r in range(mp_size)]
