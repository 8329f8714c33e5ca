"""Process-group registry: data / tensor(model) / expert / expert-data / sequence /
sequence-data / hpZ secondary / qgZ local-global all-to-all groups.

Behavioural parity with the reference's deepspeed/utils/groups.py (EP creation E+D :240-312 and
E+M+D :315-438, hpZ groups :692-719, qgZ local/global groups :515-541, mesh fallbacks :544-684),
re-expressed around one helper, :func:`_grid_groups`, that enumerates the cosets of a cartesian
rank grid; every rank calls ``new_group`` for every coset in the same order (a torch.distributed
requirement) and keeps the one it belongs to.
"""
from __future__ import annotations

from typing import Dict, List, Optional

from ..comm import comm as dist
from .logging import log_dist

mpu = None                       # external Megatron-style model-parallel unit, if any
mesh_device = None
expert_tensor_parallel_world_size = 1

_WORLD_GROUP = None
_EXPERT_PARALLEL_GROUP: Dict[str, object] = {}
_EXPERT_DATA_PARALLEL_GROUP: Dict[str, object] = {}
_EXPERT_PARALLEL_RANKS: Dict[str, List[int]] = {}
_EXPERT_DATA_PARALLEL_RANKS: Dict[str, List[int]] = {}
_MAX_EP_SIZE_NAME = None
_ZERO_PARAM_INTRA_PARALLEL_GROUP = None
_ZERO_PARAM_INTRA_PARALLEL_RANKS = None
_ALL_TO_ALL_GROUP: Dict[str, object] = {}
_DATA_PARALLEL_GROUP = None
_TENSOR_PARALLEL_GROUP = None
_SEQUENCE_PARALLEL_GROUP = None
_SEQUENCE_DATA_PARALLEL_GROUP = None
_TP_SIZE = 1
_SP_SIZE = 1


def _ensure_divisibility(num, den):
    assert num % den == 0, f"{num} is not divisible by {den}"


def _grid_groups(world: int, inner: int, contiguous: bool):
    """Rank lists of size ``inner``: contiguous blocks, or strided cosets (stride = world // inner)."""
    if contiguous:
        return [list(range(i, i + inner)) for i in range(0, world, inner)]
    stride = world // inner
    return [list(range(i, world, stride)) for i in range(stride)]


def _make(rank_lists, my_rank):
    mine, mine_ranks = None, None
    for ranks in rank_lists:
        g = dist.new_group(ranks)
        if my_rank in ranks:
            mine, mine_ranks = g, ranks
    return mine, mine_ranks


def initialize(ep_size=1, mpu_=None):
    """Deprecated reference API kept for scripts: creates the EP groups."""
    global mpu
    if mpu_ is not None:
        mpu = mpu_
    _create_expert_and_data_parallel(ep_size)


# -------------------------------------------------------------------------- tensor parallel
def _init_tp_mesh_device(tensor_model_parallel_size=1, data_parallel_size=None):
    """TP groups = contiguous blocks; DP groups = strided cosets (AutoTP training)."""
    global _DATA_PARALLEL_GROUP, _TENSOR_PARALLEL_GROUP, _TP_SIZE
    world, rank = dist.get_world_size(), dist.get_rank()
    _ensure_divisibility(world, tensor_model_parallel_size)
    _TP_SIZE = tensor_model_parallel_size
    if not dist.is_initialized():
        return None, None
    _TENSOR_PARALLEL_GROUP, _ = _make(_grid_groups(world, tensor_model_parallel_size, True), rank)
    _DATA_PARALLEL_GROUP, _ = _make(_grid_groups(world, world // tensor_model_parallel_size, False), rank)
    return _DATA_PARALLEL_GROUP, _TENSOR_PARALLEL_GROUP


def get_tensor_model_parallel_group():
    if mpu is not None:
        return _mpu_call("get_tensor_model_parallel_group", "get_model_parallel_group")
    return _TENSOR_PARALLEL_GROUP


def get_model_parallel_group():
    return get_tensor_model_parallel_group()


def get_tensor_model_parallel_world_size():
    if mpu is not None:
        return _mpu_call("get_tensor_model_parallel_world_size", "get_model_parallel_world_size")
    return _TP_SIZE


def get_model_parallel_world_size():
    return get_tensor_model_parallel_world_size()


def get_tensor_model_parallel_rank():
    if mpu is not None:
        return _mpu_call("get_tensor_model_parallel_rank", "get_model_parallel_rank")
    if globals().get("_TP_RANK_OVERRIDE") is not None:
        return globals()["_TP_RANK_OVERRIDE"]
    g = _TENSOR_PARALLEL_GROUP
    return dist.get_rank(g) if g is not None else 0


def get_model_parallel_rank():
    return get_tensor_model_parallel_rank()


def _mpu_call(*names):
    for n in names:
        if hasattr(mpu, n):
            return getattr(mpu, n)()
    raise AttributeError(f"mpu has none of {names}")


# ---------------------------------------------------------------------------- data parallel
def _clone_world_group():
    global _WORLD_GROUP
    if _WORLD_GROUP is None and dist.is_initialized():
        _WORLD_GROUP = dist.new_group(ranks=list(range(dist.get_world_size())))
    return _WORLD_GROUP


def _get_data_parallel_group():
    if mesh_device is not None:
        return mesh_device.get_group(mesh_dim="data_parallel")
    if mpu is not None:
        return mpu.get_data_parallel_group()
    if _DATA_PARALLEL_GROUP is not None:
        return _DATA_PARALLEL_GROUP
    return _clone_world_group()


def _get_data_parallel_world_size():
    if mesh_device is not None:
        return dist.get_world_size(mesh_device.get_group(mesh_dim="data_parallel"))
    if mpu is not None:
        return mpu.get_data_parallel_world_size()
    if _DATA_PARALLEL_GROUP is not None:
        return dist.get_world_size(_DATA_PARALLEL_GROUP)
    return dist.get_world_size()


def _get_data_parallel_rank():
    g = _get_data_parallel_group()
    return dist.get_rank(g) if g is not None else 0


def _get_broadcast_src_rank():
    return dist.get_global_rank(_get_sequence_data_parallel_group(), 0)


# ------------------------------------------------------------------------ sequence parallel
def _create_sequence_parallel_groups(sp_size: int):
    """Consecutive ranks form one SP group; SPxDP group is the world (reference:
    runtime/sequence_parallel/parallel_state_sp.py:21-62)."""
    global _SEQUENCE_PARALLEL_GROUP, _SEQUENCE_DATA_PARALLEL_GROUP, _SP_SIZE, _DATA_PARALLEL_GROUP
    world, rank = dist.get_world_size(), dist.get_rank()
    _ensure_divisibility(world, sp_size)
    _SP_SIZE = sp_size
    if not dist.is_initialized():
        return
    _SEQUENCE_PARALLEL_GROUP, _ = _make(_grid_groups(world, sp_size, True), rank)
    _DATA_PARALLEL_GROUP, _ = _make(_grid_groups(world, world // sp_size, False), rank)
    _SEQUENCE_DATA_PARALLEL_GROUP = _clone_world_group()


def _get_sequence_parallel_group():
    if mesh_device is not None:
        return mesh_device.get_group(mesh_dim="sequence_parallel")
    if mpu is not None and hasattr(mpu, "get_sequence_parallel_group"):
        return mpu.get_sequence_parallel_group()
    return _SEQUENCE_PARALLEL_GROUP


def _get_sequence_parallel_world_size():
    if mesh_device is not None:
        return dist.get_world_size(mesh_device.get_group(mesh_dim="sequence_parallel"))
    if mpu is not None and hasattr(mpu, "get_sequence_parallel_world_size"):
        return mpu.get_sequence_parallel_world_size()
    return _SP_SIZE


def _get_sequence_parallel_rank():
    if mpu is not None and hasattr(mpu, "get_sequence_parallel_rank"):
        return mpu.get_sequence_parallel_rank()
    g = _get_sequence_parallel_group()
    return dist.get_rank(g) if g is not None else 0


def _get_sequence_data_parallel_group():
    if mesh_device is not None:
        return _clone_world_group()
    if mpu is not None and hasattr(mpu, "get_sequence_data_parallel_group"):
        return mpu.get_sequence_data_parallel_group()
    if _SEQUENCE_DATA_PARALLEL_GROUP is not None:
        return _SEQUENCE_DATA_PARALLEL_GROUP
    return _get_data_parallel_group()


def _get_sequence_data_parallel_world_size():
    if mpu is not None and hasattr(mpu, "get_sequence_data_parallel_world_size"):
        return mpu.get_sequence_data_parallel_world_size()
    g = _get_sequence_data_parallel_group()
    return dist.get_world_size(g) if g is not None else 1


def _get_sequence_data_parallel_rank():
    g = _get_sequence_data_parallel_group()
    return dist.get_rank(g) if g is not None else 0


# --------------------------------------------------------------------------- expert parallel
def _expert_rank_lists(world: int, ep_size: int, use_data_before_expert_parallel: bool = False):
    """(expert-parallel rank lists, expert-data-parallel rank lists) for a pure DP world.

    E+D: EP groups are consecutive ranks, expert-DP groups stride by ep_size.  With
    ``use_data_before_expert_parallel`` (D+E) the roles of the two axes are swapped.
    """
    _ensure_divisibility(world, ep_size)
    if not use_data_before_expert_parallel:
        ep = _grid_groups(world, ep_size, True)
        edp = _grid_groups(world, world // ep_size, False)
    else:
        ep = _grid_groups(world, ep_size, False)
        edp = _grid_groups(world, world // ep_size, True)
    return ep, edp


def _create_expert_and_data_parallel(expert_parallel_size_: int, use_data_before_expert_parallel_: bool = False):
    global _MAX_EP_SIZE_NAME
    assert dist.is_initialized()
    world, rank = dist.get_world_size(), dist.get_rank()
    name = f"ep_size_{expert_parallel_size_}"
    if name in _EXPERT_PARALLEL_GROUP:
        return
    log_dist(f"Creating expert and data parallel groups with size {expert_parallel_size_}", ranks=[0])
    ep, edp = _expert_rank_lists(world, expert_parallel_size_, use_data_before_expert_parallel_)
    _EXPERT_DATA_PARALLEL_GROUP[name], _EXPERT_DATA_PARALLEL_RANKS[name] = _make(edp, rank)
    _EXPERT_PARALLEL_GROUP[name], _EXPERT_PARALLEL_RANKS[name] = _make(ep, rank)
    if _MAX_EP_SIZE_NAME is None or expert_parallel_size_ > int(_MAX_EP_SIZE_NAME.split("_")[-1]):
        _MAX_EP_SIZE_NAME = name


def _get_expert_parallel_ranks(world_size, tensor_parallel_size_, expert_parallel_size_, pipeline_parallel_size_=1,
                               use_data_before_expert_parallel_=False):
    """Rank lists for E + M + D (+P) layouts (reference: groups.py:315-385).

    World is viewed as [pp, dp, tp] (tp fastest).  For every (pp, tp) coordinate the dp ranks are
    cut into EP groups of ``expert_parallel_size_`` and the complementary expert-DP groups.
    """
    _ensure_divisibility(world_size, tensor_parallel_size_ * pipeline_parallel_size_)
    dp_world = world_size // (tensor_parallel_size_ * pipeline_parallel_size_)
    _ensure_divisibility(dp_world, expert_parallel_size_)
    ep_groups, edp_groups = [], []
    pp_stride = world_size // pipeline_parallel_size_
    for pp in range(pipeline_parallel_size_):
        for tp in range(tensor_parallel_size_):
            dp_ranks = [pp * pp_stride + d * tensor_parallel_size_ + tp for d in range(dp_world)]
            ep, edp = _expert_rank_lists(dp_world, expert_parallel_size_, use_data_before_expert_parallel_)
            ep_groups += [[dp_ranks[i] for i in g] for g in ep]
            edp_groups += [[dp_ranks[i] for i in g] for g in edp]
    return ep_groups, edp_groups


def _create_expert_data_and_model_parallel(expert_parallel_size_, mpu_, use_data_before_expert_parallel_=False):
    global expert_tensor_parallel_world_size, _MAX_EP_SIZE_NAME
    assert dist.is_initialized()
    tp = _mpu_size(mpu_, "get_tensor_model_parallel_world_size", "get_model_parallel_world_size")
    pp = mpu_.get_pipeline_model_parallel_world_size() if hasattr(mpu_, "get_pipeline_model_parallel_world_size") else 1
    world, rank = dist.get_world_size(), dist.get_rank()
    name = f"ep_size_{expert_parallel_size_}"
    if name in _EXPERT_PARALLEL_GROUP:
        return
    expert_tensor_parallel_world_size = tp
    ep, edp = _get_expert_parallel_ranks(world, tp, expert_parallel_size_, pp, use_data_before_expert_parallel_)
    _EXPERT_PARALLEL_GROUP[name], _EXPERT_PARALLEL_RANKS[name] = _make(ep, rank)
    _EXPERT_DATA_PARALLEL_GROUP[name], _EXPERT_DATA_PARALLEL_RANKS[name] = _make(edp, rank)
    if _MAX_EP_SIZE_NAME is None or expert_parallel_size_ > int(_MAX_EP_SIZE_NAME.split("_")[-1]):
        _MAX_EP_SIZE_NAME = name


def _mpu_size(m, *names):
    for n in names:
        if hasattr(m, n):
            return getattr(m, n)()
    return 1


def _get_max_expert_size():
    assert _EXPERT_PARALLEL_GROUP, "expert parallel groups are not initialized"
    return max(int(k.split("_")[-1]) for k in _EXPERT_PARALLEL_GROUP)


def _get_max_expert_size_name():
    return f"ep_size_{_get_max_expert_size()}"


def _get_max_expert_parallel_group():
    return _get_expert_parallel_group(_get_max_expert_size_name())


def _get_expert_parallel_group(group_name):
    assert group_name in _EXPERT_PARALLEL_GROUP, f"expert parallel group {group_name} is not initialized"
    return _EXPERT_PARALLEL_GROUP[group_name]


def _get_expert_parallel_group_dict():
    return _EXPERT_PARALLEL_GROUP


def _get_expert_data_parallel_group(group_name):
    assert group_name in _EXPERT_DATA_PARALLEL_GROUP, f"expert data parallel group {group_name} is not initialized"
    return _EXPERT_DATA_PARALLEL_GROUP[group_name]


def _get_expert_data_parallel_group_dict():
    return _EXPERT_DATA_PARALLEL_GROUP


def _get_expert_parallel_world_size(group_name):
    return dist.get_world_size(group=_get_expert_parallel_group(group_name))


def _get_expert_data_parallel_world_size(group_name):
    return dist.get_world_size(group=_get_expert_data_parallel_group(group_name))


def _get_expert_parallel_rank(group_name):
    return dist.get_rank(group=_get_expert_parallel_group(group_name))


def _get_expert_parallel_src_rank(group_name):
    g = _get_expert_parallel_group(group_name)
    return dist.get_global_rank(g, 0)


def _get_expert_data_parallel_rank(group_name):
    return dist.get_rank(group=_get_expert_data_parallel_group(group_name))


# ----------------------------------------------------------------------------- ZeRO++ groups
def _create_zero_param_parallel_group(group_size: int):
    """hpZ secondary partition groups: consecutive blocks of ``group_size`` ranks."""
    global _ZERO_PARAM_INTRA_PARALLEL_GROUP, _ZERO_PARAM_INTRA_PARALLEL_RANKS
    assert dist.is_initialized()
    if _ZERO_PARAM_INTRA_PARALLEL_GROUP is not None:
        return
    world, rank = dist.get_world_size(), dist.get_rank()
    _ensure_divisibility(world, group_size)
    _ZERO_PARAM_INTRA_PARALLEL_GROUP, _ZERO_PARAM_INTRA_PARALLEL_RANKS = _make(
        _grid_groups(world, group_size, True), rank)


def _zero_param_parallel_is_initialized():
    return _ZERO_PARAM_INTRA_PARALLEL_GROUP is not None


def _get_zero_param_intra_parallel_group():
    return _ZERO_PARAM_INTRA_PARALLEL_GROUP


def _get_zero_param_intra_parallel_rank_in_mygroup():
    return dist.get_rank(group=_ZERO_PARAM_INTRA_PARALLEL_GROUP)


def _get_zero_param_intra_parallel_group_world_size():
    return dist.get_world_size(group=_ZERO_PARAM_INTRA_PARALLEL_GROUP)


def _get_zero_param_intra_parallel_group_ranks():
    return _ZERO_PARAM_INTRA_PARALLEL_RANKS


def _get_local_all_to_all_group(devices_per_node: Optional[int] = None):
    """qgZ: ``local_i`` = ranks of node i, ``global_i`` = i-th GPU of every node."""
    assert dist.is_initialized()
    if _ALL_TO_ALL_GROUP:
        return _ALL_TO_ALL_GROUP
    import torch
    world = dist.get_world_size()
    per = devices_per_node or max(1, min(world, torch.cuda.device_count() if torch.cuda.is_available() else world))
    nodes = max(1, world // per)
    if nodes > 1:
        for i in range(per):
            _ALL_TO_ALL_GROUP[f"global_{i}"] = dist.new_group(ranks=[i + per * j for j in range(nodes)])
    for i in range(nodes):
        _ALL_TO_ALL_GROUP[f"local_{i}"] = dist.new_group(ranks=[j + per * i for j in range(per)])
    _ALL_TO_ALL_GROUP["devices_per_node"] = per      # read back by runtime/comm/coalesced_collectives.py
    return _ALL_TO_ALL_GROUP


def reset_for_tests():
    """Forget every cached group (used between spawned test worlds)."""
    global _WORLD_GROUP, _MAX_EP_SIZE_NAME, _ZERO_PARAM_INTRA_PARALLEL_GROUP, _ZERO_PARAM_INTRA_PARALLEL_RANKS
    global _DATA_PARALLEL_GROUP, _TENSOR_PARALLEL_GROUP, _SEQUENCE_PARALLEL_GROUP, _SEQUENCE_DATA_PARALLEL_GROUP
    global _TP_SIZE, _SP_SIZE, mpu, mesh_device
    _WORLD_GROUP = _MAX_EP_SIZE_NAME = None
    _ZERO_PARAM_INTRA_PARALLEL_GROUP = _ZERO_PARAM_INTRA_PARALLEL_RANKS = None
    _DATA_PARALLEL_GROUP = _TENSOR_PARALLEL_GROUP = _SEQUENCE_PARALLEL_GROUP = _SEQUENCE_DATA_PARALLEL_GROUP = None
    _TP_SIZE = _SP_SIZE = 1
    mpu = mesh_device = None
    for d in (_EXPERT_PARALLEL_GROUP, _EXPERT_DATA_PARALLEL_GROUP, _EXPERT_PARALLEL_RANKS,
              _EXPERT_DATA_PARALLEL_RANKS, _ALL_TO_ALL_GROUP):
        d.clear()


# ---- explicit overrides used by inference / hybrid code that builds its own tensor-parallel layout
_TP_RANK_OVERRIDE = None


def set_tensor_model_parallel_world_size(world_size):
    global _TP_SIZE
    _TP_SIZE = int(world_size)


def set_tensor_model_parallel_rank(rank):
    global _TP_RANK_OVERRIDE
    _TP_RANK_OVERRIDE = None if rank is None else int(rank)


def get_tensor_model_parallel_src_rank():
    """Global rank of the first member of this rank's tensor-parallel group (the broadcast source for its inputs)."""
    tp = max(1, get_tensor_model_parallel_world_size() or 1)
    return (dist.get_rank() // tp) * tp if dist.is_initialized() else 0


def get_data_parallel_rank():
    return _get_data_parallel_rank()


def get_data_parallel_group():
    """Data-parallel group of the TP mesh set up by :func:`_init_tp_mesh_device` (AutoTP training); unlike
    :func:`_get_data_parallel_group` this accessor does not fall back to the world group."""
    assert _DATA_PARALLEL_GROUP is not None, "data parallel group is not initialized"
    return _DATA_PARALLEL_GROUP


def get_data_parallel_world_size():
    return dist.get_world_size(group=get_data_parallel_group())
