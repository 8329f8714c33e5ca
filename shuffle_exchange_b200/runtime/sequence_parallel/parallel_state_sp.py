"""Process-group state for Ulysses sequence parallelism.

Reference: ``deepspeed/runtime/sequence_parallel/parallel_state_sp.py:17-90`` — consecutive ranks form one SP group
(``world / sp`` groups), ranks with the same SP rank form one "sequence data parallel" group."""
from __future__ import annotations

import torch.distributed as dist

_SEQUENCE_PARALLEL_GROUP = None
_SEQUENCE_DATA_PARALLEL_GROUP = None


def initialize_sequence_parallel(sequence_parallel_size: int) -> None:
    global _SEQUENCE_PARALLEL_GROUP, _SEQUENCE_DATA_PARALLEL_GROUP
    assert dist.is_initialized()
    world, rank = dist.get_world_size(), dist.get_rank()
    if world < sequence_parallel_size:
        raise RuntimeError(f"world_size ({world}) is less than sequence_parallel_size {sequence_parallel_size}")
    if sequence_parallel_size <= 1:
        raise ValueError(f"sequence_parallel_size must be greater than 1, got {sequence_parallel_size}")
    if world % sequence_parallel_size != 0:
        raise RuntimeError(f"world_size ({world}) is not divisible by sequence_parallel_size "
                           f"{sequence_parallel_size}")
    assert _SEQUENCE_PARALLEL_GROUP is None, "sequence parallel group is already initialized"
    data_parallel_size = world // sequence_parallel_size
    # every rank creates every group, in the same order (torch.distributed requirement)
    for i in range(data_parallel_size):
        ranks = list(range(i * sequence_parallel_size, (i + 1) * sequence_parallel_size))
        g = dist.new_group(ranks)
        if rank in ranks:
            _SEQUENCE_PARALLEL_GROUP = g
    for j in range(sequence_parallel_size):
        ranks = list(range(j, world, sequence_parallel_size))
        g = dist.new_group(ranks)
        if rank in ranks:
            _SEQUENCE_DATA_PARALLEL_GROUP = g


def _reset_for_tests() -> None:
    global _SEQUENCE_PARALLEL_GROUP, _SEQUENCE_DATA_PARALLEL_GROUP
    _SEQUENCE_PARALLEL_GROUP = _SEQUENCE_DATA_PARALLEL_GROUP = None


def get_sequence_parallel_group():
    assert _SEQUENCE_PARALLEL_GROUP is not None, "sequence parallel group is not initialized"
    return _SEQUENCE_PARALLEL_GROUP


def get_sequence_data_parallel_group():
    assert _SEQUENCE_DATA_PARALLEL_GROUP is not None, "sequence data parallel group is not initialized"
    return _SEQUENCE_DATA_PARALLEL_GROUP


def get_sequence_parallel_world_size():
    return dist.get_world_size(group=get_sequence_parallel_group())


def get_sequence_data_parallel_world_size():
    return dist.get_world_size(group=get_sequence_data_parallel_group())


def get_sequence_parallel_rank():
    return dist.get_rank(group=get_sequence_parallel_group())


def get_sequence_data_parallel_rank():
    return dist.get_rank(group=get_sequence_data_parallel_group())
