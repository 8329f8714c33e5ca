"""Helpers for re-reading a checkpoint under another parallel degree: find the shard files in their numeric order, cut
lists into equal partitions, and merge nested optimizer / module states shard by shard.  (Counterpart of
``deepspeed/checkpoint/reshape_utils.py``.)"""
import os
import re
from collections import OrderedDict
from typing import List, Sequence

import torch

from .constants import ZERO_FILE_PREFIX

_ZERO_PREFIXES = (ZERO_FILE_PREFIX, "fp16_" + ZERO_FILE_PREFIX, "bf16_" + ZERO_FILE_PREFIX)


def basic_folder_validation(dir):
    if not os.path.exists(dir):
        raise AssertionError(f"{dir} path does not exist")
    if not os.path.isdir(dir):
        raise AssertionError(f"{dir} is not a folder")


def _natural(name: str):
    """Sort key under which ``rank_10`` follows ``rank_9``."""
    return [int(tok) if tok.isdigit() else tok for tok in re.split(r"(\d+)", name)]


def get_files(dir) -> List[str]:
    return [os.path.join(root, f) for root, _, names in os.walk(dir) for f in names]


def get_files_with_prefix(all_files: Sequence[str], prefix: str) -> List[str]:
    return sorted((f for f in all_files if os.path.basename(f).startswith(prefix)), key=_natural)


def validate_files(file_list):
    for f in file_list:
        if not os.path.isfile(f):
            print(f"Error: {f} is not existent")


def get_zero_files(dir) -> List[str]:
    """Optimizer-state shard files of a checkpoint folder, whichever precision prefix they were written with."""
    everything = get_files(dir)
    return next((hit for hit in (get_files_with_prefix(everything, p) for p in _ZERO_PREFIXES) if hit), [])


def partition_data(data_list, num_partitions):
    per, left = divmod(len(data_list), num_partitions)
    if left:
        raise AssertionError(f"{len(data_list)} items cannot be split evenly into {num_partitions} partitions")
    return [data_list[i * per:(i + 1) * per] for i in range(num_partitions)]


def merge_state(state_a, state_b, key_list=()):
    """Join two shards of the same nested state: tensors are concatenated along dim 0, containers are merged element
    by element, scalars are taken from ``state_a``.  ``key_list`` is the path so far (for error messages)."""
    where = ".".join(str(k) for k in key_list)
    if type(state_a) is not type(state_b):
        raise ValueError(f"Cannot merge two states of types {type(state_a)} and type {type(state_b)} at {where}")
    if isinstance(state_a, (dict, OrderedDict)):
        out = type(state_a)()
        for k, vb in state_b.items():
            out[k] = merge_state(state_a[k], vb, (*key_list, k)) if k in state_a else vb
        return out
    if isinstance(state_a, (list, tuple)):
        if len(state_a) != len(state_b):
            raise ValueError(f"Cannot merge lists of different lengths, a = {len(state_a)} b = {len(state_b)} at {where}")
        return type(state_a)(merge_state(a, b, (*key_list, i)) for i, (a, b) in enumerate(zip(state_a, state_b)))
    if torch.is_tensor(state_a):
        return torch.cat([state_a, state_b], 0)
    return state_a


def merge_state_dict(dict_a, dict_b, key_list=()):
    return merge_state(dict_a, dict_b, key_list)


def merge_state_list(list_a, list_b, key_list=()):
    return list(merge_state(list(list_a), list(list_b), key_list))
