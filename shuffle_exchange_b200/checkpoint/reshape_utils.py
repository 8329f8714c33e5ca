"""File discovery and list partitioning used when a checkpoint is re-read under another parallel degree
(reference: ``deepspeed/checkpoint/reshape_utils.py:11-98``)."""
import os
import re
from collections import OrderedDict

import torch

from .constants import ZERO_FILE_PREFIX


def basic_folder_validation(dir):
    assert os.path.exists(dir), f"{dir} path does not exist"
    assert os.path.isdir(dir), f"{dir} is not a folder"


def _key(name):
    return [int(t) if t.isdigit() else t for t in re.split(r"(\d+)", name)]


def get_files_with_prefix(all_files, prefix):
    return sorted((f for f in all_files if os.path.basename(f).startswith(prefix)), key=_key)


def validate_files(file_list):
    for f in file_list:
        if not os.path.isfile(f):
            print(f"Error: {f} is not existent")


def get_files(dir):
    out = []
    for root, _, files in os.walk(dir):
        out += [os.path.join(root, f) for f in files]
    return out


def get_zero_files(dir):
    files = get_files(dir)
    for prefix in (ZERO_FILE_PREFIX, "fp16_" + ZERO_FILE_PREFIX, "bf16_" + ZERO_FILE_PREFIX):
        found = get_files_with_prefix(files, prefix)
        if found:
            return found
    return []


def partition_data(data_list, num_partitions):
    n = len(data_list)
    assert n % num_partitions == 0, f"{n} items cannot be split evenly into {num_partitions} partitions"
    size = n // num_partitions
    return [data_list[i:i + size] for i in range(0, n, size)]


def _key_list_to_string(key_list):
    return ".".join(key_list)


def merge_state_dict(dict_a, dict_b, key_list):
    merged = type(dict_a)({})
    for k, v in dict_b.items():
        merged[k] = merge_state(dict_a[k], v, [str(k)]) if k in dict_a else v
    return merged


def merge_state_list(list_a, list_b, key_list):
    if len(list_a) != len(list_b):
        print(f"{_key_list_to_string(key_list)}")
        raise ValueError(f"Cannot merge lists of different lengths, a = {len(list_a)} b = {len(list_b)}")
    return [merge_state(a, b, key_list) for a, b in zip(list_a, list_b)]


def merge_state(state_a, state_b, key_list=[]):
    if type(state_a) != type(state_b):
        raise ValueError(f"Cannot merge two states of types {type(state_a)} and type {type(state_b)} at {_key_list_to_string(key_list)}")
    if type(state_a) in (dict, OrderedDict):
        return merge_state_dict(state_a, state_b, key_list)
    if type(state_a) in (list, tuple):
        return type(state_a)(merge_state_list(state_a, state_b, key_list))
    if torch.is_tensor(state_a):
        return torch.cat([state_a, state_b], 0)
    return state_a
