"""Checkpoint naming helpers (reference: ``deepspeed/checkpoint/utils.py:12-62``)."""
import os

import torch

from .constants import (BF16_ZERO_FILE_PREFIX, FP16_ZERO_FILE_PREFIX, MODEL_FILE_PREFIX, MODEL_FILE_SUFFIX,
                        OPTIM_FILE_SUFFIX, ZERO_FILE_PREFIX)


def get_model_ckpt_name_for_rank(base_folder, mp_rank_str):
    return os.path.join(base_folder, MODEL_FILE_PREFIX + mp_rank_str + MODEL_FILE_SUFFIX)


def get_zero_ckpt_name_for_rank(base_folder, dp_rank, mp_rank):
    return os.path.join(base_folder, f"{ZERO_FILE_PREFIX}{dp_rank}_{MODEL_FILE_PREFIX}{mp_rank:02d}{OPTIM_FILE_SUFFIX}")


def get_layer_ckpt_name_for_rank(base_folder, layer_id, tp_rank):
    return os.path.join(base_folder, f"{layer_id}-model_{tp_rank:02d}{MODEL_FILE_SUFFIX}")


def clone_tensors_for_torch_save(item, device=torch.device("cpu")):
    """Deep-copy so that ``torch.save`` serialises only the viewed elements, not the whole flat storage behind a view
    (saving a slice of a 10 GB flat buffer otherwise writes 10 GB)."""
    if torch.is_tensor(item):
        if type(device) is str:
            device = torch.device(device)
        return item.detach().clone().to(device) if device == item.device else item.detach().to(device)
    if isinstance(item, list):
        return [clone_tensors_for_torch_save(v, device) for v in item]
    if isinstance(item, tuple):
        return tuple(clone_tensors_for_torch_save(v, device) for v in item)
    if isinstance(item, dict):
        return type(item)({k: clone_tensors_for_torch_save(v, device) for k, v in item.items()})
    return item
