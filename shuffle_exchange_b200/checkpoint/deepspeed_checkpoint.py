"""Inspector of a checkpoint directory: parallel degrees, file lists, per-layer state access.

Reference: ``deepspeed/checkpoint/deepspeed_checkpoint.py:35-317`` (Megatron-DeepSpeed 3-D layouts: ``layer_XX-model_YY``
files + ``mp_rank_XX`` + ``zero_pp_rank_X_mp_rank_XX``).  Covers what the conversion tools and the inference loaders
use: degrees, iteration, args, embedding / transformer / final-norm states merged over TP files.
"""
from __future__ import annotations

import os
import re
from typing import Dict, List

import torch

from .constants import LAYER_FILE_PREFIX, MODEL_FILE_PREFIX, ZERO_FILE_PREFIX
from .reshape_utils import basic_folder_validation, get_files, get_files_with_prefix, get_zero_files, merge_state, partition_data

EMBEDDING_LAYER_INDEX = 0
FINAL_LAYER_NORM_INDEX = -1
ARGS_KEY = "args"
CHECKPOINT_INFO_KEY = "checkpoint_info"
ITERATION_KEY = "iteration"
LAYER_FILE_PREFIX_PATTERN = r"layer_(\d+)-model_.*"
SEQUENTIAL_LAYERS = ["input_layernorm.weight", "input_layernorm.bias", "self_attention.dense.bias",
                     "post_attention_layernorm.weight", "post_attention_layernorm.bias", "mlp.dense_4h_to_h.bias",
                     "position_embeddings.weight"]
LAYER_CONCAT_DIM = {"self_attention.dense.weight": 1, "mlp.dense_4h_to_h.weight": 1}


class DeepSpeedCheckpoint:
    def __init__(self, dir, tp_degree=None, pp_degree=None, dp_degree=None, final_layer_norm_idx=FINAL_LAYER_NORM_INDEX):
        self.final_layer_norm_idx = final_layer_norm_idx
        self.dir = dir
        basic_folder_validation(dir)
        self.file_list = get_files(dir)
        self.layer_files = get_files_with_prefix(self.file_list, LAYER_FILE_PREFIX)
        self.mp_rank_files = get_files_with_prefix(self.file_list, MODEL_FILE_PREFIX)
        self.zero_files = get_zero_files(dir)
        self.layer_keys = self._get_layer_keys()
        self.layer_count = len(self.layer_keys)
        self.original_tp_degree = max(1, len(get_files_with_prefix(self.layer_files, f"{LAYER_FILE_PREFIX}01"))) \
            if self.layer_files else max(1, len(self.mp_rank_files))
        self.original_pp_degree = max(1, len(self.mp_rank_files) // self.original_tp_degree)
        self.original_dp_degree = max(1, len(self.zero_files) // max(1, self.original_tp_degree * self.original_pp_degree))
        self.tp_degree = self.original_tp_degree if tp_degree is None else tp_degree
        self.pp_degree = self.original_pp_degree if pp_degree is None else pp_degree
        self.dp_degree = self.original_dp_degree if dp_degree is None else dp_degree
        self.original_world_size = self.original_tp_degree * self.original_pp_degree * self.original_dp_degree
        self.world_size = self.tp_degree * self.pp_degree * self.dp_degree
        self.global_state: Dict = {}
        self._sanity_check()

    def is_change_tp_degree(self):
        return self.tp_degree != self.original_tp_degree

    def is_change_pp_degree(self):
        return self.pp_degree != self.original_pp_degree

    def is_change_dp_degree(self):
        return self.dp_degree != self.original_dp_degree

    def show_file_lists(self):
        for name, lst in (("layer", self.layer_files), ("mp_rank", self.mp_rank_files), ("zero", self.zero_files)):
            print(f"{name} files: {len(lst)}")

    def _load(self, path):
        return torch.load(path, map_location="cpu", weights_only=False)

    def _build_global_state(self):
        sd = self._load(self.mp_rank_files[0]) if self.mp_rank_files else {}
        self.global_state[ITERATION_KEY] = sd.get(ITERATION_KEY, 0)
        self.global_state[ARGS_KEY] = sd.get(ARGS_KEY, None)

    def get_iteration(self):
        if ITERATION_KEY not in self.global_state:
            self._build_global_state()
        return self.global_state[ITERATION_KEY]

    def get_args(self):
        if ARGS_KEY not in self.global_state:
            self._build_global_state()
        return self.global_state[ARGS_KEY]

    def get_checkpoint_info(self, info_key=CHECKPOINT_INFO_KEY):
        sd = self._load(self.mp_rank_files[0]) if self.mp_rank_files else {}
        return sd.get(info_key, None)

    def _get_layer_keys(self) -> List[str]:
        keys = set()
        for f in self.layer_files:
            m = re.search(LAYER_FILE_PREFIX_PATTERN, os.path.basename(f))
            if m:
                keys.add(m.group(1))
        return sorted(keys, key=int)

    def _layer_files_for(self, key) -> List[str]:
        return get_files_with_prefix(self.layer_files, f"{LAYER_FILE_PREFIX}{key}-")

    def _merge_tp(self, files, tp_index):
        """State of logical TP rank ``tp_index`` assembled from the original TP files it covers."""
        groups = partition_data(files, self.tp_degree) if len(files) % self.tp_degree == 0 else [files]
        mine = groups[min(tp_index, len(groups) - 1)]
        sds = [self._load(f) for f in mine]
        out = sds[0]
        for nxt in sds[1:]:
            merged = type(out)()
            for k, v in out.items():
                if not torch.is_tensor(v) or any(k.endswith(s) for s in SEQUENTIAL_LAYERS):
                    merged[k] = v
                else:
                    dim = next((d for s, d in LAYER_CONCAT_DIM.items() if k.endswith(s)), 0)
                    merged[k] = torch.cat([v, nxt[k]], dim=dim)
            out = merged
        return out

    def get_embedding_state(self, tp_index: int) -> Dict:
        assert self.layer_keys, "checkpoint has no layer files"
        return self._merge_tp(self._layer_files_for(self.layer_keys[EMBEDDING_LAYER_INDEX]), tp_index)

    def get_final_norm_state(self, tp_index: int) -> Dict:
        assert self.layer_keys
        return self._merge_tp(self._layer_files_for(self.layer_keys[self.final_layer_norm_idx]), tp_index)

    def get_transformer_state(self, tp_index: int, pp_index: int) -> List[Dict]:
        keys = self.layer_keys[1:self.final_layer_norm_idx] if self.final_layer_norm_idx < 0 else self.layer_keys[1:]
        per_stage = partition_data(keys, self.pp_degree) if len(keys) % self.pp_degree == 0 else [keys]
        return [self._merge_tp(self._layer_files_for(k), tp_index) for k in per_stage[min(pp_index, len(per_stage) - 1)]]

    def get_zero_files(self, pp_index=0, tp_index=0, dp_index=None) -> List[str]:
        return self.zero_files

    def _sanity_check(self):
        if self.layer_files:
            assert len(self.mp_rank_files) % max(1, self.original_tp_degree) == 0
