"""Inspector of a checkpoint directory: parallel degrees, file maps, per-layer state access under a (smaller) target
layout.

Reference: ``deepspeed/checkpoint/deepspeed_checkpoint.py:35-317`` (Megatron-DeepSpeed 3-D layouts: ``layer_XX-model_YY``
files + ``mp_rank_XX`` + ``zero_pp_rank_X_mp_rank_XX``).  Everything the object answers is derived once, in the
constructor, as four tables:

* ``tp_to_embedding_map`` / ``tp_to_final_norm_map``: target TP index -> the source layer files it merges;
* ``pp_to_transformer_map``: target PP index -> the transformer layer keys of that stage;
* ``transformer_file_map``: ``(tp, pp)`` -> per layer, the source files of that target rank;
* ``new_2d_map``: ``(pp, tp)`` -> source ``mp_rank`` file indices (``reshape_meg_2d_parallel``),

plus a :class:`ZeROCheckpoint` reshaped to the target 3-D layout for the optimizer files.
"""
from __future__ import annotations

import os
import re
from typing import Dict, List, Tuple

import torch

from .constants import LAYER_FILE_PREFIX, MODEL_FILE_PREFIX
from .reshape_3d_utils import model_3d_desc
from .reshape_meg_2d import meg_2d_parallel_map, reshape_meg_2d_parallel
from .reshape_utils import (basic_folder_validation, get_files, get_files_with_prefix, get_zero_files, merge_state,
                            partition_data)
from .zero_checkpoint import ZeROCheckpoint

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

        self.old_2d_map = meg_2d_parallel_map(self.original_pp_degree, self.original_tp_degree)
        self.old_2d_map.simple_init()
        self.new_2d_map = reshape_meg_2d_parallel(self.original_pp_degree, self.original_tp_degree, self.pp_degree,
                                                  self.tp_degree)
        self.zero_checkpoint = None
        if self.zero_files:
            self.zero_checkpoint = ZeROCheckpoint(dir)
            if self.is_change_pp_degree() or self.is_change_tp_degree() or self.is_change_dp_degree():
                self.zero_checkpoint.reshape(model_3d_desc(self.pp_degree, self.tp_degree, self.dp_degree))

        self.global_state: Dict = {}
        self._sanity_check()
        self.tp_to_embedding_map = self._files_by_tp(EMBEDDING_LAYER_INDEX)
        self.tp_to_final_norm_map = self._files_by_tp(self.final_layer_norm_idx)
        self.pp_to_transformer_map = self._layers_by_pp()
        self.transformer_file_map = self._files_by_tp_pp()

    # ---------------------------------------------------------------------------------- degrees
    def is_change_tp_degree(self):
        return self.tp_degree != self.original_tp_degree

    def is_change_pp_degree(self):
        return self.pp_degree != self.original_pp_degree

    def is_change_dp_degree(self):
        return self.dp_degree != self.original_dp_degree

    # ----------------------------------------------------------------------------- table building
    def _get_layer_keys(self) -> List[str]:
        keys = set()
        for f in self.layer_files:
            m = re.search(LAYER_FILE_PREFIX_PATTERN, os.path.basename(f))
            if m:
                keys.add(m.group(1))
        return sorted(keys, key=int)

    def _layer_files_for(self, key) -> List[str]:
        return get_files_with_prefix(self.layer_files, f"{LAYER_FILE_PREFIX}{key}-")

    def _transformer_keys(self) -> List[str]:
        if not self.layer_keys:
            return []
        stop = self.final_layer_norm_idx if self.final_layer_norm_idx < 0 else len(self.layer_keys)
        return self.layer_keys[1:stop]

    def _files_by_tp(self, layer_index: int) -> Dict[int, List[str]]:
        if not self.layer_keys:
            return {}
        files = self._layer_files_for(self.layer_keys[layer_index])
        assert len(files) % self.tp_degree == 0, \
            f"{len(files)} files of layer {self.layer_keys[layer_index]} do not divide over tp={self.tp_degree}"
        return dict(enumerate(partition_data(files, self.tp_degree)))

    def _layers_by_pp(self) -> Dict[int, List[str]]:
        keys = self._transformer_keys()
        if not keys:
            return {}
        assert len(keys) % self.pp_degree == 0, f"{len(keys)} transformer layers do not divide over pp={self.pp_degree}"
        return dict(enumerate(partition_data(keys, self.pp_degree)))

    def _files_by_tp_pp(self) -> Dict[Tuple[int, int], List[List[str]]]:
        table: Dict[Tuple[int, int], List[List[str]]] = {}
        for pp, keys in self.pp_to_transformer_map.items():
            for key in keys:
                for tp, files in enumerate(partition_data(self._layer_files_for(key), self.tp_degree)):
                    table.setdefault((tp, pp), []).append(files)
        return table

    # ----------------------------------------------------------------------------------- printing
    @staticmethod
    def _dump_mapping(data_map, map_tag=None):
        if map_tag is not None:
            print(map_tag)
        for k, v in data_map.items():
            print(f"{k} = {v}")

    def show_2d_mapping(self):
        print("reshaped 2d map ---- begin")
        for i in range(self.pp_degree):
            for j in range(self.tp_degree):
                print(f"[{i}, {j}] = {self.new_2d_map.get_data(pp_index=i, tp_index=j)}")
        print("reshaped 2d map ---- end")

    def show_tp_embedding_map(self):
        self._dump_mapping(self.tp_to_embedding_map, "tp_to_embedding_layers")

    def show_tp_final_norm_map(self):
        self._dump_mapping(self.tp_to_final_norm_map, "tp_to_final_norm_layers")

    def show_pp_transformer_map(self):
        self._dump_mapping(self.pp_to_transformer_map, "pp_to_transformer_layers")

    def show_transformer_file_map(self):
        self._dump_mapping(self.transformer_file_map, "rank_to_transformer_files")

    def show_file_lists(self):
        for name, lst in (("layer", self.layer_files), ("mp_rank", self.mp_rank_files), ("zero", self.zero_files)):
            print(f"{name} files: {len(lst)}")

    # ------------------------------------------------------------------------------- global state
    def _load(self, path):
        return torch.load(path, map_location="cpu", weights_only=False)

    def _build_global_state(self):
        sd = self._load(self.mp_rank_files[0]) if self.mp_rank_files else {}
        self.global_state[ITERATION_KEY] = sd.get(ITERATION_KEY, 0)
        self.global_state[ARGS_KEY] = sd.get(ARGS_KEY, None)

    def _get_checkpoint_value(self, key):
        if key not in self.global_state:
            self._build_global_state()
        return self.global_state[key]

    def get_iteration(self):
        return self._get_checkpoint_value(ITERATION_KEY)

    def get_args(self):
        return self._get_checkpoint_value(ARGS_KEY)

    def get_checkpoint_info(self, info_key=CHECKPOINT_INFO_KEY):
        sd = self._load(self.mp_rank_files[0]) if self.mp_rank_files else {}
        return sd.get(info_key, None)

    # ------------------------------------------------------------------------------ optimizer files
    def get_zero_checkpoint_state(self, pp_index, tp_index, dp_index) -> dict:
        assert self.zero_checkpoint is not None, f"no ZeRO optimizer files in {self.dir}"
        return self.zero_checkpoint.get_state_for_rank(pp_index=pp_index, tp_index=tp_index, dp_index=dp_index,
                                                       keys_to_ignore=[])

    def get_zero_files(self, pp_index=None, tp_index=None, dp_index=None) -> List[str]:
        """Optimizer files of one target rank; with no coordinates, all of them."""
        if pp_index is None and tp_index is None and dp_index is None:
            return self.zero_files
        assert self.zero_checkpoint is not None, f"no ZeRO optimizer files in {self.dir}"
        return self.zero_checkpoint.get_files_for_rank(pp_index=pp_index or 0, tp_index=tp_index or 0,
                                                       dp_index=dp_index or 0)

    # --------------------------------------------------------------------------------- layer files
    def get_embedding_layer_id(self):
        return self.layer_keys[EMBEDDING_LAYER_INDEX]

    def get_final_norm_layer_id(self):
        return self.layer_keys[self.final_layer_norm_idx]

    def get_embedding_files(self, tp_index: int) -> List[str]:
        assert tp_index in self.tp_to_embedding_map
        return self.tp_to_embedding_map[tp_index]

    def get_final_norm_files(self, tp_index: int) -> List[str]:
        assert tp_index in self.tp_to_final_norm_map
        return self.tp_to_final_norm_map[tp_index]

    def get_pp_transformer_map(self, pp_index: int) -> List[str]:
        assert pp_index < self.pp_degree
        return self.pp_to_transformer_map[pp_index]

    def _merge_layer_files(self, files: List[str]) -> Dict:
        """One state dict out of the TP slices ``files``: sequential tensors are taken from the first slice, the rest
        are concatenated along their tensor-parallel dimension."""
        sds = [self._load(f) for f in files]
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
        return self._merge_layer_files(self.get_embedding_files(tp_index))

    def get_final_norm_state(self, tp_index: int) -> Dict:
        """Norm parameters are replicated over tensor parallelism: the first slice is the state."""
        return self._load(self.get_final_norm_files(tp_index)[0])

    def get_transformer_state(self, tp_index: int, pp_index: int) -> List[Dict]:
        assert tp_index < self.tp_degree and pp_index < self.pp_degree
        return [self._merge_layer_files(files) for files in self.transformer_file_map.get((tp_index, pp_index), [])]

    # ------------------------------------------------------------------------------- mp_rank files
    def get_2d_parallel_files(self, tp_index: int, pp_index: int) -> List[str]:
        assert tp_index < self.tp_degree and pp_index < self.pp_degree
        return [self.mp_rank_files[i] for i in self.new_2d_map.get_data(pp_index=pp_index, tp_index=tp_index)]

    def get_2d_parallel_state(self, tp_index: int, pp_index: int) -> dict:
        merged = None
        for path in self.get_2d_parallel_files(tp_index, pp_index):
            sd = self._load(path)
            merged = sd if merged is None else merge_state(merged, sd)
        return merged

    # ------------------------------------------------------------------------------------- checks
    def validate_files(self):
        for f in self.file_list:
            if not os.path.isfile(f):
                print(f"Error: {f} is not existent")

    def _sanity_check(self):
        if self.layer_files:
            assert len(self.mp_rank_files) % max(1, self.original_tp_degree) == 0
            assert self.layer_count > 2, "a layered checkpoint has at least embedding, one block and the final norm"
        assert self.tp_degree <= self.original_tp_degree and self.pp_degree <= self.original_pp_degree, \
            "the target layout can only shrink the tensor / pipeline degrees"
