"""ZeRO optimizer-state files of a checkpoint folder viewed through a (possibly smaller) 3-D layout
(reference: ``deepspeed/checkpoint/zero_checkpoint.py``).

``ZeROCheckpoint(dir).reshape(model_3d_desc(pp, tp, dp))`` re-maps source ``*_optim_states.pt`` files onto the target
coordinates; ``get_state_for_rank`` then loads and merges the files of one target rank (flat fp32 partitions are
concatenated after their alignment padding is stripped)."""
from __future__ import annotations

from typing import List

import torch

from .constants import (BASE_OPTIMIZER_STATE, GROUP_PADDINGS, OPTIMIZER_STATE_DICT, PARAM_GROUPS, PARTITION_COUNT)
from .reshape_3d_utils import get_model_3d_descriptor, model_3d_desc
from .reshape_utils import basic_folder_validation, get_zero_files, merge_state

GROUP_STATE_KEY = "state"


class ZeROCheckpoint:

    def __init__(self, dir):
        basic_folder_validation(dir)
        self.dir = dir
        self.file_list = get_zero_files(dir)
        self.num_files = len(self.file_list)
        assert self.num_files > 0, f"No ZeRO files found in {dir}"
        self.src_3d = get_model_3d_descriptor(dir)
        self.target_3d = model_3d_desc(self.src_3d.pp_degree, self.src_3d.tp_degree, self.src_3d.dp_degree)
        self._3d_file_map = self.src_3d.reshape(self.target_3d)

    def get_src_world_size(self): return self.src_3d.world_size()
    def get_src_tp_degree(self): return self.src_3d.tp_degree
    def get_src_pp_degree(self): return self.src_3d.pp_degree
    def get_src_dp_degree(self): return self.src_3d.dp_degree

    def get_file_indices_for_rank(self, pp_index, tp_index, dp_index) -> List[int]:
        assert dp_index < len(self._3d_file_map), f"DP index {dp_index} >= DP degree {len(self._3d_file_map)}"
        return self._3d_file_map[dp_index].get_data(pp_index, tp_index)

    def get_files_for_rank(self, pp_index, tp_index, dp_index) -> List[str]:
        return [self.file_list[i] for i in self.get_file_indices_for_rank(pp_index, tp_index, dp_index)]

    def get_state_for_rank(self, pp_index, tp_index, dp_index, keys_to_ignore=(), strip_tensor_paddings=True):
        merged = None
        for path in self.get_files_for_rank(pp_index, tp_index, dp_index):
            sd = torch.load(path, map_location="cpu", weights_only=False)
            for k in keys_to_ignore:
                sd.pop(k, None)
            if strip_tensor_paddings:
                self._strip_tensor_paddings(sd)
            merged = sd if merged is None else merge_state(merged, sd)
            self._update_partition_count(merged)
            if strip_tensor_paddings:
                self._clear_group_paddings(merged)
        return merged

    def print_3d_index_map(self, tag=None):
        if tag:
            print(f"3D index map: {tag}")
        for dp, m in enumerate(self._3d_file_map):
            m.print_data(f"dp = {dp}")

    def print_3d_file_map(self, tag=None):
        if tag:
            print(f"3D file map: {tag}")
        for dp, m in enumerate(self._3d_file_map):
            for pp in range(m.pp_degree):
                for tp in range(m.tp_degree):
                    print(f"{pp}, {tp}, {dp} => {[self.file_list[i] for i in m.get_data(pp, tp)]}")

    def reshape(self, target_3d_desc: model_3d_desc):
        self.target_3d = target_3d_desc
        self._3d_file_map = self.src_3d.reshape(self.target_3d)

    # ---------------------------------------------------------------------------------------------- internals
    def _strip_tensor_paddings(self, sd):
        paddings = self._get_optimizer_state(sd, GROUP_PADDINGS)
        groups = self._get_param_group_states(sd)
        if paddings is None or groups is None:
            return
        for gid, state in groups.items():
            pad = paddings[gid] if gid < len(paddings) else 0
            if pad <= 0:
                continue
            for name, t in list(state.items()):
                if torch.is_tensor(t) and t.dim() == 1 and t.numel() > pad:
                    state[name] = t[: t.numel() - pad].clone()

    def _clear_group_paddings(self, sd):
        paddings = self._get_optimizer_state(sd, GROUP_PADDINGS)
        if paddings:
            sd[OPTIMIZER_STATE_DICT][GROUP_PADDINGS] = [0] * len(paddings)

    def _get_optimizer_state(self, sd, state_key):
        osd = sd.get(OPTIMIZER_STATE_DICT)
        return None if osd is None else osd.get(state_key)

    def _get_param_group_states(self, sd):
        osd = sd.get(OPTIMIZER_STATE_DICT)
        base = None if osd is None else osd.get(BASE_OPTIMIZER_STATE)
        if base is None:
            return None
        if isinstance(base, dict):
            return base.get(GROUP_STATE_KEY)
        return {i: s for i, s in enumerate(base)} if isinstance(base, (list, tuple)) else None

    def _update_partition_count(self, sd):
        counts = self._get_optimizer_state(sd, PARTITION_COUNT)
        if counts:
            n = self.target_3d.dp_degree
            sd[OPTIMIZER_STATE_DICT][PARTITION_COUNT] = [n] * len(counts) if isinstance(counts, (list, tuple)) else n
