"""(pipeline, tensor, data)-parallel layout descriptors of a checkpoint folder and their reshaping
(reference: ``deepspeed/checkpoint/reshape_3d_utils.py``)."""
from __future__ import annotations

from typing import List

from .constants import LAYER_FILE_PREFIX, MODEL_FILE_PREFIX, ZERO_FILE_PREFIX
from .reshape_meg_2d import meg_2d_parallel_map, reshape_meg_2d_parallel
from .reshape_utils import basic_folder_validation, get_files, get_files_with_prefix, get_zero_files, partition_data

PP_DIM, TP_DIM, DP_DIM = "PP", "TP", "DP"


class model_3d_desc:

    def __init__(self, pp_degree=1, tp_degree=1, dp_degree=1):
        self.pp_degree, self.tp_degree, self.dp_degree = pp_degree, tp_degree, dp_degree

    def reshape(self, target_3d_desc: "model_3d_desc", verbose=False) -> List[meg_2d_parallel_map]:
        """One 2-D (pp × tp) map per TARGET data-parallel index; entries are global source ranks
        ``dp * (pp*tp) + pp_i * tp + tp_j`` of this (source) layout."""
        ok, errs = self.can_reshape(target_3d_desc)
        assert ok, ",".join(errs)
        tgt_2d = reshape_meg_2d_parallel(self.pp_degree, self.tp_degree, target_3d_desc.pp_degree,
                                         target_3d_desc.tp_degree, verbose)
        flat = flatten_dp_dimension(tgt_2d, self.pp_degree * self.tp_degree, self.dp_degree)
        return unflatten_dp_dimension(flat, target_3d_desc.dp_degree)

    def get_desc(self) -> str:
        return f"{PP_DIM},{TP_DIM},{DP_DIM} = ({self.pp_degree}, {self.tp_degree}, {self.dp_degree})"

    def world_size(self) -> int:
        return self.pp_degree * self.tp_degree * self.dp_degree

    def is_valid(self, pp_index, tp_index, dp_index):
        errs = [f"{name} index {i} >= degree {deg}" for name, i, deg in
                ((PP_DIM, pp_index, self.pp_degree), (TP_DIM, tp_index, self.tp_degree), (DP_DIM, dp_index, self.dp_degree))
                if i >= deg]
        return not errs, errs

    def can_reshape(self, target_3d_desc: "model_3d_desc"):
        errs = [f"Expansion reshape not supported - {name}: {mine} ---> {theirs}" for name, mine, theirs in
                ((PP_DIM, self.pp_degree, target_3d_desc.pp_degree), (TP_DIM, self.tp_degree, target_3d_desc.tp_degree),
                 (DP_DIM, self.dp_degree, target_3d_desc.dp_degree)) if theirs > mine]
        return not errs, errs


def get_model_3d_descriptor(dir) -> model_3d_desc:
    """Infer the layout from file counts: ``layer_01-model_*`` files = tp; ``mp_rank_*`` files = tp·pp;
    ``*zero_pp_rank_*`` files = tp·pp·dp."""
    basic_folder_validation(dir)
    files = get_files(dir)
    zero_files = get_zero_files(dir)
    n_mp = len(get_files_with_prefix(files, MODEL_FILE_PREFIX))
    if get_files_with_prefix(files, LAYER_FILE_PREFIX):
        tp = len(get_files_with_prefix(files, f"{LAYER_FILE_PREFIX}01"))
        pp = max(1, n_mp // max(1, tp))
        dp = max(1, len(zero_files) // max(1, pp * tp))
    else:
        tp, pp = max(1, n_mp), 1
        dp = max(1, len(zero_files) // tp)
    return model_3d_desc(pp, tp, dp)


def flatten_dp_dimension(meg_2d_map: meg_2d_parallel_map, src_2d_size: int, dp_degree: int) -> meg_2d_parallel_map:
    """Each (pp, tp) entry gains the same ranks of every data-parallel replica (offset ``dp * src_2d_size``), ordered
    replica-major so consecutive chunks belong to consecutive dp indices."""
    new = meg_2d_parallel_map(meg_2d_map.pp_degree, meg_2d_map.tp_degree)
    for i in range(meg_2d_map.pp_degree):
        for j in range(meg_2d_map.tp_degree):
            base = meg_2d_map.get_data(i, j)
            for d in range(dp_degree):
                new.add_data(i, j, [r + d * src_2d_size for r in base])
    return new


def unflatten_dp_dimension(meg_2d_map: meg_2d_parallel_map, dp_degree: int) -> List[meg_2d_parallel_map]:
    maps = [meg_2d_parallel_map(meg_2d_map.pp_degree, meg_2d_map.tp_degree) for _ in range(dp_degree)]
    for i in range(meg_2d_map.pp_degree):
        for j in range(meg_2d_map.tp_degree):
            for d, ranks in enumerate(partition_data(meg_2d_map.get_data(i, j), dp_degree)):
                maps[d].add_data(i, j, ranks)
    return maps
