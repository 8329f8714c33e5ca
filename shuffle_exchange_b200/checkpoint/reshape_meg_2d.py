"""Re-partition a (pipeline × tensor) grid of checkpoint shards onto a smaller grid.

Reference: ``deepspeed/checkpoint/reshape_meg_2d.py``.  A ``meg_2d_parallel_map`` records, for every ``(pp, tp)``
coordinate of the TARGET grid, the list of SOURCE ranks (or files) whose shards are merged there.  Shrinking a dimension
by an integer factor concatenates contiguous groups of source coordinates; growing is not supported (a shard cannot be
split without knowing each tensor's partition dimension — that is what the universal checkpoint is for).
"""
from __future__ import annotations

from typing import Dict, List

from .reshape_utils import partition_data


class meg_2d_parallel_map:

    def __init__(self, pp_degree: int, tp_degree: int):
        self.pp_degree, self.tp_degree = pp_degree, tp_degree
        self.map: Dict[str, List] = {}

    def simple_init(self):
        """Identity map: coordinate (i, j) holds source rank ``i * tp + j``."""
        self.map = {self._make_key(i // self.tp_degree, i % self.tp_degree): [i]
                    for i in range(self.pp_degree * self.tp_degree)}

    def add_data(self, pp_index: int, tp_index: int, data: List):
        self._validate_indices(pp_index, tp_index)
        assert isinstance(data, list)
        self.map.setdefault(self._make_key(pp_index, tp_index), []).extend(data)

    def get_data(self, pp_index=None, tp_index=None) -> List:
        self._validate_indices(pp_index, tp_index)
        pps = range(self.pp_degree) if pp_index is None else [pp_index]
        tps = range(self.tp_degree) if tp_index is None else [tp_index]
        out: List = []
        for i in pps:
            for j in tps:
                out += self.map[self._make_key(i, j)]
        return out

    def print_data(self, tag):
        print(f"{tag}")
        for key, value in self.map.items():
            print(f"{key} = {value}")

    def _validate_indices(self, pp_index, tp_index):
        assert pp_index is None or 0 <= pp_index < self.pp_degree
        assert tp_index is None or 0 <= tp_index < self.tp_degree

    def _make_key(self, i, j):
        return f"{i},{j}"


def _reshape_tp_dimension(old_2d_map: meg_2d_parallel_map, new_tp_degree: int) -> meg_2d_parallel_map:
    new_map = meg_2d_parallel_map(old_2d_map.pp_degree, new_tp_degree)
    for i in range(old_2d_map.pp_degree):
        groups = partition_data(old_2d_map.get_data(pp_index=i), new_tp_degree)
        for j, ranks in enumerate(groups):
            new_map.add_data(i, j, ranks)
    return new_map


def _reshape_pp_dimension(old_2d_map: meg_2d_parallel_map, new_pp_degree: int) -> meg_2d_parallel_map:
    new_map = meg_2d_parallel_map(new_pp_degree, old_2d_map.tp_degree)
    for j in range(old_2d_map.tp_degree):
        groups = partition_data(old_2d_map.get_data(tp_index=j), new_pp_degree)
        for i, ranks in enumerate(groups):
            new_map.add_data(i, j, ranks)
    return new_map


def reshape_meg_2d_parallel(old_pp_degree, old_tp_degree, new_pp_degree, new_tp_degree, verbose=False):
    assert new_pp_degree <= old_pp_degree and new_tp_degree <= old_tp_degree, "2-D reshape can only shrink"
    cur = meg_2d_parallel_map(old_pp_degree, old_tp_degree)
    cur.simple_init()
    if verbose:
        cur.print_data("original_2d_map:")
    if old_tp_degree != new_tp_degree:
        cur = _reshape_tp_dimension(cur, new_tp_degree)
        if verbose:
            cur.print_data("after_tp_reshape:")
    if old_pp_degree != new_pp_degree:
        cur = _reshape_pp_dimension(cur, new_pp_degree)
        if verbose:
            cur.print_data("after_pp_reshape:")
    return cur


def get_mpu_ranks(tp_size=1, pp_size=1, dp_size=1, virtual_pp_size=None):
    """Megatron's rank-grid convention (tp fastest, then dp, then pp): returns ``(tp_groups, pp_groups, dp_groups)``."""
    world = tp_size * pp_size * dp_size
    per_stage = world // pp_size
    dp_groups, tp_groups, pp_groups = [], [], []
    for i in range(pp_size):
        lo, hi = i * per_stage, (i + 1) * per_stage
        for j in range(tp_size):
            dp_groups.append(list(range(lo + j, hi, tp_size)))
    for i in range(world // tp_size):
        tp_groups.append(list(range(i * tp_size, (i + 1) * tp_size)))
    for i in range(per_stage):
        pp_groups.append(list(range(i, world, per_stage)))
    return tp_groups, pp_groups, dp_groups


def reshape(src, tgt):
    """``src`` / ``tgt``: objects with ``tp_size``, ``pp_size``, ``dp_size`` — prints the rank groups of both layouts
    (the reference's debugging entry point) and returns them."""
    out = {}
    for name, d in (("src", src), ("tgt", tgt)):
        tp, pp, dp = get_mpu_ranks(d.tp_size, d.pp_size, d.dp_size)
        out[name] = {"tp": tp, "pp": pp, "dp": dp}
        print(f"{name}: tp={tp} pp={pp} dp={dp}")
    return out
