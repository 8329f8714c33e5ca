"""Cartesian process topologies and the pipeline/data/model process-group grid.

Reference: deepspeed/runtime/pipe/topology.py — ``ProcessTopology`` :12-215 (rank <-> coordinate),
``PipeDataParallelTopology`` :232, ``PipeModelDataParallelTopology`` :244, ``PipelineParallelGrid``
:251-472.  Ranks are laid out row-major over the axes in the order given (last axis fastest).
"""
from __future__ import annotations

from collections import namedtuple
from itertools import product
from typing import List

from ...comm import comm as dist


class ProcessTopology:
    def __init__(self, axes: List[str], dims: List[int]):
        self.axes = list(axes)
        self.dims = list(dims)
        self.ProcessCoord = namedtuple("ProcessCoord", axes)
        self.mapping = {}
        for rank, coord in enumerate(product(*[range(d) for d in dims])):
            self.mapping[self.ProcessCoord(*coord)] = rank

    def get_rank(self, **coord_kwargs) -> int:
        if len(coord_kwargs) != len(self.axes):
            raise ValueError("get_rank() does not support slices. Use filter_match())")
        key = self.ProcessCoord(**coord_kwargs)
        assert key in self.mapping, f"key {coord_kwargs} invalid"
        return self.mapping[key]

    def get_axis_names(self):
        return self.axes

    def get_rank_repr(self, rank, omit_axes=("data", "pipe"), inner_sep="_", outer_sep="-") -> str:
        omit = frozenset(omit_axes)
        coord = self.get_coord(rank)
        return outer_sep.join(f"{ax}{inner_sep}{getattr(coord, ax):02d}" for ax in self.axes if ax not in omit)

    def get_dim(self, axis) -> int:
        return self.dims[self.axes.index(axis)] if axis in self.axes else 0

    def get_coord(self, rank):
        for coord, r in self.mapping.items():
            if r == rank:
                return coord
        raise ValueError(f"rank {rank} not found in topology.")

    def get_axis_comm_lists(self, axis) -> List[List[int]]:
        """Rank lists that vary only along ``axis`` (one list per combination of the other axes)."""
        if axis not in self.axes:
            return []
        others = [a for a in self.axes if a != axis]
        lists = []
        for combo in product(*[range(self.get_dim(a)) for a in others]):
            fixed = dict(zip(others, combo))
            lists.append([self.mapping[self.ProcessCoord(**fixed, **{axis: i})] for i in range(self.get_dim(axis))])
        return lists

    def filter_match(self, **filter_kwargs) -> List[int]:
        return [r for c, r in self.mapping.items()
                if all(getattr(c, k) == v for k, v in filter_kwargs.items())]

    def get_axis_list(self, axis, idx) -> List[int]:
        ax = self.axes.index(axis)
        return [r for c, r in self.mapping.items() if c[ax] == idx]

    def world_size(self) -> int:
        return len(self.mapping)

    def __str__(self):
        return str(self.mapping)


def _prime_factors(n: int) -> List[int]:
    out, d = [], 2
    while d * d <= n:
        while n % d == 0:
            out.append(d)
            n //= d
        d += 1
    if n > 1:
        out.append(n)
    return out


class PipeDataParallelTopology(ProcessTopology):
    """axes (pipe, data): adjacent ranks are data-parallel peers (fast all-reduce), pipeline stages
    are strided."""

    def __init__(self, num_pp, num_dp):
        super().__init__(axes=["pipe", "data"], dims=[num_pp, num_dp])


class PipeModelDataParallelTopology(ProcessTopology):
    def __init__(self, num_pp, num_mp, num_dp):
        super().__init__(axes=["pipe", "data", "model"], dims=[num_pp, num_dp, num_mp])


class PipelineParallelGrid:
    """Process groups for a (pipe, data[, model]) topology; also serves as an ``mpu``."""

    def __init__(self, topology=None, process_group=None):
        self.global_rank = dist.get_rank()
        self.world_size = dist.get_world_size()
        if topology is not None:
            self._topo = topology
        else:
            num_pp = num_dp = 1
            for idx, prime in enumerate(_prime_factors(self.world_size)):
                if idx % 2 == 0:
                    num_pp *= prime
                else:
                    num_dp *= prime
            self._topo = PipeDataParallelTopology(num_dp=num_dp, num_pp=num_pp)
        self.data_parallel_size = max(self._topo.get_dim("data"), 1)
        self.pipe_parallel_size = max(self._topo.get_dim("pipe"), 1)
        self.model_parallel_size = max(self._topo.get_dim("model"), 1)
        self.slice_parallel_size = self.model_parallel_size
        assert self._is_grid_valid(), "Invalid Grid"
        self.stage_id = self.get_stage_id()
        self.data_parallel_id = self.get_data_parallel_id()

        def build(lists):
            mine, mine_ranks = None, None
            for ranks in lists:
                g = dist.new_group(ranks=ranks) if dist.is_initialized() else None
                if self.global_rank in ranks:
                    mine, mine_ranks = g, ranks
            return mine, mine_ranks

        self.ds_model_proc_group = None
        self.ds_model_rank = -1
        for dp in range(self.data_parallel_size):
            ranks = sorted(self._topo.get_axis_list(axis="data", idx=dp))
            g = dist.new_group(ranks=ranks) if dist.is_initialized() else None
            if self.global_rank in ranks:
                self.ds_model_proc_group = g
                self.ds_model_world_size = len(ranks)
                self.ds_model_rank = ranks.index(self.global_rank)
        assert self.ds_model_rank > -1
        self.dp_groups = self._topo.get_axis_comm_lists("data")
        self.dp_proc_group, self.dp_group = build(self.dp_groups)
        self.is_first_stage = self.stage_id == 0
        self.is_last_stage = self.stage_id == self.pipe_parallel_size - 1
        self.p2p_groups = self._build_p2p_groups()
        self.pp_groups = self._topo.get_axis_comm_lists("pipe")
        self.pp_proc_group, self.pp_group = build(self.pp_groups)
        self.pp_group = self.pp_group or [self.global_rank]
        if self.model_parallel_size > 1:
            self.slice_proc_group, self.slice_group = build(self._topo.get_axis_comm_lists("model"))
        else:
            self.slice_proc_group, self.slice_group = build([[r] for r in range(self.world_size)])

    def get_stage_id(self):
        return self._topo.get_coord(rank=self.global_rank).pipe

    def get_data_parallel_id(self):
        return self._topo.get_coord(rank=self.global_rank).data

    def _build_p2p_groups(self):
        """[rank, next-stage rank] pairs along every pipe line (wrapping)."""
        pairs = []
        for ranks in self._topo.get_axis_comm_lists("pipe"):
            assert len(ranks) == self.pipe_parallel_size
            for i, r in enumerate(ranks):
                pairs.append([r, ranks[(i + 1) % self.pipe_parallel_size]])
        return pairs

    def _is_grid_valid(self):
        ranks = 1
        for ax in self._topo.get_axis_names():
            ranks *= self._topo.get_dim(ax)
        return ranks == dist.get_world_size()

    def stage_to_global(self, stage_id, **kwargs):
        me = self._topo.get_coord(self.global_rank)
        transform = me._replace(pipe=stage_id, **kwargs)._asdict()
        return self._topo.get_rank(**transform)

    def topology(self):
        return self._topo

    # ---- mpu interface ----------------------------------------------------------------------
    def get_global_rank(self): return self.global_rank
    def get_pipe_parallel_rank(self): return self.get_stage_id()
    def get_pipe_parallel_world_size(self): return self.pipe_parallel_size
    def get_pipe_parallel_group(self): return self.pp_proc_group
    def get_data_parallel_rank(self): return self.data_parallel_id
    def get_data_parallel_world_size(self): return self.data_parallel_size
    def get_data_parallel_group(self): return self.dp_proc_group
    def get_model_parallel_rank(self): return self.ds_model_rank
    def get_model_parallel_world_size(self): return self.ds_model_world_size
    def get_model_parallel_group(self): return self.ds_model_proc_group
    def get_slice_parallel_rank(self):
        return self._topo.get_coord(self.global_rank).model if "model" in self._topo.get_axis_names() else 0
    def get_slice_parallel_world_size(self): return self.slice_parallel_size
    # Megatron-style aliases
    def get_pipeline_model_parallel_rank(self): return self.get_pipe_parallel_rank()
    def get_pipeline_model_parallel_world_size(self): return self.get_pipe_parallel_world_size()
    def get_tensor_model_parallel_rank(self): return self.get_slice_parallel_rank()
    def get_tensor_model_parallel_world_size(self): return self.get_slice_parallel_world_size()
    def get_data_parallel_group_ranks(self):
        return next((list(g) for g in self.dp_groups if self.global_rank in g), [self.global_rank])
    def get_slice_parallel_group(self): return self.slice_proc_group
