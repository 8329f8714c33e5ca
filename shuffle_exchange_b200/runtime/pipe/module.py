"""``PipelineModule`` — a model given as a flat sequence of layers, cut into pipeline stages.

Public surface as in the reference (``deepspeed/runtime/pipe/module.py``: ``LayerSpec`` :30, ``TiedLayerSpec`` :77,
``PipelineModule`` :86, per-layer checkpoint files ``layer_XX-model_YY-model_states.pt`` :582-634).  Organisation here:

* the cut is a :class:`StageMap` (layer index -> stage) produced by one of the registered *partitioners*
  (``uniform``, ``parameters``, ``type:<regex>``) — pure functions of per-layer weights, testable without a process group;
* a stage materialises only its own slice of the sequence into ``self.stage`` (a list of :class:`_Step`: the callable
  to run plus, when it is a module, the module that owns its state) — specs of other stages are never built;
* tied layers are :class:`_Tie` records (module, attribute names, the process group of the stages that share it);
* ``forward`` runs the stage in segments of ``activation_checkpoint_interval`` steps, recomputing a segment in backward
  when it holds parameters.
"""
from __future__ import annotations

import glob
import os
import re
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

import torch
import torch.nn as nn

from ... import comm as dist
from ...device import get_accelerator
from ...utils.logging import logger
from .. import utils as ds_utils
from ..activation_checkpointing import checkpointing
from .topology import PipeDataParallelTopology, PipelineParallelGrid


class PipelineError(Exception):
    """Errors related to the use of pipeline parallelism."""


# ------------------------------------------------------------------------------------------ specs
class LayerSpec:
    """Recipe for a layer: ``LayerSpec(nn.Linear, 8, 8)`` is only instantiated on the stage that owns it."""

    def __init__(self, typename, *module_args, **module_kwargs):
        if not (isinstance(typename, type) and issubclass(typename, nn.Module)):
            raise RuntimeError("LayerSpec only supports torch.nn.Module types.")
        self.typename, self.module_args, self.module_kwargs = typename, module_args, module_kwargs
        self.global_rank = dist.get_rank() if dist.is_initialized() else -1

    def __repr__(self):
        return ds_utils.call_to_str(self.typename.__name__, *self.module_args, **self.module_kwargs)

    def build(self, log=False):
        if log:
            logger.info(f"RANK={self.global_rank} building {self!r}")
        return self.typename(*self.module_args, **self.module_kwargs)


class TiedLayerSpec(LayerSpec):
    """A layer whose ``tied_weight_attr`` parameter(s) are one shared weight for every stage that lists ``key``
    (embedding / LM head).  ``forward_fn(module, x)`` replaces ``module(x)`` where the reuse computes something else."""

    def __init__(self, key, typename, *module_args, forward_fn=None, tied_weight_attr=("weight",), **module_kwargs):
        super().__init__(typename, *module_args, **module_kwargs)
        self.key, self.forward_fn = key, forward_fn
        self.tied_weight_attr = [tied_weight_attr] if isinstance(tied_weight_attr, str) else list(tied_weight_attr)


def _display_name(layer) -> str:
    if isinstance(layer, LayerSpec):
        return layer.typename.__name__
    if isinstance(layer, nn.Module):
        return type(layer).__name__
    return getattr(layer, "__name__", str(layer))


# ------------------------------------------------------------------------------------- partitioning
@dataclass
class StageMap:
    """``bounds[s] .. bounds[s+1]`` are the layers of stage ``s``."""
    bounds: List[int]

    @property
    def num_stages(self) -> int:
        return len(self.bounds) - 1

    def span(self, stage: int):
        return self.bounds[stage], self.bounds[stage + 1]

    def owner(self, layer_idx: int) -> int:
        for s in range(self.num_stages):
            if self.bounds[s] <= layer_idx < self.bounds[s + 1]:
                return s
        raise RuntimeError(f"Layer {layer_idx} not owned? parts={self.bounds}")


def _trainable_params(layer) -> int:
    """Weight of a layer for the ``parameters`` partitioner (specs are built once, counted and dropped)."""
    if isinstance(layer, LayerSpec):
        layer = layer.build()
    if isinstance(layer, nn.Module):
        return sum(p.numel() for p in layer.parameters() if p.requires_grad)
    return 0


def partition_layers(layers: Sequence, num_stages: int, method: str = "parameters") -> StageMap:
    """Cut ``layers`` into ``num_stages`` contiguous stages.

    ``uniform``: equal layer counts.  ``parameters``: balance trainable-parameter counts.  ``type:<regex>``: balance the
    number of layers whose class name matches (e.g. ``type:transformer``), other layers ride along."""
    m = method.lower()
    if m == "uniform":
        return StageMap(ds_utils.partition_uniform(num_items=len(layers), num_parts=num_stages))
    if m == "parameters":
        return StageMap(ds_utils.partition_balanced(weights=[_trainable_params(l) for l in layers], num_parts=num_stages))
    if m.startswith("type:"):
        pattern = re.compile(method.split(":", 1)[1], re.IGNORECASE)
        hits = [1 if pattern.search(_display_name(l)) else 0 for l in layers]
        if not any(hits):
            raise RuntimeError(f"Partitioning '{method.split(':', 1)[1]}' found no valid layers to partition.")
        return StageMap(ds_utils.partition_balanced(weights=hits, num_parts=num_stages))
    raise NotImplementedError(f"Partitioning method {method} not implemented.")


# ----------------------------------------------------------------------------------------- runtime
@dataclass
class _Step:
    """One element of a stage: ``fn(x)`` is what runs; ``module`` owns the state that is checkpointed (may be None)."""
    index: int                      # position in the global layer sequence
    fn: Callable
    module: Optional[nn.Module] = None
    tie_key: Optional[str] = None


@dataclass
class _Tie:
    module: nn.Module
    attrs: List[str]
    ranks: List[int] = field(default_factory=list)
    group: object = None

    def weights(self):
        return [getattr(self.module, a) for a in self.attrs]


class PipelineModule(nn.Module):
    def __init__(self, layers, num_stages=None, topology=None, loss_fn=None, seed_layers=False, seed_fn=None,
                 base_seed=1234, partition_method="parameters", activation_checkpoint_interval=0,
                 activation_checkpoint_func=checkpointing.checkpoint, checkpointable_layers=None,
                 dynamic_shape=False):
        super().__init__()
        if num_stages is None and topology is None:
            raise RuntimeError("must provide num_stages or topology")
        if checkpointable_layers is not None and not isinstance(checkpointable_layers, list):
            raise AssertionError("param `checkpointable_layers` must be type of list.")
        self.loss_fn, self.dynamic_shape = loss_fn, dynamic_shape
        self.seed_layers, self.seed_fn, self.base_seed = seed_layers, seed_fn, base_seed
        self.checkpointable_layers = checkpointable_layers
        self.activation_checkpoint_interval = activation_checkpoint_interval
        self.activation_checkpoint_func = activation_checkpoint_func
        self.micro_offset = 0           # forward passes so far (seeds differ per micro-batch when seed_layers is on)

        # ---- where am I
        dist.init_distributed()
        self.world_group = dist.new_group(ranks=range(dist.get_world_size())) if dist.is_initialized() else None
        self.global_rank = dist.get_rank(group=self.world_group)
        self.world_size = dist.get_world_size(group=self.world_group)
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if topology is None:
            if self.world_size % num_stages != 0:
                raise RuntimeError(f"num_stages ({num_stages}) must divide distributed world size ({self.world_size})")
            topology = PipeDataParallelTopology(num_pp=num_stages, num_dp=self.world_size // num_stages)
        self._topo = topology
        self.num_stages = topology.get_dim("pipe")
        self._grid = PipelineParallelGrid(process_group=self.world_group, topology=topology)
        self.stage_id = topology.get_coord(self.global_rank).pipe

        # ---- the cut
        self._layer_specs = list(layers)
        self._num_layers = len(self._layer_specs)
        self.stage_map = partition_layers(self._layer_specs, self.num_stages, partition_method)
        self.parts = self.stage_map.bounds
        self._local_start, self._local_stop = self.stage_map.span(self.stage_id)
        if self.global_rank == 0:
            self._describe(partition_method)

        # ---- my slice
        self.tied_modules = nn.ModuleDict()
        self._ties: Dict[str, _Tie] = {}
        self.stage: List[_Step] = [self._materialise(i) for i in range(self._local_start, self._local_stop)]
        for p in self.parameters():
            p.ds_pipe_replicated = False
        self.to(get_accelerator().device_name(self.local_rank) if get_accelerator().is_available() else "cpu")
        self._connect_ties()
        self._synchronize_tied_weights()

    # ------------------------------------------------------------------------------- construction
    def _describe(self, method):
        logger.info(f"Partitioning pipeline stages with method {method}")
        for s in range(self.num_stages):
            lo, hi = self.stage_map.span(s)
            print(f"stage={s} layers={hi - lo}")
            for i in range(lo, hi):
                print(f"    {i:2d}: {_display_name(self._layer_specs[i])}")
        if self.loss_fn:
            print(f"  loss: {getattr(self.loss_fn, '__name__', type(self.loss_fn).__name__)}")

    def _seed(self, value: int) -> None:
        (self.seed_fn or ds_utils.set_random_seed)(value)

    def _materialise(self, idx: int) -> _Step:
        spec = self._layer_specs[idx]
        if self.seed_layers:
            self._seed(self.base_seed + idx)
        if isinstance(spec, PipelineModule):
            raise NotImplementedError("nested PipelineModule layers are not supported")
        if isinstance(spec, TiedLayerSpec):
            if spec.key not in self.tied_modules:
                self.tied_modules[spec.key] = spec.build()
                self._ties[spec.key] = _Tie(self.tied_modules[spec.key], spec.tied_weight_attr)
            mod = self.tied_modules[spec.key]
            fn = mod if spec.forward_fn is None else (lambda x, f=spec.forward_fn, m=mod: f(m, x))
            return _Step(idx, fn, mod, tie_key=spec.key)
        if isinstance(spec, LayerSpec):
            spec = spec.build()
        if isinstance(spec, nn.Module):
            self.add_module(str(idx), spec)            # state-dict keys are "<global layer index>.<param>"
            return _Step(idx, spec, spec)
        return _Step(idx, spec)                        # plain callable, e.g. a reshape between blocks

    @property
    def forward_funcs(self) -> List[Callable]:
        return [s.fn for s in self.stage]

    @property
    def tied_comms(self) -> Dict[str, dict]:
        return {k: {"ranks": t.ranks, "group": t.group, "weight_attr": t.attrs, "module": t.module}
                for k, t in self._ties.items() if t.group is not None}

    # --------------------------------------------------------------------------------- forward
    def set_checkpoint_interval(self, interval: int) -> None:
        """Every ``interval`` consecutive steps form one recomputed segment (0 = keep all activations)."""
        assert interval >= 0
        self.activation_checkpoint_interval = interval

    def _segment(self, lo: int, hi: int, pass_no: int) -> Callable:
        steps = self.stage[lo:hi]

        def run(*xs):
            x = xs[0] if len(xs) == 1 else xs
            for st in steps:
                self.curr_layer = st.index
                if self.seed_layers:
                    self._seed(self.base_seed * pass_no + st.index)
                x = st.fn(x)
            return x
        return run

    def _recompute_ok(self, lo: int, hi: int) -> bool:
        steps = self.stage[lo:hi]
        if type(self).__name__ in ("GPTModelPipe", "GPT2ModelPipe"):
            return all("ParallelTransformerLayerPipe" in type(s.fn).__name__ for s in steps)
        if self.checkpointable_layers is not None:
            return all(type(s.fn).__name__ in self.checkpointable_layers for s in steps)
        return any(s.module is not None and any(True for _ in s.module.parameters()) for s in steps)

    def forward(self, forward_input):
        self.micro_offset += 1
        pass_no = self.micro_offset + 1
        n, every = len(self.stage), self.activation_checkpoint_interval
        if every == 0:
            return self._segment(0, n, pass_no)(forward_input)
        x = forward_input
        for lo in range(0, n, every):
            hi = min(lo + every, n)
            args = x if isinstance(x, tuple) else (x,)
            seg = self._segment(lo, hi, pass_no)
            x = self.activation_checkpoint_func(seg, *args) if self._recompute_ok(lo, hi) else seg(*args)
        return x

    # ------------------------------------------------------------------------------ tied weights
    def _connect_ties(self) -> None:
        """One process group per (tie key, data-parallel index, model-slice index) over the stages that use the key.
        Every rank takes part in every ``new_group`` call (a torch.distributed requirement)."""
        if self.num_stages == 1:
            return
        uses: Dict[str, set] = {}
        for i, spec in enumerate(self._layer_specs):
            if isinstance(spec, TiedLayerSpec):
                uses.setdefault(spec.key, set()).add(self.stage_map.owner(i))
        slices = self._grid.get_slice_parallel_world_size()
        for key in sorted(uses):
            for dp in range(self._grid.data_parallel_size):
                for mp in range(slices):
                    coords = dict(data=dp, model=mp) if slices > 1 else dict(data=dp)
                    ranks = [self._grid.stage_to_global(stage_id=s, **coords) for s in sorted(uses[key])]
                    group = dist.new_group(ranks=ranks)
                    if self.global_rank in ranks:
                        tie = self._ties[key]
                        tie.ranks, tie.group = ranks, group
                        if self.global_rank != ranks[0]:
                            for w in tie.weights():
                                w.ds_pipe_replicated = True      # counted once (on the first stage) in norms

    def _synchronize_tied_weights(self) -> None:
        for tie in self._ties.values():
            if tie.group is not None:
                for w in tie.weights():
                    dist.broadcast(w.data, src=min(tie.ranks), group=tie.group)

    def get_tied_weights_and_groups(self):
        return [(w, t.group) for t in self._ties.values() if t.group is not None for w in t.weights()]

    def allreduce_tied_weight_gradients(self):
        for w, group in self.get_tied_weights_and_groups():
            if w.grad is not None:
                dist.all_reduce(w.grad, group=group)

    # ---------------------------------------------------------------------------------- queries
    def partitions(self):
        return self.parts

    def stage_owner(self, layer_idx):
        assert 0 <= layer_idx < self._num_layers
        return self.stage_map.owner(layer_idx)

    def topology(self):
        return self._topo

    def mpu(self):
        return self._grid

    def num_pipeline_stages(self):
        return self.num_stages

    def get_additional_losses(self):
        return None

    # -------------------------------------------------------------------------------- checkpoint
    def ckpt_prefix(self, checkpoints_path, tag):
        coord = self._topo.get_coord(rank=self.global_rank)
        name = "module" + "".join(f"-{ax}_{getattr(coord, ax):02d}" for ax in self._topo.get_axis_names() if ax != "data")
        return os.path.join(checkpoints_path, str(tag), name)

    def _layer_file_stem(self, ckpt_dir, local_layer_idx) -> str:
        return os.path.join(ckpt_dir, f"layer_{local_layer_idx + self._local_start:02d}")

    def ckpt_layer_path(self, ckpt_dir, local_layer_idx):
        slice_name = self._topo.get_rank_repr(rank=self.global_rank) or "model_00"
        return f"{self._layer_file_stem(ckpt_dir, local_layer_idx)}-{slice_name}-model_states.pt"

    def ckpt_layer_path_list(self, ckpt_dir, local_layer_idx):
        return sorted(glob.glob(self._layer_file_stem(ckpt_dir, local_layer_idx) + "-*model_states.pt"))

    def _layers_i_write(self) -> range:
        """Data-parallel replicas hold identical layers: replica 0 writes them all, unless parallel writing is on —
        then the replicas split the stage's layers between them."""
        dp_rank, dp_size = self._grid.data_parallel_id, self._grid.data_parallel_size
        n = len(self.stage)
        if getattr(self, "checkpoint_parallel_write_pipeline", False):
            cut = ds_utils.partition_uniform(n, dp_size)
            return range(cut[dp_rank], cut[dp_rank + 1])
        return range(n) if dp_rank == 0 else range(0)

    def save_state_dict(self, save_dir, checkpoint_engine, exclude_frozen_params=False):
        mine = self._layers_i_write()
        if len(mine) == 0:
            return
        os.makedirs(save_dir, exist_ok=True)
        for local_idx in mine:
            mod = self.stage[local_idx].module
            if mod is None:
                continue
            state = mod.state_dict()
            if exclude_frozen_params:
                frozen = {n for n, p in mod.named_parameters() if not p.requires_grad}
                state = {k: v for k, v in state.items() if k not in frozen}
            checkpoint_engine.save({k: v.clone() for k, v in state.items()}, self.ckpt_layer_path(save_dir, local_idx))

    def load_state_dir(self, load_dir, checkpoint_engine, strict=True):
        mp_rank = self._grid.get_slice_parallel_rank()
        for local_idx, st in enumerate(self.stage):
            if st.module is None:
                continue
            files = self.ckpt_layer_path_list(load_dir, local_idx)
            if files:
                st.module.load_state_dict(checkpoint_engine.load(files[min(mp_rank, len(files) - 1)], map_location="cpu"),
                                          strict=strict)
        self._synchronize_tied_weights()
