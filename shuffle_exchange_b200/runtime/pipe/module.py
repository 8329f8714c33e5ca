"""``PipelineModule`` — a model expressed as a flat list of layers, partitioned over pipeline stages.

Reference: deepspeed/runtime/pipe/module.py — ``LayerSpec`` :30, ``TiedLayerSpec`` :77,
``PipelineModule`` :86; partitioning ``uniform`` / ``parameters`` / ``type:regex`` :393-445; activation
checkpoint interval execution :340-391; tied-weight all-reduce :454-476; per-layer checkpoint files
``layer_XX-model_YY-model_states.pt`` :582-634.
"""
from __future__ import annotations

import glob
import os
import re as regex
from collections import defaultdict
from functools import partial

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


class LayerSpec:
    """Deferred construction of a layer: ``LayerSpec(nn.Linear, 8, 8)`` builds only on the owning stage."""

    def __init__(self, typename, *module_args, **module_kwargs):
        self.typename = typename
        self.module_args = module_args
        self.module_kwargs = module_kwargs
        if not issubclass(typename, nn.Module):
            raise RuntimeError("LayerSpec only supports torch.nn.Module types.")
        self.global_rank = dist.get_rank() if dist.is_initialized() else -1

    def __repr__(self):
        return ds_utils.call_to_str(self.typename.__name__, *self.module_args, **self.module_kwargs)

    def build(self, log=False):
        if log:
            logger.info(f"RANK={self.global_rank} building {repr(self)}")
        return self.typename(*self.module_args, **self.module_kwargs)


class TiedLayerSpec(LayerSpec):
    """A layer whose ``tied_weight_attr`` parameter(s) are shared by every stage that lists ``key``."""

    def __init__(self, key, typename, *module_args, forward_fn=None, tied_weight_attr=("weight",), **module_kwargs):
        super().__init__(typename, *module_args, **module_kwargs)
        self.key = key
        self.forward_fn = forward_fn
        self.tied_weight_attr = [tied_weight_attr] if isinstance(tied_weight_attr, str) else list(tied_weight_attr)


class PipelineModule(nn.Module):
    def __init__(self, layers, num_stages=None, topology=None, loss_fn=None, seed_layers=False, seed_fn=None,
                 base_seed=1234, partition_method="parameters", activation_checkpoint_interval=0,
                 activation_checkpoint_func=checkpointing.checkpoint, checkpointable_layers=None,
                 dynamic_shape=False):
        super().__init__()
        if num_stages is None and topology is None:
            raise RuntimeError("must provide num_stages or topology")
        self.micro_offset = 0
        self.loss_fn = loss_fn
        self.checkpointable_layers = checkpointable_layers
        if checkpointable_layers is not None:
            assert isinstance(checkpointable_layers, list), "param `checkpointable_layers` must be type of list."
        self.seed_layers = seed_layers
        self.seed_fn = seed_fn
        self.base_seed = base_seed
        self.dynamic_shape = dynamic_shape
        if dist.is_initialized() and dist.get_rank() == 0:
            try:
                logger.info(f"SEED_LAYERS={self.seed_layers} BASE_SEED={self.base_seed} SEED_FN={self.seed_fn}")
            except Exception:
                pass
        dist.init_distributed()
        self.world_group = dist.new_group(ranks=range(dist.get_world_size())) if dist.is_initialized() else None
        self.global_rank = dist.get_rank(group=self.world_group)
        self.world_size = dist.get_world_size(group=self.world_group)
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if topology:
            self._topo = topology
            self.num_stages = self._topo.get_dim("pipe")
        else:
            self.num_stages = num_stages
            if self.world_size % self.num_stages != 0:
                raise RuntimeError(f"num_stages ({self.num_stages}) must divide distributed world size ({self.world_size})")
            self._topo = PipeDataParallelTopology(num_pp=self.num_stages, num_dp=self.world_size // self.num_stages)
        self._grid = PipelineParallelGrid(process_group=self.world_group, topology=self._topo)
        self.stage_id = self._topo.get_coord(self.global_rank).pipe
        self._layer_specs = list(layers)
        self._num_layers = len(self._layer_specs)
        self._local_start = 0
        self._local_stop = None
        self._partition_layers(method=partition_method)
        self.forward_funcs = []
        self.fwd_map = {}
        self.tied_modules = nn.ModuleDict()
        self.tied_weight_attrs = {}
        self._build()
        self.to(get_accelerator().device_name(self.local_rank) if get_accelerator().is_available() else "cpu")
        self.tied_comms = self._index_tied_modules()
        self._synchronize_tied_weights()
        self.activation_checkpoint_interval = activation_checkpoint_interval
        self.activation_checkpoint_func = activation_checkpoint_func
        self.is_checkpointable_results = []
        self.is_checkpointable_results_interval = None

    # ------------------------------------------------------------------------------------ build
    def _build(self):
        specs = self._layer_specs
        for local_idx, layer in enumerate(specs[self._local_start:self._local_stop]):
            layer_idx = local_idx + self._local_start
            if self.seed_layers:
                if self.seed_fn:
                    self.seed_fn(self.base_seed + layer_idx)
                else:
                    ds_utils.set_random_seed(self.base_seed + layer_idx)
            if isinstance(layer, PipelineModule):
                raise NotImplementedError("RECURSIVE BUILD NOT YET IMPLEMENTED")
            elif isinstance(layer, nn.Module):
                name = str(layer_idx)
                self.forward_funcs.append(layer)
                self.fwd_map.update({name: len(self.forward_funcs) - 1})
                self.add_module(name, layer)
            elif isinstance(layer, TiedLayerSpec):
                if layer.key not in self.tied_modules:
                    self.tied_modules[layer.key] = layer.build()
                    self.tied_weight_attrs[layer.key] = layer.tied_weight_attr
                if layer.forward_fn is None:
                    self.forward_funcs.append(self.tied_modules[layer.key])
                else:
                    self.forward_funcs.append(partial(layer.forward_fn, self.tied_modules[layer.key]))
            elif isinstance(layer, LayerSpec):
                module = layer.build()
                name = str(layer_idx)
                self.forward_funcs.append(module)
                self.fwd_map.update({name: len(self.forward_funcs) - 1})
                self.add_module(name, module)
            else:
                self.forward_funcs.append(layer)     # plain callable (e.g. a lambda reshaping activations)
        for p in self.parameters():
            p.ds_pipe_replicated = False

    def _count_layer_params(self):
        counts = [0] * len(self._layer_specs)
        for idx, layer in enumerate(self._layer_specs):
            if isinstance(layer, LayerSpec):
                l = layer.build()
                counts[idx] = sum(p.numel() for p in l.parameters() if p.requires_grad)
            elif isinstance(layer, nn.Module):
                counts[idx] = sum(p.numel() for p in layer.parameters() if p.requires_grad)
        return counts

    def _find_layer_type(self, layername):
        idxs = []
        typeregex = regex.compile(layername, regex.IGNORECASE)
        for idx, layer in enumerate(self._layer_specs):
            name = None
            if isinstance(layer, LayerSpec):
                name = layer.typename.__name__
            elif isinstance(layer, nn.Module):
                name = layer.__class__.__name__
            else:
                try:
                    name = layer.__name__
                except AttributeError:
                    continue
            if typeregex.search(name):
                idxs.append(idx)
        if len(idxs) == 0:
            raise RuntimeError(f"Partitioning '{layername}' found no valid layers to partition.")
        return idxs

    def _partition_layers(self, method="uniform"):
        num_stages = self._topo.get_dim("pipe")
        stage_id = self._topo.get_coord(self.global_rank).pipe
        if self.global_rank == 0:
            logger.info(f"Partitioning pipeline stages with method {method}")
        method = method.lower()
        if method == "uniform":
            self.parts = ds_utils.partition_uniform(num_items=len(self._layer_specs), num_parts=num_stages)
        elif method == "parameters":
            self.parts = ds_utils.partition_balanced(weights=self._count_layer_params(), num_parts=num_stages)
        elif method.startswith("type:"):
            layertype = method.split(":")[1]
            binary_weights = [0] * len(self._layer_specs)
            for idx in self._find_layer_type(layertype):
                binary_weights[idx] = 1
            self.parts = ds_utils.partition_balanced(weights=binary_weights, num_parts=num_stages)
        elif method == "profile":
            raise NotImplementedError(f"Partitioning method {method} not implemented.")
        else:
            raise NotImplementedError(f"Partitioning method {method} not implemented.")
        if self.global_rank == 0:
            for stage in range(num_stages):
                start, stop = self.parts[stage], self.parts[stage + 1]
                print(f"stage={stage} layers={stop - start}")
                for idx, layer in enumerate(self._layer_specs[start:stop]):
                    name = str(layer)
                    if isinstance(layer, LayerSpec):
                        name = layer.typename.__name__
                    elif isinstance(layer, nn.Module):
                        name = layer.__class__.__name__
                    else:
                        name = getattr(layer, "__name__", name)
                    print(f"    {idx + start:2d}: {name}")
            if self.loss_fn:
                print(f"  loss: {getattr(self.loss_fn, '__name__', self.loss_fn.__class__.__name__)}")
        self._set_bounds(start=self.parts[stage_id], stop=self.parts[stage_id + 1])

    def _set_bounds(self, start=None, stop=None):
        self._local_start = start
        self._local_stop = stop

    # ---------------------------------------------------------------------------------- forward
    def set_checkpoint_interval(self, interval: int) -> None:
        """Recompute granularity: every ``interval`` consecutive layers form one checkpointed segment (0 = off)."""
        assert interval >= 0
        self.activation_checkpoint_interval = interval

    def forward(self, forward_input):
        self.micro_offset += 1

        def exec_range_func(start, end):
            local_micro_offset = self.micro_offset + 1

            def exec_func(*inputs):
                if len(inputs) == 1:
                    inputs = inputs[0]
                for idx, layer in enumerate(self.forward_funcs[start:end]):
                    self.curr_layer = idx + self._local_start
                    if self.seed_layers:
                        new_seed = (self.base_seed * local_micro_offset) + self.curr_layer
                        if self.seed_fn:
                            self.seed_fn(new_seed)
                        else:
                            ds_utils.set_random_seed(new_seed)
                    inputs = layer(inputs)
                return inputs

            return exec_func

        if self.activation_checkpoint_interval == 0:
            return exec_range_func(0, len(self.forward_funcs))(forward_input)
        num_layers = len(self.forward_funcs)
        x = forward_input
        for start_idx in range(0, num_layers, self.activation_checkpoint_interval):
            end_idx = min(start_idx + self.activation_checkpoint_interval, num_layers)
            funcs = self.forward_funcs[start_idx:end_idx]
            if not isinstance(x, tuple):
                x = (x,)
            if self._is_checkpointable(funcs):
                x = self.activation_checkpoint_func(exec_range_func(start_idx, end_idx), *x)
            else:
                x = exec_range_func(start_idx, end_idx)(*x)
        return x

    def _is_checkpointable(self, funcs):
        if self.__class__.__name__ in ("GPTModelPipe", "GPT2ModelPipe"):
            return all("ParallelTransformerLayerPipe" in f.__class__.__name__ for f in funcs)
        if self.checkpointable_layers is not None:
            return all(f.__class__.__name__ in self.checkpointable_layers for f in funcs)
        params = [f.parameters() for f in funcs if isinstance(f, torch.nn.Module)]
        return any(len(list(p)) > 0 for p in params)

    # ------------------------------------------------------------------------------ tied weights
    def allreduce_tied_weight_gradients(self):
        for key, comm in self.tied_comms.items():
            for attr_name in comm["weight_attr"]:
                weight = getattr(self.tied_modules[key], attr_name)
                if weight.grad is not None:
                    dist.all_reduce(weight.grad, group=comm["group"])

    def get_tied_weights_and_groups(self):
        out = []
        for key, comm in self.tied_comms.items():
            for attr_name in comm["weight_attr"]:
                out.append((getattr(self.tied_modules[key], attr_name), comm["group"]))
        return out

    def _synchronize_tied_weights(self):
        for key, comm in self.tied_comms.items():
            for attr_name in comm["weight_attr"]:
                dist.broadcast(getattr(comm["module"], attr_name).data, src=min(comm["ranks"]), group=comm["group"])

    def _index_tied_modules(self):
        tied_comms = {}
        if self._topo.get_dim("pipe") == 1:
            return tied_comms
        specs = self._layer_specs
        tie_keys = set(s.key for s in specs if isinstance(s, TiedLayerSpec))
        for key in sorted(tie_keys):
            tied_layers = [idx for idx, layer in enumerate(specs) if isinstance(layer, TiedLayerSpec) and layer.key == key]
            tied_stages = set(self.stage_owner(idx) for idx in tied_layers)
            for dp in range(self._grid.data_parallel_size):
                for mp in range(self._grid.get_slice_parallel_world_size()):
                    tied_ranks = []
                    for s in sorted(tied_stages):
                        if self._grid.get_slice_parallel_world_size() > 1:
                            tied_ranks.append(self._grid.stage_to_global(stage_id=s, data=dp, model=mp))
                        else:
                            tied_ranks.append(self._grid.stage_to_global(stage_id=s, data=dp))
                    group = dist.new_group(ranks=tied_ranks)
                    if self.global_rank in tied_ranks:
                        assert key in self.tied_modules
                        tied_comms[key] = {"ranks": tied_ranks, "group": group,
                                           "weight_attr": self.tied_weight_attrs[key], "module": self.tied_modules[key]}
                        if self.global_rank != tied_ranks[0]:
                            for attr in self.tied_weight_attrs[key]:
                                for p in [getattr(self.tied_modules[key], attr)]:
                                    p.ds_pipe_replicated = True
        return tied_comms

    # ---------------------------------------------------------------------------------- queries
    def partitions(self): return self.parts

    def stage_owner(self, layer_idx):
        assert 0 <= layer_idx < self._num_layers
        for stage in range(self._topo.get_dim("pipe")):
            if self.parts[stage] <= layer_idx < self.parts[stage + 1]:
                return stage
        raise RuntimeError(f"Layer {layer_idx} not owned? parts={self.parts}")

    def topology(self): return self._topo
    def mpu(self): return self._grid
    def num_pipeline_stages(self): return self._topo.get_dim("pipe")

    def get_additional_losses(self):
        return None

    # -------------------------------------------------------------------------------- checkpoint
    def ckpt_prefix(self, checkpoints_path, tag):
        rank_name = "module"
        omit_dims = frozenset(["data"])
        axes = [a for a in self._grid._topo.get_axis_names() if a not in omit_dims]
        for dim in axes:
            rank = getattr(self._grid._topo.get_coord(rank=self.global_rank), dim)
            rank_name += f"-{dim}_{rank:02d}"
        return os.path.join(checkpoints_path, str(tag), rank_name)

    def ckpt_layer_path(self, ckpt_dir, local_layer_idx):
        idx = local_layer_idx + self._local_start
        layer_ckpt_path = os.path.join(ckpt_dir, f"layer_{idx:02d}")
        rank_repr = self._grid._topo.get_rank_repr(rank=self.global_rank)
        if rank_repr != "":
            layer_ckpt_path += f"-{rank_repr}"
        else:
            layer_ckpt_path += "-model_00"
        return layer_ckpt_path + "-model_states.pt"

    def ckpt_layer_path_list(self, ckpt_dir, local_layer_idx):
        idx = local_layer_idx + self._local_start
        files = glob.glob(os.path.join(ckpt_dir, f"layer_{idx:02d}-") + "*model_states.pt")
        files.sort()
        return files

    def save_state_dict(self, save_dir, checkpoint_engine, exclude_frozen_params=False):
        dp_rank, dp_size = self._grid.data_parallel_id, self._grid.data_parallel_size
        num_layers = len(self.forward_funcs)
        if getattr(self, "checkpoint_parallel_write_pipeline", False):
            offsets = ds_utils.partition_uniform(num_layers, dp_size)
            start, end = offsets[dp_rank], offsets[dp_rank + 1]
        else:
            if dp_rank != 0:
                return
            start, end = 0, num_layers
        os.makedirs(save_dir, exist_ok=True)
        for idx, layer in list(enumerate(self.forward_funcs))[start:end]:
            if not hasattr(layer, "state_dict"):
                continue
            sd = layer.state_dict()
            if exclude_frozen_params:
                frozen = {n for n, p in layer.named_parameters() if not p.requires_grad}
                sd = {k: v for k, v in sd.items() if k not in frozen}
            checkpoint_engine.save({k: v.clone() for k, v in sd.items()}, self.ckpt_layer_path(save_dir, idx))

    def load_state_dir(self, load_dir, checkpoint_engine, strict=True):
        for idx, layer in enumerate(self.forward_funcs):
            if not hasattr(layer, "load_state_dict"):
                continue
            files = self.ckpt_layer_path_list(load_dir, idx)
            if not files:
                continue
            mp_rank = self._grid.get_slice_parallel_rank()
            path = files[min(mp_rank, len(files) - 1)]
            layer.load_state_dict(checkpoint_engine.load(path, map_location="cpu"), strict=strict)
        self._synchronize_tied_weights()
