"""ZeRO stage 3 — parameters, gradients and optimizer state sharded over the data-parallel group.

What the reference spreads over ``stage3.py`` (3263 lines), ``partitioned_param_coordinator.py``,
``parameter_offload.py`` and most of ``partition_parameters.py`` is organised here around a static
notion of **unit**:

* a unit is a module whose parameters are fetched and released together (by default every child
  of a ``nn.ModuleList`` — i.e. one transformer block — plus one *root* unit for what is left:
  embeddings, final norm, LM head; ``set_z3_leaf_modules`` / ``unit_modules=`` override this);
* each (unit, optimizer param group) pair is one :class:`FlatGroup` whose flat low-precision
  buffer is cut into ``dp`` shards.  Fetch = ONE all-gather per unit (coalesced over its flat
  groups), release = dropping the gathered buffer; gradient reduction = ONE reduce-scatter per
  unit launched when the unit's last gradient lands, followed — when no global clipping / overflow
  check forces a deferred step — by the fused Adam update of that unit's shard *inside backward*
  (kernel K1 of SURVEY.md §8: reduce-scatter ⊕ scale ⊕ Adam ⊕ bf16 write-back, overlapped with the
  rest of backward);
* the fetch order is recorded on the first iteration and replayed as a prefetch schedule on a
  side stream (``prefetch_bucket_size`` / ``max_live_parameters`` bound the look-ahead), forward and
  backward, including the recompute pass of activation checkpointing (no double gather).

Reference anchors: sub-groups & defragmented partitions ``stage3.py:720-983``; grad hooks and IPG
buckets ``:1175-1417``; ``partition_grads`` ``:1478-1540``; per-sub-group step ``:2112-2174``;
fetch/release coordinator ``partitioned_param_coordinator.py:285-559``; module hooks
``parameter_offload.py:282-510``; state dict keys ``stage3.py:2656-2669``.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, List, Optional

import os

import torch
import torch.nn as nn

from ...comm import comm as dist
from ...device import get_accelerator
from ...ops import functional as _F
from ...utils.logging import logger
from ...version import __version__
from .flat_group import FlatGroup
from .sharded_optimizer import (CLIP_GRAD, DS_VERSION, LOSS_SCALER, PARTITION_COUNT, ZERO_STAGE, ShardedOptimizerBase)

FP32_FLAT_GROUPS = "fp32_flat_groups"
OPTIMIZER_STATE_DICT = "optimizer_state_dict"


# Look-ahead (default on; SXB_Z3_LOOKAHEAD=0 restores the strictly bounded behaviour): the unit that follows the
# current one is always fetched ahead of time — one unit of look-ahead is what hides a gather behind compute — and
# gathers issued inside backward skip the cross-rank shard barrier.  The skip is safe because a peer rewrites its
# shard of unit u only in the fused step that follows the reduction of u, which cannot start before EVERY rank
# announced its gradients of u, i.e. after every rank's gather of u (stream-ordered before its backward of u)
# completed.  The root unit (embedding + LM head) is live for the whole iteration and is not charged against
# ``stage3_max_live_parameters``: the limit bounds what is fetched ahead of use.
_Z3_LOOKAHEAD = os.environ.get("SXB_Z3_LOOKAHEAD", "1") == "1"
_Z3_GATED = os.environ.get("SXB_Z3_GATED", "1") == "1"

class Unit:
    """A module whose parameters are gathered / released / reduced / stepped together."""

    def __init__(self, name: str, module: Optional[nn.Module], index: int):
        self.name = name
        self.module = module
        self.index = index
        self.groups: List[FlatGroup] = []
        self.persistent = False
        self.available = False
        self.fetch_event = None
        self.grads_ready = 0
        self.n_trainable = 0
        self.reduced = False
        self.stepped = False
        self.numel = 0
        self.live_users = 0

    def __repr__(self):
        return f"Unit({self.index}:{self.name}, numel={self.numel:,}, groups={len(self.groups)})"


def discover_units(module: nn.Module, unit_modules: Optional[List[nn.Module]] = None) -> List[nn.Module]:
    """Outermost children of ``nn.ModuleList`` containers (transformer blocks), or modules flagged as
    ZeRO-3 leaf modules (``utils/z3_leaf_module.py``), in registration order."""
    if unit_modules:
        return list(unit_modules)
    found: List[nn.Module] = []

    def walk(m: nn.Module):
        if getattr(m, "_z3_leaf", False) and m is not module:
            found.append(m)
            return
        if isinstance(m, (nn.ModuleList, nn.Sequential)) and m is not module and len(m) > 1 \
                and all(any(True for _ in c.parameters()) for c in m):
            found.extend(list(m))
            return
        for c in m.children():
            walk(c)

    walk(module)
    return found


class DeepSpeedZeroOptimizer_Stage3(ShardedOptimizerBase):
    zero_stage = 3

    def __init__(self, module, init_optimizer, param_names=None, timers=None, ds_config=None, static_loss_scale=1.0,
                 dynamic_loss_scale=False, dynamic_loss_args=None, verbose=True, contiguous_gradients=True,
                 reduce_bucket_size=500000000, prefetch_bucket_size=50000000, max_reuse_distance=1000000000,
                 max_live_parameters=1000000000, param_persistence_threshold=100000,
                 model_persistence_threshold=None, dp_process_group=None, reduce_scatter=True, overlap_comm=True,
                 offload_optimizer_config=None, offload_param_config=None, sub_group_size=1000000000000,
                 offload_ratio=0.0, mpu=None, clip_grad=0.0, gradient_accumulation_dtype=torch.float32,
                 communication_data_type=torch.float16, postscale_gradients=True, gradient_predivide_factor=1.0,
                 gradient_accumulation_steps=1, elastic_checkpoint=False, aio_config=None, all2all_process_group=None,
                 zero_hpz_partition_size=1, zero_quantized_weights=False, zero_quantized_nontrainable_weights=False,
                 zero_module_granularity_threshold=0, zeropp_loco_param=None, zero_quantized_gradients=False, log_trace_cache_warnings=False,
                 check_grad_overflow=None, dtype=None, fused_step=True, unit_modules=None, fused_collectives=None):
        first = next(p for g in init_optimizer.param_groups for p in g["params"])
        lp_dtype = dtype or (first.dtype if first.dtype != torch.float32 or getattr(first, "ds_tensor", None) is None
                             else first.ds_tensor.dtype)
        super().__init__(init_optimizer, timers=timers, dtype=lp_dtype, static_loss_scale=static_loss_scale,
                         dynamic_loss_scale=dynamic_loss_scale, dynamic_loss_args=dynamic_loss_args, clip_grad=clip_grad,
                         dp_process_group=dp_process_group, mpu=mpu,
                         gradient_accumulation_steps=gradient_accumulation_steps,
                         gradient_accumulation_dtype=gradient_accumulation_dtype,
                         communication_data_type=communication_data_type,
                         gradient_predivide_factor=gradient_predivide_factor, postscale_gradients=postscale_gradients,
                         check_grad_overflow=check_grad_overflow)
        self.module = module
        self.param_names = param_names or {p: n for n, p in module.named_parameters()}
        self.world = dist.get_world_size(dp_process_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(dp_process_group) if dist.is_initialized() else 0
        self.partition_count = self.world
        self.prefetch_bucket_size = int(prefetch_bucket_size)
        self.max_live_parameters = int(max_live_parameters)
        self.max_reuse_distance = int(max_reuse_distance)
        self.param_persistence_threshold = int(param_persistence_threshold)
        self.overlap_comm = bool(overlap_comm) and get_accelerator().is_available() and self.world > 1
        self.offload_optimizer = bool(offload_optimizer_config is not None and
                                      str(getattr(offload_optimizer_config, "device", "none")).endswith("cpu"))
        self.offload_optimizer_pin = bool(self.offload_optimizer and getattr(offload_optimizer_config, "pin_memory", False))
        self.offload_param = bool(offload_param_config is not None and
                                  str(getattr(offload_param_config, "device", "none")).endswith("cpu"))
        self.offload_param_pin = bool(self.offload_param and getattr(offload_param_config, "pin_memory", False))
        # ---- ZeRO-Infinity / ZeRO++ switches (every accepted key either acts or raises — none is dropped silently)
        opt_dev = str(getattr(offload_optimizer_config, "device", "none")) if offload_optimizer_config is not None else "none"
        self.offload_optimizer_nvme = opt_dev.endswith("nvme")
        if self.offload_optimizer_nvme:
            self.offload_optimizer = True          # masters + optimizer state in host buffers between swaps
            if getattr(offload_optimizer_config, "nvme_path", None) is None:
                raise ValueError("offload_optimizer.device = nvme needs offload_optimizer.nvme_path")
        self._offload_optimizer_config, self._aio_config = offload_optimizer_config, aio_config
        self.offload_ratio = float(getattr(offload_optimizer_config, "ratio", 1.0)) if self.offload_optimizer else 0.0
        self._param_store = None
        if offload_param_config is not None and str(getattr(offload_param_config, "device", "none")).endswith("nvme"):
            # ZeRO-Infinity parameter offload: shards live in files, at most buffer_count of them resident in host memory
            nvme_path = getattr(offload_param_config, "nvme_path", None)
            if nvme_path is None:
                raise ValueError("offload_param.device = nvme needs offload_param.nvme_path")
            from .nvme_shards import NvmeShardStore
            aio = dict(aio_config) if isinstance(aio_config, dict) else {
                k: getattr(aio_config, k) for k in ("block_size", "queue_depth", "single_submit", "overlap_events",
                                                    "intra_op_parallelism") if hasattr(aio_config, k)}
            self.offload_param = True
            self.offload_param_pin = bool(getattr(offload_param_config, "pin_memory", False))
            self._param_store = NvmeShardStore(
                os.path.join(str(nvme_path), "zero_stage_3", f"{str(self.dtype).split('.')[-1]}params", f"rank{self.rank}"),
                self.dtype, buffer_count=int(getattr(offload_param_config, "buffer_count", 5) or 5), aio_config=aio,
                pin_memory=self.offload_param_pin, max_in_cpu=int(getattr(offload_param_config, "max_in_cpu", 0) or 0))
        self.hpz_size = int(zero_hpz_partition_size or 1)
        self.quantized_weights = bool(zero_quantized_weights)
        self.quantized_gradients = bool(zero_quantized_gradients)
        self.loco_param = zeropp_loco_param
        if zero_quantized_nontrainable_weights:
            raise NotImplementedError("zero_quantized_nontrainable_weights: frozen weights are gathered like trainable "
                                      "ones; use zero_quantized_weights")
        if int(sub_group_size) < 10 ** 12 and dist.get_rank() == 0:
            logger.info("ZeRO-3: sub_group_size is accepted for compatibility; the sub-groups of this design are the "
                        "parameter units (one flat group per transformer block)")
        self.elastic_checkpoint = elastic_checkpoint
        self._is_boundary = True
        self.fused_step_enabled = bool(fused_step)
        self.allgather_stream = get_accelerator().Stream() if self.overlap_comm else None
        self.reduce_stream = get_accelerator().Stream() if self.overlap_comm else None
        self._pending_reduce: List[Unit] = []
        self._live_numel = 0
        self._trace: List[int] = []
        self._trace_complete = False
        self._trace_pos = 0
        self._in_backward = False
        self._hooks = []

        self.fused = None
        self._build_units(unit_modules)
        from .fused_collectives import maybe_create
        self.fused = maybe_create(self, dp_process_group, fused_collectives)
        self._install_hooks()
        from ...utils.tensor_fragment import link_group_params
        for fg in self.groups:
            link_group_params(fg, self)
        self.fp32_partitioned_groups_flat = [g.hp_param for g in self.groups]
        if verbose and dist.get_rank() == 0:
            tot = sum(u.numel for u in self.units)
            logger.info(f"ZeRO stage 3: {len(self.units)} unit(s), {len(self.groups)} flat group(s), {tot:,} parameters, "
                        f"dp={self.world}, fused_step={self._fused_step_active()}")

    # ------------------------------------------------------------------------------ construction
    def _build_units(self, unit_modules):
        pg_of: Dict[int, int] = {}
        for gi, pg in enumerate(self.optimizer.param_groups):
            for p in pg["params"]:
                pg_of[id(p)] = gi
        unit_mods = discover_units(self.module, unit_modules)
        self._total_numel = sum(getattr(p, "ds_numel", p.numel()) for p in self.module.parameters() if id(p) in pg_of)
        claimed = set()
        self.units: List[Unit] = []
        for um in unit_mods:
            u = Unit(self._module_name(um), um, len(self.units))
            u._params = [p for p in um.parameters() if id(p) in pg_of and id(p) not in claimed]
            claimed.update(id(p) for p in u._params)
            if u._params:
                self.units.append(u)
        root = Unit("<root>", self.module, len(self.units))
        root._params = [p for p in self.module.parameters() if id(p) in pg_of and id(p) not in claimed]
        if root._params:
            self.units.append(root)
        self.root_unit = root if root._params else None
        # one flat group per (unit, param group); rewrite the base optimizer's groups accordingly
        new_param_groups = []
        keep_full = self.world == 1
        for u in self.units:
            by_pg: "OrderedDict[int, list]" = OrderedDict()
            for p in u._params:
                by_pg.setdefault(pg_of[id(p)], []).append(p)
            for gi, params in by_pg.items():
                params = self._materialize_pre_partitioned(params)
                names = [self.param_names.get(p, f"unit{u.index}.p{j}") for j, p in enumerate(params)]
                # partial optimizer offload (offload_optimizer.ratio): the first `ratio` of the parameters (in unit
                # order) keep master + state on the host, the rest step on the GPU (reference stage3.py:258-272)
                seen = sum(g.unpadded_numel for g in self.groups)
                on_cpu = self.offload_optimizer and (self.offload_ratio >= 1.0 or seen < self.offload_ratio * self._total_numel)
                fg = FlatGroup(params, names, self.world, self.rank, self.dp_process_group, self.device, self.dtype,
                               grad_dtype=self.dtype, group_index=len(self.groups), keep_full_params=keep_full,
                               hp_on_cpu=on_cpu, pin=self.offload_optimizer_pin or self.offload_param_pin,
                               lp_on_cpu=self.offload_param and not keep_full,
                               lp_store=None if keep_full else self._param_store)
                self._configure_zeropp(fg)
                fg.lazy_grad_zero = os.environ.get("SXB_LAZY_GRAD_ZERO", "1") != "0"
                src_pg = self.optimizer.param_groups[gi]
                pg = {k: v for k, v in src_pg.items() if k != "params"}
                hp = self._register_group(fg, pg)
                pg["params"] = [hp]
                fg.unit = u
                fg.source_param_group = gi
                new_param_groups.append(pg)
                u.groups.append(fg)
                u.numel += fg.unpadded_numel
                u.n_trainable += fg.n_trainable
            u.persistent = keep_full or (u.numel <= self.param_persistence_threshold)
            u.available = keep_full
            if u.persistent and not keep_full:
                for fg in u.groups:
                    fg.keep_full_params_backup = False
                    fg.gather_params()
                u.available = True
            del u._params
        self.optimizer.param_groups = new_param_groups
        self.optimizer.state.clear()

    def _configure_zeropp(self, fg: FlatGroup):
        """qwZ / hpZ / qgZ on the NCCL collectives of this flat group (the fused symmetric-memory path is not used with
        them: ``fused_collectives.maybe_create``)."""
        if self.world <= 1:
            return
        fg.quantized_weights = self.quantized_weights
        if self.hpz_size > 1:
            from ...utils import groups as _groups
            if not _groups._zero_param_parallel_is_initialized():
                _groups._create_zero_param_parallel_group(self.hpz_size)
            fg.hpz_group = _groups._get_zero_param_intra_parallel_group()
        if self.quantized_gradients:
            from ...utils import groups as _groups
            fg.quantized_gradients = (_groups._get_local_all_to_all_group(), self.loco_param)

    def _swapper(self):
        """NVMe swapper for masters + optimizer state (offload_optimizer.device = nvme), created on first use."""
        sw = getattr(self, "optimizer_swapper", None)
        if sw is None and self.offload_optimizer_nvme:
            from ..swap_tensor import PartitionedOptimizerSwapper
            from ..swap_tensor.aio_config import get_aio_config
            aio = self._aio_config if isinstance(self._aio_config, dict) else get_aio_config({})
            cfg = self._offload_optimizer_config
            largest = max(fg.part for fg in self.groups)
            sw = self.optimizer_swapper = PartitionedOptimizerSwapper(cfg, aio, str(cfg.nvme_path), self.optimizer, largest,
                                                                      "cpu", torch.float32, None)
            self._swapped = set()
        return sw

    def _swap_in(self, fg: FlatGroup):
        sw = self._swapper()
        if sw is not None and id(fg) in self._swapped:
            sw.swap_in_optimizer_state(fg.hp_param)
            self._swapped.discard(id(fg))

    def _swap_out(self, fg: FlatGroup):
        sw = self._swapper()
        if sw is not None and fg.hp_on_cpu and id(fg) not in self._swapped:
            sw.swap_out_optimizer_state(fg.hp_param)
            self._swapped.add(id(fg))

    def swap_in_all(self):
        for fg in self.groups:
            self._swap_in(fg)

    def swap_out_all(self):
        for fg in self.groups:
            self._swap_out(fg)

    def _step_group(self, fg: FlatGroup, grad_scale: float, write_lp: bool = True):
        self._swap_in(fg)
        super()._step_group(fg, grad_scale, write_lp)
        self._swap_out(fg)

    def _materialize_pre_partitioned(self, params):
        """Parameters created under ``zero.Init`` hold only ``ds_tensor`` (their own 1/world slice);
        rebuild the full tensor once so the unit-flat layout can be formed (peak = one unit)."""
        for p in params:
            shard = getattr(p, "ds_tensor", None)
            if shard is None or p.numel() == getattr(p, "ds_numel", p.numel()):
                continue
            n = p.ds_numel
            per = shard.numel()
            full = torch.empty(per * self.world, dtype=shard.dtype, device=self.device)
            if self.world > 1:
                dist.all_gather_into_tensor(full, shard.to(self.device), group=self.dp_process_group)
            else:
                full.copy_(shard)
            p.data = full[:n].view(p.ds_shape)
            p.ds_tensor = None
        return params

    def _module_name(self, m):
        for n, mod in self.module.named_modules():
            if mod is m:
                return n or "<root>"
        return type(m).__name__

    def _fused_step_active(self) -> bool:
        return (self.fused_step_enabled and self._fused and self.clip_grad <= 0.0 and not self.check_overflow_enabled
                and self.loss_scale == 1.0)

    # ------------------------------------------------------------------------------------ hooks
    def _install_hooks(self):
        for u in self.units:
            if u.module is self.module or u is self.root_unit:
                continue
            self._hooks.append(u.module.register_forward_pre_hook(lambda m, a, u=u: self._pre_forward(u)))
            self._hooks.append(u.module.register_forward_hook(lambda m, a, o, u=u: self._post_forward(u)))
            self._hooks.append(u.module.register_full_backward_pre_hook(lambda m, g, u=u: self._pre_backward_unit(u)))
        self._hooks.append(self.module.register_forward_pre_hook(lambda m, a: self._pre_forward_root()))
        for u in self.units:
            for fg in u.groups:
                self._install_ready_hooks(u, fg)

    def _install_ready_hooks(self, u: Unit, fg: FlatGroup):
        def on_ready(param, u=u):
            # one call per USE of the parameter (a tied embedding / LM-head weight delivers twice): the unit is
            # complete when every distinct parameter received all the contributions its forward uses announced
            if u.reduced:
                raise RuntimeError(
                    f"ZeRO-3: a gradient contribution for a parameter of {u} arrived after the unit's reduction was "
                    "launched (a parameter used by both main-grad ops and plain autograd ops, or a forward that "
                    "ran outside the counted window); its gradient would be dropped")
            param._sxb_calls = getattr(param, "_sxb_calls", 0) + 1
            if param._sxb_calls != _F.expected_uses(param):
                return
            u.grads_ready += 1
            if u.grads_ready >= u.n_trainable:
                self._unit_backward_done(u)

        for i, p in enumerate(fg.params):
            if not fg.trainable[i]:
                continue
            p._sxb_grad_ready = on_ready

            def autograd_hook(param, fg=fg, idx=i):
                if param.grad is None:
                    return
                gv = fg.grad_view(idx)
                if getattr(param, "_sxb_grad_written", False):
                    gv.add_(param.grad.reshape(gv.shape).to(gv.dtype))
                else:
                    gv.copy_(param.grad.reshape(gv.shape))
                param._sxb_grad_written = True
                param.grad = None
                on_ready(param)

            self._hooks.append(p.register_post_accumulate_grad_hook(autograd_hook))

    # ---------------------------------------------------------------------------- fetch / release
    def _gather_unit(self, u: Unit, async_on_side_stream: bool):
        if u.available or u.fetch_event is not None:
            return
        if self.fused is not None and self.fused._shards_dirty and (
                not _Z3_LOOKAHEAD or not self._in_backward or u.stepped):
            # first gather after an optimizer step: wait for my own shard updates, then one cross-rank barrier.
            # (With SXB_Z3_LOOKAHEAD=1 gathers inside backward skip it: a unit's shard is only rewritten by the fused
            # step that follows ITS backward on every rank, i.e. after every rank gathered it.)
            side = self.allgather_stream if self.allgather_stream is not None else get_accelerator().current_stream()
            if self.reduce_stream is not None:
                side.wait_stream(self.reduce_stream)
            side.wait_stream(get_accelerator().current_stream())
            with get_accelerator().stream(side):
                self.fused.shards_ready_barrier()
            if side is not get_accelerator().current_stream() and not async_on_side_stream:
                get_accelerator().current_stream().wait_stream(side)
        gated = self._unit_is_gated(u)
        if (async_on_side_stream or gated) and self.allgather_stream is not None:
            self.allgather_stream.wait_stream(get_accelerator().current_stream())
            with get_accelerator().stream(self.allgather_stream):
                small = [self._gather_fg(fg, gated) for fg in u.groups]
                ev = get_accelerator().Event()
                ev.record(self.allgather_stream)
            u.fetch_event = ev
            # gated units: the compute stream only waits for the non-matrix parameters; the GEMMs gate themselves
            u.small_events = [e for e in small if e is not None] if gated else None
        else:
            for fg in u.groups:
                self._gather_fg(fg, False)
            u.available = True
        if u is not self.root_unit:
            self._live_numel += u.numel

    def _unit_is_gated(self, u: Unit) -> bool:
        """K2 applies to units whose module declares that all of its matrix parameters are consumed through
        ``ops.functional.linear`` (``_sxb_linear_only_weights``), on the fused copy/pull gather.  SXB_Z3_GATED=0
        turns it off."""
        return (_Z3_GATED and self.fused is not None and self.allgather_stream is not None and u.module is not None
                and getattr(u.module, "_sxb_linear_only_weights", False)
                and all((fg.part * 2) % 256 == 0 for fg in u.groups))     # flag chunks are multiples of 256 bytes

    def _gather_fg(self, fg: FlatGroup, gated: bool = False):
        if self.fused is not None and getattr(fg, "_symm_shard", None) is not None:
            return self.fused.gather(fg, gated)
        fg.gather_params(backward=self._in_backward)
        return None

    def _wait_unit(self, u: Unit, full: bool = True):
        """``full=False`` (module hooks of a gated unit): wait only for the non-matrix parameters; the completion
        event of the whole gather is kept in ``u.full_event`` and honoured before the buffer is released or handed
        to anything but a gated GEMM."""
        cur = get_accelerator().current_stream()
        if u.fetch_event is not None:
            small = getattr(u, "small_events", None)
            if small is not None and not full:
                for e in small:
                    cur.wait_event(e)
                u.full_event = u.fetch_event
            else:
                cur.wait_event(u.fetch_event)
                u.full_event = None
            for fg in u.groups:
                if fg.flat_lp is not None:
                    fg.flat_lp.record_stream(cur)
            u.fetch_event = None
            u.small_events = None
            u.available = True
        elif full and getattr(u, "full_event", None) is not None:
            cur.wait_event(u.full_event)
            u.full_event = None

    def _fetch(self, u: Unit, full: bool = True):
        if u.persistent:
            return
        if not u.available and u.fetch_event is None:
            self._gather_unit(u, async_on_side_stream=False)
        self._wait_unit(u, full)

    def _release(self, u: Unit):
        if u.persistent or not u.available:
            return
        if getattr(u, "full_event", None) is not None:
            # every copy of the gather must have landed before the buffer goes back to the allocator
            get_accelerator().current_stream().wait_event(u.full_event)
            u.full_event = None
        for fg in u.groups:
            fg.release_params()
        u.available = False
        if u is not self.root_unit:
            self._live_numel -= u.numel

    def _prefetch_after(self, u: Unit, forward: bool):
        """Issue gathers for the units that follow ``u`` in the recorded order."""
        if not self._trace_complete or self.world == 1:
            return
        table = getattr(self, "_compiled_prefetch", None)
        if table is not None and (u.index, forward) in table:
            # schedule produced by the schedule compiler (compile/backend.py: apply_plan): explicit gather list
            for nxt in table[(u.index, forward)]:
                nu = self.units[nxt]
                if nu.persistent or nu.available or nu.fetch_event is not None:
                    continue
                if self._live_numel + nu.numel > self.max_live_parameters:
                    break
                self._gather_unit(nu, async_on_side_stream=True)
            return
        order = self._trace if forward else list(reversed(self._trace))
        try:
            pos = order.index(u.index)
        except ValueError:
            return
        budget = self.prefetch_bucket_size
        issued = 0
        for nxt in order[pos + 1:]:
            nu = self.units[nxt]
            if nu.persistent or nu.available:
                continue
            if nu.fetch_event is not None:
                issued += 1                  # already in flight: counts as look-ahead
                continue
            # The next unit is always fetched ahead (one unit of look-ahead is what hides the gather behind the current
            # unit's compute; the root unit alone — embedding + head, live for the whole iteration — can exceed
            # max_live_parameters); the limits bound everything beyond it.
            if (issued > 0 or not _Z3_LOOKAHEAD) and (
                    budget <= 0 or self._live_numel + nu.numel > self.max_live_parameters):
                break
            self._gather_unit(nu, async_on_side_stream=True)
            issued += 1
            budget -= nu.numel
            if budget <= 0:
                break

    # ----------------------------------------------------------------------------- module hooks
    def _pre_forward_root(self):
        if self.root_unit is not None:
            self._fetch(self.root_unit)
        if not torch.is_grad_enabled() or not self.module.training:
            return

    def _pre_forward(self, u: Unit):
        if not self._trace_complete and not self._in_backward:
            self._trace.append(u.index)
        self._fetch(u, full=False)
        self._prefetch_after(u, forward=not self._in_backward)

    def _post_forward(self, u: Unit):
        if self._in_backward:
            return      # recompute pass of activation checkpointing: backward of this unit follows
        self._release(u)

    def _pre_backward_unit(self, u: Unit):
        self._fetch(u, full=False)
        self._attach_unit(u)
        self._prefetch_after(u, forward=False)

    def _attach_unit(self, u: Unit):
        """Allocate the unit's flat gradient buffer and point its parameters' main-grad views at it —
        done lazily, right before the unit's backward, so only ~one unit of gradients is alive."""
        if getattr(u, "attached", False):
            return
        u.attached = True
        if self.fused is not None:
            if not hasattr(self, "_arena_free_events"):
                self._arena_free_events = {}
            u._arena_key = self.fused.arena_key(u)
            ev = self._arena_free_events.get(u._arena_key)
            if ev is not None:
                get_accelerator().current_stream().wait_event(ev)   # my own reduce of this arena's last occupant
            self.fused.attach_unit_grads(u)
            return
        for fg in u.groups:
            if (fg.grad_shard is not None and fg.flat_grad is not None
                    and fg.grad_shard.data_ptr() == fg.flat_grad.data_ptr()):
                fg.grad_shard = fg.grad_shard.clone()   # un-alias before the buffer is overwritten
            fg.attach_main_grads()
            fg.mark_grads_unwritten()

    # ---------------------------------------------------------------------------------- backward
    def set_gradient_accumulation_boundary(self, is_boundary: bool):
        self._is_boundary = is_boundary

    def _pre_backward(self):
        self._in_backward = True
        _F.close_use_window()
        if self._trace and not self._trace_complete:
            self._trace_complete = True
        for u in self.units:
            u.grads_ready = 0
            u.reduced = False
            u.stepped = False
            u.attached = False
            for fg in u.groups:
                fg.mark_grads_unwritten()
                for p in fg.params:
                    p._sxb_calls = 0
            if u is self.root_unit or u.module is self.module:
                # no module-level backward hook for the root unit: make sure it is gathered (a pipeline stage runs the
                # backward of an earlier micro-batch after another micro-batch's backward released it) and attached
                self._fetch(u)
                self._attach_unit(u)

    def _post_backward(self):
        # units whose gradients never (all) arrived — unused parameters — are flushed here
        for u in self.units:
            if not u.reduced and u.n_trainable > 0:
                self._unit_backward_done(u, force=True)
        if self.reduce_stream is not None and self._pending_reduce:
            get_accelerator().current_stream().wait_stream(self.reduce_stream)
            self._pending_reduce = []
        for u in self.units:
            self._release(u)
            for fg in u.groups:
                fg.micro_steps_accumulated += 1
        self._in_backward = False
        _F.open_use_window()

    def _unit_backward_done(self, u: Unit, force: bool = False):
        if u.reduced:
            return
        u.reduced = True
        accumulate = any(fg.micro_steps_accumulated > 0 for fg in u.groups)
        fuse = self._is_boundary and self._fused_step_active()
        stream = self.reduce_stream
        if self.fused is not None:
            if not getattr(u, "attached", False):
                self._attach_unit(u)
            # announce on the compute stream, wait + reduce on the reduce stream
            for fg in u.groups:
                fg.zero_unwritten_grads()
            epochs = [self.fused.announce(fg) for fg in u.groups]
            if stream is not None:
                stream.wait_stream(get_accelerator().current_stream())
            with get_accelerator().stream(stream):
                for fg, ep in zip(u.groups, epochs):
                    if self.fused.reduce_and_step(fg, fuse, accumulate, announced_epoch=ep):
                        u.stepped = True
                    fg.flat_grad = None
                    fg.detach_main_grads()
                ev = get_accelerator().Event()
                ev.record(stream if stream is not None else get_accelerator().current_stream())
            self._arena_free_events[u._arena_key] = ev
            if stream is not None:
                self._pending_reduce.append(u)
            if not u.persistent and not force:
                self._release(u)
            return
        if stream is not None:
            stream.wait_stream(get_accelerator().current_stream())
        with get_accelerator().stream(stream):
            for fg in u.groups:
                if fg.flat_grad is None:
                    fg.ensure_flat_grad()
                if self.world == 1 and not accumulate:
                    fg.zero_unwritten_grads()
                    fg.grad_shard = fg.flat_grad[: fg.part]     # alias: no copy on a single rank
                else:
                    fg.reduce_grads(self.gradient_predivide_factor, accumulate_into_shard=accumulate)
                if fuse:
                    self._step_group(fg, 1.0)
                    fg.grad_shard = None
                    u.stepped = True
                if not (self.world == 1 and not accumulate and not fuse):
                    fg.release_flat_grad()
                else:
                    fg.detach_main_grads()
        if stream is not None:
            self._pending_reduce.append(u)
        if not u.persistent and not force:
            self._release(u)

    def reduce_gradients(self, pipeline_parallel=False):
        pass

    def overlapping_partition_gradients_reduce_epilogue(self):
        pass

    independent_gradient_partition_epilogue = overlapping_partition_gradients_reduce_epilogue

    # -------------------------------------------------------------------------------------- step
    def step(self, closure=None):
        self.micro_step_id = -1
        pending = [fg for fg in self.groups if fg.grad_shard is not None]
        if not pending:      # everything was stepped inside backward (fused path)
            self._global_grad_norm = 0.0
            self.overflow = False
            self._finish_step()
            return
        norm, overflow = self.compute_grad_norm_and_overflow()
        if not self.check_overflow_enabled:
            overflow = False
        self.overflow = overflow
        self._global_grad_norm = norm
        prev_scale = self.loss_scale
        self._update_scale(overflow)
        if overflow:
            self.skipped_steps += 1
            self._finish_step()
            return
        scale = (1.0 / prev_scale) * (self.clip_grad / (norm + 1e-6) if (self.clip_grad > 0 and norm > self.clip_grad) else 1.0)
        for fg in pending:
            self._step_group(fg, scale)
        self._finish_step()

    def _finish_step(self):
        if self.fused is not None:
            self.fused._shards_dirty = True
        for fg in self.groups:
            fg.micro_steps_accumulated = 0
            fg.grad_shard = None
            fg.release_flat_grad()
            fg.hpz_secondary = None          # hpZ: the secondary partition is rebuilt by the next forward gather
        for u in self.units:
            if u.persistent and self.world > 1:
                for fg in u.groups:      # persistent units keep a gathered copy: refresh it
                    fg.refresh_full_from_shard()

    def get_global_grad_norm(self):
        return self._global_grad_norm

    # ------------------------------------------------------------------------- gather for users
    def gather_all(self):
        for u in self.units:
            self._fetch(u)

    def release_all(self):
        for u in self.units:
            self._release(u)

    def unit_of(self, param) -> Optional[Unit]:
        lk = getattr(param, "_sxb_link", None)
        return None if lk is None else lk.group.unit

    def empty_partition_cache(self):
        self.release_all()

    def make_persistent(self, u: Unit):
        """Promote a unit to persistent at run time (schedule compiler's selective gather): gather it once, then it
        is never released and its gathered copy is refreshed after every step like the build-time persistent units."""
        if u.persistent:
            return
        self._fetch(u)
        u.persistent = True

    def destroy(self):
        for h in self._hooks:
            h.remove()
        self._hooks = []

    # --------------------------------------------------------------------------------- state dict
    def _lean(self, fg: FlatGroup, t: torch.Tensor) -> torch.Tensor:
        return t.detach().clone()

    def state_dict(self):
        self.swap_in_all()
        try:
            return self._state_dict()
        finally:
            self.swap_out_all()

    def _state_dict(self):
        sd = {}
        sd[ZERO_STAGE] = 3
        sd[LOSS_SCALER] = self.loss_scaler
        sd["dynamic_loss_scale"] = self.dynamic_loss_scale
        sd["overflow"] = self.overflow
        sd[PARTITION_COUNT] = self.partition_count
        base = {"state": {}, "param_groups": [{k: v for k, v in pg.items() if k != "params"}
                                                for pg in self.optimizer.param_groups]}
        for i, fg in enumerate(self.groups):
            base["state"][i] = {k: (v.detach().clone() if torch.is_tensor(v) else v)
                                for k, v in self._optimizer_state_for(fg).items()}
            base["param_groups"][i]["params"] = [i]
        sd[OPTIMIZER_STATE_DICT] = base
        sd[FP32_FLAT_GROUPS] = [fg.hp_param.data.detach().clone() for fg in self.groups]
        sd[CLIP_GRAD] = self.clip_grad
        sd[DS_VERSION] = __version__
        sd["param_offsets"] = [dict(zip(fg.names, zip(fg.offsets, fg.numels))) for fg in self.groups]
        sd["param_shapes"] = [dict(zip(fg.names, fg.shapes)) for fg in self.groups]
        sd["group_paddings"] = [fg.padding for fg in self.groups]
        return sd

    def load_state_dict(self, state_dict_list, load_optimizer_states=True, load_from_fp32_weights=False,
                        checkpoint_folder=None, load_serial=None, param_shapes=None):
        if isinstance(state_dict_list, dict):
            state_dict_list = [state_dict_list]
        saved_world = state_dict_list[0][PARTITION_COUNT]
        if saved_world != self.world:
            if len(state_dict_list) != saved_world:
                raise NotImplementedError(
                    f"ZeRO-3 checkpoint was saved with dp={saved_world}, current dp={self.world}: pass the state of "
                    "all saved ranks (or use the universal checkpoint) to re-partition")
            return self._load_repartitioned(state_dict_list, load_optimizer_states)
        mine = state_dict_list[min(self.rank, len(state_dict_list) - 1)]
        self.loss_scaler = mine.get(LOSS_SCALER, self.loss_scaler)
        self.dynamic_loss_scale = mine.get("dynamic_loss_scale", self.dynamic_loss_scale)
        self.overflow = mine.get("overflow", False)
        for i, fg in enumerate(self.groups):
            fg.hp_param.data.copy_(mine[FP32_FLAT_GROUPS][i].to(fg.hp_param.device))
            fg.lp_shard.copy_(fg.hp_param.data.to(fg.lp_shard.device))
            fg.lp_shard_written()
            if load_optimizer_states:
                st = self.optimizer.state[fg.hp_param]
                for k, v in mine[OPTIMIZER_STATE_DICT]["state"].get(i, {}).items():
                    st[k] = v.to(fg.hp_param.device) if torch.is_tensor(v) and v.dim() > 0 else v
                for k, v in mine[OPTIMIZER_STATE_DICT]["param_groups"][i].items():
                    if k != "params":
                        fg.param_group[k] = v
        self._refresh_gathered()

    def _load_repartitioned(self, sds, load_optimizer_states):
        for i, fg in enumerate(self.groups):
            def cut(pieces):
                full = torch.cat([p.reshape(-1) for p in pieces])
                out = torch.zeros(fg.part, dtype=full.dtype)
                lo = fg.rank * fg.part
                hi = min(lo + fg.part, full.numel())
                if hi > lo:
                    out[: hi - lo] = full[lo:hi]
                return out
            fg.hp_param.data.copy_(cut([sd[FP32_FLAT_GROUPS][i] for sd in sds]).to(fg.hp_param.device))
            fg.lp_shard.copy_(fg.hp_param.data.to(fg.lp_shard.device))
            fg.lp_shard_written()
            if load_optimizer_states:
                st = self.optimizer.state[fg.hp_param]
                for k, v0 in sds[0][OPTIMIZER_STATE_DICT]["state"].get(i, {}).items():
                    if torch.is_tensor(v0) and v0.dim() > 0:
                        st[k] = cut([sd[OPTIMIZER_STATE_DICT]["state"][i][k] for sd in sds]).to(fg.hp_param.device)
                    else:
                        st[k] = v0
        self._refresh_gathered()

    def _refresh_gathered(self):
        for u in self.units:
            if self.world == 1:
                continue
            if u.persistent:
                for fg in u.groups:
                    fg.refresh_full_from_shard()


def estimate_zero3_model_states_mem_needs(total_params, largest_layer_params, num_gpus_per_node=1, num_nodes=1,
                                          cpu_offload=True, cpu_offload_params=True, zero_init=True,
                                          additional_buffer_factor=1.5):
    """(cpu_bytes, gpu_bytes, largest_layer_bytes) (reference: stage3.py:3120-3160)."""
    total_gpus = num_nodes * num_gpus_per_node
    gpus_factor = 1 / num_nodes
    largest_layer_memory = 4 * largest_layer_params
    if cpu_offload:
        if cpu_offload_params:
            gpu_mem = largest_layer_memory
            cpu_mem = total_params * (18 * gpus_factor if zero_init else max(4 * num_gpus_per_node, 18 * gpus_factor)) * additional_buffer_factor
        else:
            gpu_mem = largest_layer_memory + int(2 * total_params / total_gpus)
            cpu_mem = total_params * (16 * gpus_factor if zero_init else max(4 * num_gpus_per_node, 16 * gpus_factor)) * additional_buffer_factor
    else:
        gpu_mem = largest_layer_memory + int(18 * total_params / total_gpus)
        cpu_mem = largest_layer_params * 4 * num_gpus_per_node * additional_buffer_factor if zero_init \
            else total_params * 4 * num_gpus_per_node * additional_buffer_factor
    return int(cpu_mem), int(gpu_mem), largest_layer_memory


def estimate_zero3_model_states_mem_needs_all_live(model, num_gpus_per_node=1, num_nodes=1, additional_buffer_factor=1.5):
    total = sum(p.numel() for p in model.parameters())
    largest = max((sum(p.numel() for p in m.parameters(recurse=False)) for m in model.modules()), default=0)
    estimate_zero3_model_states_mem_needs_all_cold(total, largest, num_gpus_per_node, num_nodes, additional_buffer_factor)


def estimate_zero3_model_states_mem_needs_all_cold(total_params, largest_layer_params, num_gpus_per_node=1, num_nodes=1,
                                                   additional_buffer_factor=1.5):
    print("Estimated memory needed for params, optim states and gradients for a:\n"
          f"HW: Setup with {num_nodes} node(s), {num_gpus_per_node} GPU(s) per node.\n"
          f"SW: Model with {int(total_params / 1e6)}M total params, {int(largest_layer_params / 1e6)}M largest layer params.")
    print("  per CPU  |  per GPU |   Options")
    for cpu_offload in (True, False):
        for cpu_offload_params in ((True, False) if cpu_offload else (False,)):
            for zero_init in (True, False):
                c, g, _ = estimate_zero3_model_states_mem_needs(total_params, largest_layer_params, num_gpus_per_node,
                                                                num_nodes, cpu_offload, cpu_offload_params, zero_init,
                                                                additional_buffer_factor)
                opts = (f"offload_param={'cpu' if cpu_offload_params else 'none'}, "
                        f"offload_optimizer={'cpu' if cpu_offload else 'none'}, zero_init={int(zero_init)}")
                print(f" {c / 2**30:7.2f}GB | {g / 2**30:6.2f}GB | {opts}")
