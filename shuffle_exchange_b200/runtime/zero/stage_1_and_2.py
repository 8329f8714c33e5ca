"""ZeRO stage 1 / 2 (optimizer-state and gradient sharding) + shuffle-exchange replica averaging.

Capabilities of the reference ``DeepSpeedZeroOptimizer`` (deepspeed/runtime/zero/stage_1_and_2.py):
flat bit16 groups with fp32 master partitions (:431-534), gradient reduction into the owner's
partition overlapped with backward (:1090-1342), the partitioned step (:2058-2268), elastic / rigid
checkpoint state (:2472-2743), MoE param groups sharded over their expert-data-parallel group
(:797-822), CPU offload of the optimizer state (:1347-1523), and the fork's ``method / rings /
shuffle_step / slice_count`` decentralised averaging with the user-called trio
``shuffle_exchange() / synchronization() / reset_rings()`` (:692-734, :2185-2250).

Design differences: storage is :class:`FlatGroup`; gradients are written by the
wgrad GEMMs directly into the flat gradient buffer ("main grad"), the reduction of a group is ONE
reduce-scatter launched on a side stream as soon as its last gradient lands, and the update is one
fused kernel per group that also emits the bf16 shard.  Nothing on this path names a device
literal, so it runs under gloo on CPU (the reference cannot: hard-coded ``cuda`` at :199, :2071).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from ...comm import comm as dist
from ...device import get_accelerator
from ...ops import functional as _F
from ...utils import groups as groups_mod
from ...utils.logging import logger
from ...version import __version__
from .flat_group import FlatGroup
from .replica_averaging import ReplicaAverager
from .sharded_optimizer import (BASE_OPTIMIZER_STATE, BASE_OPTIMIZER_STATE_STEP, CLIP_GRAD, DS_VERSION, GROUP_PADDINGS,
                                LOSS_SCALER, PARAM_SLICE_MAPPINGS, PARTITION_COUNT, SINGLE_PARTITION_OF_FP32_GROUPS,
                                ZERO_STAGE, ShardedOptimizerBase)

OPTIMIZER_ALLGATHER_TIMER = "optimizer_allgather"
OPTIMIZER_GRADIENTS_TIMER = "optimizer_gradients"
OPTIMIZER_STEP_TIMER = "optimizer_step"
OPTIMIZER_TIMERS = [OPTIMIZER_ALLGATHER_TIMER, OPTIMIZER_GRADIENTS_TIMER, OPTIMIZER_STEP_TIMER]


def _is_moe_group(pg: dict) -> bool:
    return bool(pg.get("moe", False))


class DeepSpeedZeroOptimizer(ShardedOptimizerBase):
    def __init__(self, init_optimizer, param_names=None, timers=None, static_loss_scale=1.0, dynamic_loss_scale=False,
                 dynamic_loss_args=None, verbose=True, contiguous_gradients=True, reduce_bucket_size=500000000,
                 use_multi_rank_bucket_allreduce=True, allgather_bucket_size=5000000000, dp_process_group=None,
                 expert_parallel_group=None, expert_data_parallel_group=None, reduce_scatter=True, overlap_comm=False,
                 offload_optimizer_config=None, mpu=None, clip_grad=0.0, gradient_accumulation_dtype=torch.float32,
                 communication_data_type=torch.float16, postscale_gradients=True, gradient_predivide_factor=1.0,
                 gradient_accumulation_steps=1, ignore_unused_parameters=True, partition_grads=True,
                 round_robin_gradients=False, has_moe_layers=False, fp16_master_weights_and_gradients=False,
                 elastic_checkpoint=False, check_grad_overflow=None, dtype=None,
                 method="RR", rings=4, shuffle_step=1, slice_count=2, seed=1234, replica_averaging=True,
                 fused_replica_averaging=None, shard_optimizer=True):
        first = next(p for g in init_optimizer.param_groups for p in g["params"])
        lp_dtype = dtype or first.dtype
        super().__init__(init_optimizer, timers=timers, dtype=lp_dtype, static_loss_scale=static_loss_scale,
                         dynamic_loss_scale=dynamic_loss_scale, dynamic_loss_args=dynamic_loss_args, clip_grad=clip_grad,
                         dp_process_group=dp_process_group, mpu=mpu,
                         gradient_accumulation_steps=gradient_accumulation_steps,
                         gradient_accumulation_dtype=gradient_accumulation_dtype,
                         communication_data_type=communication_data_type,
                         gradient_predivide_factor=gradient_predivide_factor, postscale_gradients=postscale_gradients,
                         check_grad_overflow=check_grad_overflow)
        self.zero_stage = 2 if partition_grads else 1
        self.partition_gradients = partition_grads
        self.contiguous_gradients = contiguous_gradients
        self.reduce_bucket_size = int(reduce_bucket_size)
        self.allgather_bucket_size = int(allgather_bucket_size)
        self.overlap_comm = overlap_comm and get_accelerator().is_available()
        self.elastic_checkpoint = elastic_checkpoint
        self.has_moe_layers = has_moe_layers
        self.param_names = param_names or {}
        self.round_robin_gradients = round_robin_gradients
        self.ignore_unused_parameters = ignore_unused_parameters
        self.cpu_offload = bool(offload_optimizer_config is not None and
                                str(getattr(offload_optimizer_config, "device", "none")).endswith("cpu"))
        self.cpu_offload_pin_memory = bool(self.cpu_offload and getattr(offload_optimizer_config, "pin_memory", False))
        self.expert_dp_process_group = expert_data_parallel_group
        self.ep_process_group = expert_parallel_group

        # ---- the fork: slices + replica averaging -------------------------------------------------
        self.real_dp_process_group = dp_process_group
        self.averager: Optional[ReplicaAverager] = None
        world = dist.get_world_size(dp_process_group) if dist.is_initialized() else 1
        if replica_averaging and dist.is_initialized() and world > 1 and slice_count < world:
            self.averager = ReplicaAverager(method=method, rings=rings, shuffle_step=shuffle_step,
                                            slice_count=slice_count, seed=seed, fused=fused_replica_averaging)
            self.dp_process_group = self.averager.slice_pg
        self.method, self.rings, self.shuffle_step, self.slice_count = method, rings, shuffle_step, slice_count
        if shard_optimizer and dist.is_initialized():
            self.partition_count_value = dist.get_world_size(self.dp_process_group)
            self.partition_rank = dist.get_rank(self.dp_process_group)
        else:
            self.partition_count_value, self.partition_rank = 1, 0

        # ---- flat groups ------------------------------------------------------------------------
        self.bit16_groups: List[List[torch.nn.Parameter]] = []
        self.reduce_stream = get_accelerator().Stream() if self.overlap_comm else None
        self._pending_reduce = []
        for i, pg in enumerate(self.optimizer.param_groups):
            params = [p for p in pg["params"]]
            trainable = [p for p in params if p.requires_grad]
            if round_robin_gradients:
                trainable = self._round_robin_reorder(trainable, self.partition_count_value)
            self.bit16_groups.append(trainable)
            if _is_moe_group(pg) and self.expert_dp_process_group is not None:
                egrp = self.expert_dp_process_group[pg["name"]] if isinstance(self.expert_dp_process_group, dict) \
                    else self.expert_dp_process_group
                w, r, grp = dist.get_world_size(egrp), dist.get_rank(egrp), egrp
            else:
                w, r, grp = self.partition_count_value, self.partition_rank, self.dp_process_group
            names = [self.param_names.get(p, f"group{i}.param{j}") for j, p in enumerate(trainable)]
            fg = FlatGroup(trainable, names, w, r, grp, self.device, lp_dtype,
                           grad_dtype=(gradient_accumulation_dtype if not partition_grads and gradient_accumulation_steps > 1
                                       else lp_dtype) if gradient_accumulation_dtype is not None else lp_dtype,
                           group_index=i, keep_full_params=True, hp_on_cpu=self.cpu_offload,
                           pin=self.cpu_offload_pin_memory)
            hp = self._register_group(fg, pg)
            pg["params"] = [hp]
            fg.frozen_params = [p for p in params if not p.requires_grad]
            self._install_ready_hooks(fg)
        self.single_partition_of_fp32_groups = [g.hp_param for g in self.groups]
        self.averaged_gradients: Dict[int, torch.Tensor] = {}
        self._link_params()
        if verbose and dist.get_rank() == 0:
            tot = sum(g.total for g in self.groups)
            logger.info(f"ZeRO stage {self.zero_stage}: {len(self.groups)} flat group(s), {tot:,} elements, "
                        f"partition world {self.partition_count_value}"
                        + (f", replica averaging {method} over {self.averager.n_replicas} replicas"
                           if self.averager else ""))

    # ------------------------------------------------------------------------------- utilities
    @staticmethod
    def _round_robin_reorder(tensors, num_partitions):
        buckets = {i: [] for i in range(num_partitions)}
        for i, t in enumerate(tensors):
            buckets[i % num_partitions].append(t)
        out = []
        for i in range(num_partitions):
            out.extend(buckets[i])
        return out

    def _link_params(self):
        """lp <-> hp linkage used by ``safe_get_full_fp32_param`` & friends (utils/tensor_fragment.py)."""
        from ...utils.tensor_fragment import link_group_params
        for fg in self.groups:
            link_group_params(fg, self)

    def _install_ready_hooks(self, fg: FlatGroup):
        fg.grads_ready = 0

        def on_ready(param, fg=fg):
            # called once per USE of the parameter (tied embedding / LM head: twice); the group is complete when
            # every distinct parameter received all the contributions its forward uses announced
            if getattr(fg, "_reduced_this_micro_step", False) and self._reduce_now():
                raise RuntimeError(
                    f"ZeRO-{self.zero_stage}: a gradient contribution for flat group {fg.index} arrived after its "
                    "reduction was launched (a parameter used by both main-grad ops and plain autograd ops, or a "
                    "forward that ran outside the counted window); its gradient would be dropped")
            param._sxb_calls = getattr(param, "_sxb_calls", 0) + 1
            if param._sxb_calls != _F.expected_uses(param):
                return
            fg.grads_ready += 1
            if fg.grads_ready == fg.n_trainable and self._reduce_now():
                self._launch_group_reduce(fg)

        for i, p in enumerate(fg.params):
            if not fg.trainable[i]:
                continue
            p._sxb_grad_ready = on_ready

            def autograd_hook(param, fg=fg, idx=i):
                # gradient produced by an op that does not implement the main-grad protocol
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

            p.register_post_accumulate_grad_hook(autograd_hook)

    def _reduce_now(self) -> bool:
        """Reduce in this micro-step?  stage 2: always; stage 1: only at the accumulation boundary."""
        return self.partition_gradients or self._is_boundary

    _is_boundary = True

    def set_gradient_accumulation_boundary(self, is_boundary: bool):
        self._is_boundary = is_boundary

    # --------------------------------------------------------------------------------- backward
    def _pre_backward(self):
        _F.close_use_window()
        for g in self.groups:
            g.grads_ready = 0
            for p in g.params:
                p._sxb_calls = 0
            g.attach_main_grads()
            if self.partition_gradients or self._first_micro_step():
                g.mark_grads_unwritten()

    def _first_micro_step(self) -> bool:
        return all(g.micro_steps_accumulated == 0 for g in self.groups)

    def _post_backward(self):
        _F.open_use_window()
        for g in self.groups:
            g.micro_steps_accumulated += 1

    def _launch_group_reduce(self, fg: FlatGroup):
        if getattr(fg, "_reduced_this_micro_step", False):
            return
        fg._reduced_this_micro_step = True
        accumulate = self.partition_gradients and fg.micro_steps_accumulated > 0
        if self.reduce_stream is not None:
            self.reduce_stream.wait_stream(get_accelerator().current_stream())
            with get_accelerator().stream(self.reduce_stream):
                fg.reduce_grads(self.gradient_predivide_factor, accumulate_into_shard=accumulate)
            self._pending_reduce.append(fg)
        else:
            fg.reduce_grads(self.gradient_predivide_factor, accumulate_into_shard=accumulate)

    def reduce_gradients(self, pipeline_parallel=False):
        self.overlapping_partition_gradients_reduce_epilogue()

    def overlapping_partition_gradients_reduce_epilogue(self):
        """Flush: reduce every group that has not been reduced in this micro-step, join the stream."""
        if self._reduce_now():
            for fg in self.groups:
                if not getattr(fg, "_reduced_this_micro_step", False):
                    if fg.flat_grad is None:
                        fg.ensure_flat_grad()
                    self._launch_group_reduce(fg)
        if self.reduce_stream is not None and self._pending_reduce:
            get_accelerator().current_stream().wait_stream(self.reduce_stream)
            self._pending_reduce = []
        for fg in self.groups:
            if getattr(fg, "_reduced_this_micro_step", False):
                fg._reduced_this_micro_step = False
                if self.partition_gradients:
                    fg.release_flat_grad()
        for i, fg in enumerate(self.groups):
            if fg.grad_shard is not None:
                self.averaged_gradients[i] = fg.grad_shard

    independent_gradient_partition_epilogue = overlapping_partition_gradients_reduce_epilogue

    # ------------------------------------------------------------------------------------- step
    def step(self, closure=None):
        self.micro_step_id = -1
        timers = self.timers
        if timers:
            timers(OPTIMIZER_GRADIENTS_TIMER).start()
        norm, overflow = self.compute_grad_norm_and_overflow()
        if not self.check_overflow_enabled:
            overflow = overflow and False
        if self.averager is not None and self.check_overflow_enabled:
            # replicas must agree on skipping: any overflow anywhere skips everywhere (reference :2067-2073)
            flag = torch.tensor([1.0 if overflow else 0.0], device=self.device)
            dist.all_reduce(flag, op=dist.ReduceOp.MAX)
            overflow = bool(flag.item() > 0)
        self.overflow = overflow
        self._global_grad_norm = norm
        if timers:
            timers(OPTIMIZER_GRADIENTS_TIMER).stop()
        prev_scale = self.loss_scale
        self._update_scale(overflow)
        if overflow:
            self.skipped_steps += 1
            self._reset_grads()
            return
        if self.averager is not None and self.averager.method == "Gossip":
            # push-sum merge goes into the fp32 master (the reference merges into the bit16 partition,
            # which its own fp32->bit16 copy a few lines later overwrites: stage_1_and_2.py:2092-2108,2176)
            self.averager.gossip_merge([fg.hp_param.data for fg in self.groups])
        if timers:
            timers(OPTIMIZER_STEP_TIMER).start()
        scale = (1.0 / prev_scale) * (self.clip_grad / (norm + 1e-6) if (self.clip_grad > 0 and norm > self.clip_grad) else 1.0)
        for fg in self.groups:
            if fg.grad_shard is None:
                continue
            self._step_group(fg, scale)
        if timers:
            timers(OPTIMIZER_STEP_TIMER).stop()
            timers(OPTIMIZER_ALLGATHER_TIMER).start()
        # cross-replica averaging of the freshly updated partition, then intra-slice all-gather
        if self.averager is not None:
            self.averager.average([fg.lp_shard for fg in self.groups])
        for fg in self.groups:
            fg.refresh_full_from_shard()
        if timers:
            timers(OPTIMIZER_ALLGATHER_TIMER).stop()
        self._reset_grads()

    def _reset_grads(self):
        for fg in self.groups:
            fg.micro_steps_accumulated = 0
            fg.grads_ready = 0
            if fg.grad_shard is not None:
                fg.grad_shard = None if self.partition_gradients else fg.grad_shard
            fg.release_flat_grad()
            fg.grad_shard = None
        self.averaged_gradients = {}

    def zero_grad(self, set_to_none=True):
        super().zero_grad(set_to_none)

    # ------------------------------------------------------------------------- fork's user API
    def shuffle_exchange(self):
        if self.averager is not None:
            return self.averager.shuffle_exchange()
        return False

    def synchronization(self):
        if self.averager is None:
            return
        # like the reference (:722-728) only the bit16 model weights are averaged; fp32 masters keep
        # their own trajectory
        self.averager.synchronization([fg.flat_lp for fg in self.groups])

    def reset_rings(self, rings):
        if self.averager is not None:
            self.averager.reset_rings(rings)
            self.rings = rings

    # ------------------------------------------------------------------------------- misc API
    def update_lp_params(self):
        for fg in self.groups:
            fg.lp_shard.copy_(fg.hp_param.data.to(fg.lp_shard.device))
            fg.refresh_full_from_shard()

    def refresh_fp32_params(self):
        for fg in self.groups:
            fg.hp_param.data.copy_(fg.lp_shard.to(fg.hp_param.device))

    _restore_from_bit16_weights = refresh_fp32_params

    def destroy(self):
        for fg in self.groups:
            for p in fg.params:
                for a in ("_sxb_grad_ready", "_sxb_main_grad", "_sxb_grad_written"):
                    if hasattr(p, a):
                        delattr(p, a)

    def get_global_grad_norm(self):
        return self._global_grad_norm

    # --------------------------------------------------------------------------- checkpointing
    def _lean(self, fg: FlatGroup, t: torch.Tensor) -> torch.Tensor:
        """Strip the tail padding that only the last partition carries."""
        lo = fg.rank * fg.part
        valid = max(0, min(fg.part, fg.unpadded_numel - lo))
        return t[:valid].detach().clone()

    def state_dict(self):
        sd = {}
        sd[LOSS_SCALER] = self.loss_scaler
        sd["dynamic_loss_scale"] = self.dynamic_loss_scale
        sd["overflow"] = self.overflow
        sd[CLIP_GRAD] = self.clip_grad
        base = []
        steps = []
        for fg in self.groups:
            st = self._optimizer_state_for(fg)
            lean = {}
            for k, v in st.items():
                if torch.is_tensor(v) and v.numel() == fg.part:
                    lean[k] = self._lean(fg, v)
                else:
                    lean[k] = v
            base.append(lean)
            steps.append(st.get("step", 0))
        sd[BASE_OPTIMIZER_STATE] = base
        sd[BASE_OPTIMIZER_STATE_STEP] = steps
        sd["base_optimizer_param_groups"] = [{k: v for k, v in pg.items() if k != "params"}
                                             for pg in self.optimizer.param_groups]
        sd[SINGLE_PARTITION_OF_FP32_GROUPS] = [self._lean(fg, fg.hp_param.data) for fg in self.groups]
        sd[ZERO_STAGE] = self.zero_stage
        sd[GROUP_PADDINGS] = [fg.padding if fg.rank == fg.world - 1 else 0 for fg in self.groups]
        sd[PARTITION_COUNT] = [fg.world for fg in self.groups]
        sd[DS_VERSION] = __version__
        sd[PARAM_SLICE_MAPPINGS] = [fg.param_slice_mappings() for fg in self.groups]
        sd["param_offsets"] = [dict(zip(fg.names, zip(fg.offsets, fg.numels))) for fg in self.groups]
        sd["param_shapes"] = [dict(zip(fg.names, fg.shapes)) for fg in self.groups]
        if self.averager is not None:
            sd["replica_averaging"] = self.averager.state_dict()
        return sd

    def _repartition(self, fg: FlatGroup, pieces: List[torch.Tensor]) -> torch.Tensor:
        """Concatenate all saved (lean) partitions of a group and cut out this rank's slice."""
        full = torch.cat([p.reshape(-1) for p in pieces])
        out = torch.zeros(fg.part, dtype=full.dtype)
        lo = fg.rank * fg.part
        hi = min(lo + fg.part, full.numel())
        if hi > lo:
            out[: hi - lo] = full[lo:hi]
        return out

    def load_state_dict(self, state_dict_list, load_optimizer_states=True, load_from_fp32_weights=False,
                        checkpoint_folder=None, load_serial=None, param_shapes=None):
        """``state_dict_list``: the optimizer state of EVERY saved partition rank, in rank order (a single
        dict is accepted when the partition count is unchanged).  Re-partitions when the count differs."""
        if isinstance(state_dict_list, dict):
            state_dict_list = [state_dict_list]
        mine_idx = min(self.partition_rank, len(state_dict_list) - 1)
        mine = state_dict_list[mine_idx]
        self.loss_scaler = mine.get(LOSS_SCALER, self.loss_scaler)
        self.dynamic_loss_scale = mine.get("dynamic_loss_scale", self.dynamic_loss_scale)
        self.overflow = mine.get("overflow", False)
        self.clip_grad = mine.get(CLIP_GRAD, self.clip_grad)
        saved_world = mine[PARTITION_COUNT][0] if isinstance(mine[PARTITION_COUNT], (list, tuple)) else mine[PARTITION_COUNT]
        same = saved_world == self.groups[0].world and len(state_dict_list) in (1, saved_world)
        if not same and len(state_dict_list) != saved_world:
            raise RuntimeError(f"checkpoint was saved with {saved_world} partitions; need all of them to "
                               f"re-partition to {self.groups[0].world} (got {len(state_dict_list)})")
        for gi, fg in enumerate(self.groups):
            if same:
                src = state_dict_list[mine_idx if len(state_dict_list) > 1 else 0]
                hp = torch.zeros(fg.part, dtype=torch.float32)
                piece = src[SINGLE_PARTITION_OF_FP32_GROUPS][gi].reshape(-1)
                hp[: piece.numel()] = piece
            else:
                hp = self._repartition(fg, [sd[SINGLE_PARTITION_OF_FP32_GROUPS][gi] for sd in state_dict_list])
            if load_from_fp32_weights:
                fg.hp_param.data.copy_(hp.to(fg.hp_param.device))
            else:
                # zero_optimization.load_from_fp32_weights = false: the master is rebuilt from the bit16 module
                # weights the engine loaded just before (reference: stage_1_and_2.py _restore_from_bit16_weights)
                # (the parameters alias the flat bit16 buffer, so its shard already holds the loaded values)
                fg.hp_param.data.copy_(fg.lp_shard.to(fg.hp_param.device, dtype=torch.float32))
            if load_optimizer_states:
                if same:
                    saved = (state_dict_list[mine_idx if len(state_dict_list) > 1 else 0])[BASE_OPTIMIZER_STATE][gi]
                    merged = {}
                    for k, v in saved.items():
                        if torch.is_tensor(v) and v.dim() > 0:
                            t = torch.zeros(fg.part, dtype=v.dtype)
                            t[: v.numel()] = v.reshape(-1)
                            merged[k] = t
                        else:
                            merged[k] = v
                else:
                    merged = {}
                    keys = state_dict_list[0][BASE_OPTIMIZER_STATE][gi].keys()
                    for k in keys:
                        v0 = state_dict_list[0][BASE_OPTIMIZER_STATE][gi][k]
                        if torch.is_tensor(v0) and v0.dim() > 0:
                            merged[k] = self._repartition(fg, [sd[BASE_OPTIMIZER_STATE][gi][k] for sd in state_dict_list])
                        else:
                            merged[k] = v0
                st = self.optimizer.state[fg.hp_param]
                for k, v in merged.items():
                    st[k] = v.to(fg.hp_param.device) if torch.is_tensor(v) and v.dim() > 0 else v
                pgs = mine.get("base_optimizer_param_groups")
                if pgs is not None and gi < len(pgs):
                    for k, v in pgs[gi].items():
                        fg.param_group[k] = v
            # bit16 weights follow the restored master
            fg.lp_shard.copy_(fg.hp_param.data.to(fg.lp_shard.device))
            fg.refresh_full_from_shard()
        if self.averager is not None and "replica_averaging" in mine:
            self.averager.load_state_dict(mine["replica_averaging"])


def estimate_zero2_model_states_mem_needs(total_params, num_gpus_per_node=1, num_nodes=1, cpu_offload=True,
                                          additional_buffer_factor=1.5):
    """(cpu_bytes, gpu_bytes) for model states under ZeRO-2 (reference: stage_1_and_2.py:2758-2790)."""
    total_gpus = num_nodes * num_gpus_per_node
    if cpu_offload:
        gpu_mem = 2 * total_params
        cpu_mem = total_params * max(4 * total_gpus, 16) * additional_buffer_factor
    else:
        gpu_mem = 4 * total_params + int(16 * total_params / total_gpus)
        cpu_mem = total_params * 4 * num_gpus_per_node * additional_buffer_factor
    return int(cpu_mem), int(gpu_mem)


def model_to_params(model):
    return sum(p.numel() for p in model.parameters())


def estimate_zero2_model_states_mem_needs_all_live(model, num_gpus_per_node=1, num_nodes=1, additional_buffer_factor=1.5):
    estimate_zero2_model_states_mem_needs_all_cold(model_to_params(model), num_gpus_per_node, num_nodes,
                                                   additional_buffer_factor)


def estimate_zero2_model_states_mem_needs_all_cold(total_params, num_gpus_per_node=1, num_nodes=1,
                                                   additional_buffer_factor=1.5):
    def fmt(cpu_offload):
        return f"offload_optimizer={'cpu' if cpu_offload else 'none'}"
    nodes_str = "nodes" if num_nodes > 1 else "node"
    gpus_str = "GPUs" if num_gpus_per_node > 1 else "GPU"
    print("Estimated memory needed for params, optim states and gradients for a:\n"
          f"HW: Setup with {num_nodes} {nodes_str}, {num_gpus_per_node} {gpus_str} per node.\n"
          f"SW: Model with {int(total_params / 1e6)}M total params.")
    print("  per CPU  |  per GPU |   Options")
    for cpu_offload in (True, False):
        cpu_mem, gpu_mem = estimate_zero2_model_states_mem_needs(total_params, num_gpus_per_node, num_nodes,
                                                                 cpu_offload, additional_buffer_factor)
        print(f" {cpu_mem / 2**30:7.2f}GB | {gpu_mem / 2**30:6.2f}GB | {fmt(cpu_offload)}")
