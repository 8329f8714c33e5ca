"""PipelineEngine — runs ``PipeSchedule`` instruction streams (1F1B training, forward-only evaluation) over the local
stage of a ``PipelineModule``.

Same user surface as the reference engine (``deepspeed/runtime/pipe/engine.py``: ``train_batch`` :338, ``eval_batch``
:428, ``set_dataloader / set_dataiterator / set_batch_fn``, ``forward / backward / step`` disabled).  How it is built:

* each neighbouring stage is a :class:`_Wire` — NCCL (or gloo) ``send`` / ``recv`` of a *tuple of tensors* whose
  shapes and dtypes travel once, as ONE fixed-size header message, unless ``dynamic_shape`` asks for every time
  (north star: pipeline traffic stays on NCCL p2p);
* micro-batches in flight live in :class:`_Slot` records (inputs / labels / outputs), indexed by the schedule's buffer id;
* every schedule instruction type maps to one bound handler (``self._handlers``);
* unlike the reference (which asserts ZeRO stage < 2, :77-78) the stage's data-parallel group may run ZeRO-2 / ZeRO-3:
  every micro-batch backward is one bracketed ``_pre_backward`` / ``_post_backward`` of the sharded optimizer, gradient
  shards accumulate across micro-batches and the step is taken at the batch boundary (parameters are
  gathered per micro-batch pass inside the stage's data-parallel group).
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence, Tuple, Union

import torch

from ... import comm as dist
from ...device import get_accelerator
from ...ops import functional as _F
from ...utils.logging import logger
from ...utils.timer import ThroughputTimer
from ..engine import MEMORY_OPT_ALLREDUCE_SIZE, DeepSpeedEngine
from . import p2p, schedule
from .module import PipelineError, PipelineModule

TARGET_ID = -2
LOG_STAGE = -2
DATA_PARALLEL_ID = -2

_DTYPES = [torch.float32, torch.float16, torch.bfloat16, torch.int64, torch.int32, torch.bool, torch.uint8,
           torch.int8, torch.float64, torch.int16]
_HEADER_INTS = 96          # fixed-size shape header: [is_tuple, n, (dtype, ndim, dims...)*]

Tensors = Union[torch.Tensor, Tuple[torch.Tensor, ...]]


def is_even(n):
    return n % 2 == 0


def _as_list(x: Tensors) -> List[torch.Tensor]:
    return [x] if torch.is_tensor(x) else list(x)


class _Wire:
    """Point-to-point channel to ONE neighbouring stage for tuples of tensors."""

    def __init__(self, peer_stage: int, device, resend_header_every_time: bool):
        self.peer, self.device, self.always = peer_stage, device, resend_header_every_time
        self._sent_header = False
        self._recv_like: Optional[List[torch.Tensor]] = None
        self._recv_is_tuple = False

    def reset(self):
        self._sent_header, self._recv_like = False, None

    # ---- header
    def _encode(self, x: Tensors) -> torch.Tensor:
        items = _as_list(x)
        words = [0 if torch.is_tensor(x) else 1, len(items)]
        for t in items:
            words += [_DTYPES.index(t.dtype), t.dim(), *t.shape]
        if len(words) > _HEADER_INTS:
            raise PipelineError(f"activation tuple too large for the shape header ({len(words)} > {_HEADER_INTS} ints)")
        return torch.tensor(words + [0] * (_HEADER_INTS - len(words)), dtype=torch.long, device=self.device)

    def _decode(self, header: torch.Tensor):
        w = header.tolist()
        self._recv_is_tuple, n = bool(w[0]), w[1]
        pos, like = 2, []
        for _ in range(n):
            dt, nd = _DTYPES[w[pos]], w[pos + 1]
            like.append(torch.empty(w[pos + 2: pos + 2 + nd], dtype=dt, device=self.device))
            pos += 2 + nd
        self._recv_like = like

    # ---- payload
    def send(self, x: Tensors) -> None:
        if self.always or not self._sent_header:
            p2p.send(self._encode(x), self.peer)
            self._sent_header = True
        for t in _as_list(x):
            p2p.send(t.contiguous(), self.peer)

    def recv(self) -> Tensors:
        if self.always or self._recv_like is None:
            header = torch.empty(_HEADER_INTS, dtype=torch.long, device=self.device)
            p2p.recv(header, self.peer)
            self._decode(header)
        got = []
        for like in self._recv_like:
            buf = torch.empty_like(like)
            p2p.recv(buf, self.peer)
            got.append(buf)
        return tuple(got) if self._recv_is_tuple else got[0]

    def recv_like(self, templates: Sequence[torch.Tensor]) -> List[torch.Tensor]:
        """Receive one tensor per template (gradients: shapes are those of what was sent the other way)."""
        out = []
        for t in templates:
            buf = torch.empty_like(t)
            p2p.recv(buf, self.peer)
            out.append(buf)
        return out


class _Slot:
    """One micro-batch in flight on this stage."""
    __slots__ = ("inputs", "labels", "outputs", "out_grads")

    def __init__(self):
        self.inputs = self.labels = self.outputs = self.out_grads = None


class PipelineEngine(DeepSpeedEngine):
    ID_TO_DTYPE = _DTYPES
    DTYPE_TO_ID = {d: i for i, d in enumerate(_DTYPES)}

    def __init__(self, has_bool_tensors=False, *super_args, **super_kwargs):
        super().__init__(*super_args, **super_kwargs)
        if not isinstance(self.module, PipelineModule):
            raise AssertionError("model must base PipelineModule")
        self.has_bool_tensors = has_bool_tensors
        self.enable_backward_allreduce = False          # reductions are schedule instructions here
        self.pipeline_enable_backward_allreduce = True
        self.grid = self.module._grid
        self.global_rank = self.grid.get_global_rank()
        self.micro_batch_size = self.train_micro_batch_size_per_gpu()
        self.micro_batches = self.gradient_accumulation_steps()
        if self.dp_world_size != self.grid.data_parallel_size:
            raise AssertionError(f"engine dp size {self.dp_world_size} != grid dp size {self.grid.data_parallel_size}")
        if self.train_batch_size() != self.micro_batch_size * self.micro_batches * self.grid.data_parallel_size:
            raise AssertionError("train_batch_size != micro_batch_size * gradient_accumulation_steps * data-parallel size")
        self.num_stages, self.stage_id = self.grid.pipe_parallel_size, self.grid.get_stage_id()
        self.prev_stage, self.next_stage = self.stage_id - 1, self.stage_id + 1
        self.is_pipe_parallel = self.num_stages > 1
        self.is_data_parallel = self.grid.data_parallel_size > 1
        self.is_model_parallel = self.grid.model_parallel_size > 1
        pc = self._config.pipeline
        self.is_pipe_partitioned = self.is_model_parallel and pc["pipe_partitioned"]
        self.is_grad_partitioned = self.is_model_parallel and pc["grad_partitioned"]
        self.dynamic_shape = self.module.dynamic_shape
        if pc["activation_checkpoint_interval"]:
            self.module.activation_checkpoint_interval = pc["activation_checkpoint_interval"]
            if pc.get("use_reentrant", True) is False:
                # ``pipeline.use_reentrant: false``: recompute through saved-tensor hooks instead of a re-entrant
                # autograd function (no dummy requires_grad inputs, works with keyword tensors)
                from ..activation_checkpointing import checkpointing as ds_checkpointing
                self.module.activation_checkpoint_func = ds_checkpointing.non_reentrant_checkpoint
                if self.grid.get_global_rank() == 0:
                    logger.info("CONFIG: activation_checkpoint_func=non_reentrant_checkpoint")
        self.module.checkpoint_parallel_write_pipeline = self._config.checkpoint_parallel_write_pipeline
        self.has_attention_mask = type(self.module).__name__ == "GPT2ModelPipe"
        self.batch_timer = ThroughputTimer(self._config.timers_config.throughput, batch_size=self.train_batch_size(),
                                           logging_fn=self.tput_log, monitor_memory=False,
                                           steps_per_output=self.steps_per_print())
        # ---- data
        self.data_iterator, self.batch_fn = None, None
        if self.training_data:
            self._build_data_iter(self.training_data)
        # ---- wires to the neighbours
        self._up: Optional[_Wire] = None       # towards stage - 1
        self._down: Optional[_Wire] = None     # towards stage + 1
        if self.is_pipe_parallel:
            p2p.init_process_groups(self.grid)
            if not self.is_first_stage():
                self._up = _Wire(self.prev_stage, self.device, self.dynamic_shape)
            if not self.is_last_stage():
                self._down = _Wire(self.next_stage, self.device, self.dynamic_shape)
            self._warm_up_links()
        # ---- per-batch state
        self._slots: List[_Slot] = []
        self._force_grad_boundary = False
        self._compute_loss = True
        self.eval_return_logits, self.outputs = False, None
        self.loss = torch.zeros((), device=self.device)
        self.total_loss = None
        self.agg_train_loss = None
        self.dp_group_loss = torch.zeros((), device=self.device)
        self.fwd_outputs: list = []
        self.log_batch_step_id = -1
        if self.is_last_stage():
            self.loss_model = self.module.loss_fn
        S = schedule
        self._handlers = {
            S.LoadMicroBatch: self._load_micro_batch, S.ForwardPass: self._forward_pass,
            S.BackwardPass: self._backward_pass, S.SendActivation: self._send_activations,
            S.RecvActivation: self._recv_activations, S.SendGrad: self._send_grads, S.RecvGrad: self._recv_grads,
            S.ReduceTiedGrads: self._reduce_tied_grads, S.ReduceGrads: self._reduce_grads,
            S.OptimizerStep: self._optimizer_step,
        }

    def _warm_up_links(self):
        """First message on every link, in an order that cannot deadlock (even stages talk downstream first): builds
        the p2p communicators before the schedule needs them."""
        probe = torch.zeros(1, device=self.device)
        send_first = is_even(self.stage_id)
        for do_send in ((True, False) if send_first else (False, True)):
            if do_send and self._down is not None:
                p2p.send(probe, self.next_stage)
            if not do_send and self._up is not None:
                p2p.recv(probe, self.prev_stage)

    # ----------------------------------------------------------------------------------- basics
    def set_has_attention_mask(self, value):
        assert isinstance(value, bool)
        self.has_attention_mask = value

    def _build_data_iter(self, dataset):
        from ..dataloader import RepeatingLoader
        sampler = torch.utils.data.distributed.DistributedSampler(dataset, num_replicas=self.dp_world_size,
                                                                  rank=self.mpu.get_data_parallel_rank(), shuffle=False)
        self.set_dataloader(RepeatingLoader(self.deepspeed_io(dataset, data_sampler=sampler)))

    def is_first_stage(self):
        return self.stage_id == 0

    def is_last_stage(self):
        return self.stage_id == self.num_stages - 1

    def is_gradient_accumulation_boundary(self):
        return self._force_grad_boundary

    def _feeds_data(self) -> bool:
        return self.is_first_stage() or self.is_last_stage()      # inputs enter at the top, labels at the bottom

    def set_dataloader(self, loader):
        if self._feeds_data():
            self.training_dataloader, self.data_iterator = loader, iter(loader)

    def set_dataiterator(self, iterator):
        if self._feeds_data():
            self.training_dataloader, self.data_iterator = None, iterator

    def set_batch_fn(self, fn):
        self.batch_fn = fn

    def get_additional_losses(self):
        return getattr(self, "agg_additional_losses", None)

    def tput_log(self, *msg):
        if self.global_rank == 0:
            print(*msg)

    def log_for_device(self, *msg):
        """Print from global rank 0, or from the ranks listed in ``SXB_PIPE_LOG_RANKS``."""
        ranks = os.environ.get("SXB_PIPE_LOG_RANKS")
        if self.global_rank in ({int(r) for r in ranks.split(",")} if ranks else {0}):
            print(f"RANK={self.global_rank} PIPE-ID={self.stage_id} DATA-ID={self.grid.data_parallel_id} "
                  f"STEP-ID={self.log_batch_step_id} ::", *msg, flush=True)

    def mem_status(self, msg, print_rank=-1, reset_max=False):
        """Device memory now, and the change since the previous call, on ``print_rank`` (-1: rank 0)."""
        acc = get_accelerator()
        if not acc.is_available() or self.global_rank != (0 if print_rank == -1 else print_rank):
            return
        acc.synchronize()
        if reset_max:
            acc.reset_max_memory_cached()
            acc.reset_max_memory_allocated()
        now = (acc.memory_allocated(), acc.memory_cached())
        before = getattr(self, "_mem_status_prev", (0, 0))
        self._mem_status_prev = now
        gb = 1024 ** 3
        print(f"RANK={self.global_rank} STAGE={self.stage_id} STEP={self.global_steps} MEMSTATS {msg} "
              f"alloc={now[0] / gb:0.4f}GB ({(now[0] - before[0]) / gb:+0.4f}) max={acc.max_memory_allocated() / gb:0.4f}GB "
              f"cache={now[1] / gb:0.4f}GB ({(now[1] - before[1]) / gb:+0.4f}) max={acc.max_memory_cached() / gb:0.4f}GB",
              flush=True)

    def forward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def backward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def step(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def reset_activation_shape(self):
        """Forget the cached shape headers (call when the micro-batch shape changes without ``dynamic_shape``)."""
        for w in (self._up, self._down):
            if w is not None:
                w.reset()

    # ---------------------------------------------------------------------------- train / eval
    def train_batch(self, data_iter=None):
        """One batch = ``micro_batches`` forward/backward passes in 1F1B order, gradient reduction, optimizer step.
        Returns the batch loss (mean over micro-batches and over the data-parallel group) on every stage."""
        if not torch.is_grad_enabled():
            raise RuntimeError("train_batch() requires gradients enabled. Use eval_batch() instead.")
        if self.curriculum_enabled_legacy():
            self.curriculum_scheduler_legacy.update_difficulty(self.global_steps + 1)
        if data_iter is not None:
            self.set_dataiterator(data_iter)
        self.module.train()
        self.total_loss, self._compute_loss = None, True
        self.timers("train_batch").start()
        self._run(schedule.TrainSchedule(micro_batches=self.micro_batches, stages=self.num_stages, stage_id=self.stage_id))
        with torch.no_grad():
            self.agg_train_loss = self._share_batch_loss()
        self.timers("train_batch").stop()
        if self.global_rank == 0 and self.global_steps % self.steps_per_print() == 0:
            secs = self.timers("train_batch").elapsed(reset=True) / 1000.0
            if secs > 0:
                n = self.steps_per_print()
                print(f"steps: {self.global_steps} loss: {float(self.agg_train_loss):0.4f} iter time (s): {secs / n:0.3f} "
                      f"samples/sec: {self.train_batch_size() * n / secs:0.3f}")
        if self.monitor.enabled and self.global_rank == 0:
            self.monitor.write_events([("Train/Samples/train_loss", float(self.agg_train_loss), self.global_samples)])
        return self.agg_train_loss

    def eval_batch(self, data_iter, return_logits=False, compute_loss=True, reduce_output="avg", bcast_loss=True,
                   num_micro_batches=None):
        self.eval_return_logits, self._compute_loss = return_logits, compute_loss
        self.module.eval()
        saved_iter = self.data_iterator
        self.set_dataiterator(data_iter)
        n = self.micro_batches if num_micro_batches is None else num_micro_batches
        dist.barrier()
        with torch.no_grad():
            self._run(schedule.InferenceSchedule(micro_batches=n, stages=self.num_stages, stage_id=self.stage_id))
        result = None
        if self.is_last_stage():
            result = self._reduce_outputs(self.fwd_outputs, reduce=reduce_output, micro_batches=n)
        if compute_loss and (bcast_loss or self.monitor.enabled):
            result = self._bcast_pipe_scalar(result)
        if self.monitor.enabled and self.global_rank == 0 and result is not None:
            self.monitor.write_events([("Train/Samples/eval_loss", float(result), self.global_samples)])
        self.set_dataiterator(saved_iter)
        self.module.train()
        if return_logits:
            logits, self.outputs, self.eval_return_logits = self.outputs, None, False
            return result, logits
        return result

    def _run(self, plan) -> None:
        need = plan.num_pipe_buffers()
        while len(self._slots) < need:
            self._slots.append(_Slot())
        self.num_pipe_buffers = len(self._slots)
        self.fwd_outputs = []
        for commands in plan:
            for cmd in commands:
                handler = self._handlers.get(type(cmd))
                if handler is None:
                    raise RuntimeError(f"{type(self).__name__} does not understand instruction {cmd!r}")
                handler(**cmd.kwargs)

    # ------------------------------------------------------------------------------- reductions
    def _reduce_outputs(self, outputs, reduce="avg", reduce_dp=True, micro_batches=None):
        if reduce is None:
            return outputs
        if reduce.lower() != "avg":
            raise NotImplementedError(f"reduction type {reduce} not supported.")
        n = micro_batches or self.micro_batches
        single = torch.is_tensor(outputs[0])
        columns = [[o] for o in outputs] if single else [list(o) for o in outputs]
        means = [sum(col[i] for col in columns) / n for i in range(len(columns[0]))]
        if reduce_dp and self.is_data_parallel:
            for m in means:
                dist.all_reduce(m, group=self.mpu.get_data_parallel_group())
                m /= self.dp_world_size
        return means[0] if single else means

    def _bcast_pipe_scalar(self, data, src_rank=None, dtype=torch.float32):
        src = self.grid.stage_to_global(self.num_stages - 1) if src_rank is None else src_rank
        assert src in self.grid.pp_group
        value = (data.clone().detach().type(dtype).to(self.device) if self.global_rank == src
                 else torch.zeros((), dtype=dtype, device=self.device))
        dist.broadcast(tensor=value, src=src, group=self.mpu.get_pipe_parallel_group())
        return value

    def _share_batch_loss(self):
        """[this replica's loss, data-parallel mean] computed on the last stage and broadcast up the pipe."""
        pair = torch.zeros(2, device=self.device)
        last = self.grid.stage_to_global(self.num_stages - 1)
        if self.is_last_stage():
            mine = self._scale_loss_by_gas(self.total_loss).clone().detach().float()
            mean = mine.clone()
            if self.is_data_parallel:
                dist.all_reduce(mean, group=self.mpu.get_data_parallel_group())
                mean /= self.dp_world_size
            pair = torch.stack([mine, mean])
        if self.is_pipe_parallel:
            dist.broadcast(tensor=pair, src=last, group=self.mpu.get_pipe_parallel_group())
        self.dp_group_loss = pair[0].clone()
        return pair[1].clone()

    # -------------------------------------------------------------------- instruction handlers
    def _next_batch(self):
        batch = next(self.data_iterator) if self.data_iterator is not None else None
        return self.batch_fn(batch) if self.batch_fn else batch

    def _to_device(self, x: Tensors, track_grad: bool) -> Tensors:
        def one(t):
            t = t.clone().detach().to(self.device)
            if track_grad:
                t.requires_grad = t.is_floating_point()
            return t
        return one(x) if torch.is_tensor(x) else tuple(one(t) for t in x)

    def _load_micro_batch(self, buffer_id):
        batch = self._next_batch()
        slot = self._slots[buffer_id]
        if self.is_first_stage():
            slot.inputs = self._to_device(batch[0], track_grad=True)
        if self.is_last_stage():
            slot.labels = self._to_device(batch[1], track_grad=False)

    def _forward_pass(self, buffer_id):
        self.tput_timer.start()
        slot = self._slots[buffer_id]
        x = slot.inputs
        x = tuple(t.clone() for t in x) if isinstance(x, tuple) else x.clone()
        _F.open_use_window()           # gradient-use accounting counts ONE forward per backward (ZeRO ready hooks)
        slot.outputs = DeepSpeedEngine.forward(self, x)
        if not self.is_last_stage():
            return
        if self._compute_loss and self.module.loss_fn is not None:
            self.loss = self.module.loss_fn(slot.outputs, slot.labels)
        else:
            self.loss = slot.outputs
        if self.eval_return_logits:
            self.outputs = slot.outputs
        parts = _as_list(self.loss)
        if not self.module.training:
            self.fwd_outputs.append(self.loss.detach() if torch.is_tensor(self.loss) else [l.detach() for l in parts])
        if self.total_loss is None:
            self.total_loss = torch.zeros_like(self.loss) if torch.is_tensor(self.loss) else [torch.zeros_like(l) for l in parts]
        if torch.is_tensor(self.loss):
            self.total_loss += self.loss.detach()
        else:
            for acc, l in zip(self.total_loss, parts):
                acc += l.detach()

    def _backward_pass(self, buffer_id):
        if self.optimizer is None:
            raise AssertionError("must provide optimizer during init in order to use backward")
        slot = self._slots[buffer_id]
        opt = self.optimizer
        sharded = hasattr(opt, "_pre_backward")
        if hasattr(opt, "set_gradient_accumulation_boundary"):
            opt.set_gradient_accumulation_boundary(False)       # the step is an explicit schedule instruction
        if self.is_last_stage():
            loss = self._scale_loss_by_gas(self.loss) if self.micro_batches > 1 else self.loss
            if sharded or (hasattr(opt, "backward") and not isinstance(opt, torch.optim.Optimizer)):
                opt.backward(loss)
            else:
                loss.backward()
            self._close_micro_step()
            return
        outs = [t for t in _as_list(slot.outputs) if torch.is_tensor(t) and t.is_floating_point()]
        grads = slot.out_grads
        if len(outs) != len(grads):
            raise PipelineError(f"{len(outs)} differentiable outputs but {len(grads)} gradients arrived")
        if sharded:
            opt.micro_step_id += 1
            opt._pre_backward()
        torch.autograd.backward(tensors=outs, grad_tensors=list(grads))
        if sharded:
            opt._post_backward()
        self._close_micro_step()
        slot.outputs = slot.out_grads = None

    def _close_micro_step(self):
        """ZeRO-2 reduces every micro-batch's gradients into its shards: flush what the ready hooks have not launched
        yet and re-arm them for the next micro-batch (stage 3 does this per unit inside ``_post_backward``)."""
        if self.zero_optimization_stage() == 2:
            self.optimizer.overlapping_partition_gradients_reduce_epilogue()

    def _send_activations(self, buffer_id):
        self._down.send(self._slots[buffer_id].outputs)

    def _recv_activations(self, buffer_id):
        got = self._up.recv()

        def leaf(t):
            t = t.detach()
            t.requires_grad = t.is_floating_point()
            return t
        self._slots[buffer_id].inputs = leaf(got) if torch.is_tensor(got) else tuple(leaf(t) for t in got)

    def _send_grads(self, buffer_id):
        slot = self._slots[buffer_id]
        for t in _as_list(slot.inputs):
            if t.is_floating_point():
                p2p.send((t.grad if t.grad is not None else torch.zeros_like(t)).contiguous(), self.prev_stage)
        slot.inputs = None

    def _recv_grads(self, buffer_id):
        slot = self._slots[buffer_id]
        templates = [t for t in _as_list(slot.outputs) if torch.is_tensor(t) and t.is_floating_point()]
        slot.out_grads = self._down.recv_like(templates)

    def _reduce_tied_grads(self):
        for weight, group in self.module.get_tied_weights_and_groups():
            grad = getattr(weight, "_sxb_main_grad", None)
            grad = grad if grad is not None else weight.grad
            if grad is not None:
                dist.all_reduce(grad, group=group)

    def _reduce_grads(self):
        self._force_grad_boundary = True
        if self.pipeline_enable_backward_allreduce:
            if hasattr(self.optimizer, "set_gradient_accumulation_boundary"):
                self.optimizer.set_gradient_accumulation_boundary(True)
            self.allreduce_gradients(bucket_size=MEMORY_OPT_ALLREDUCE_SIZE)
        self._force_grad_boundary = False

    def _optimizer_step(self, lr_kwargs=None):
        self._force_grad_boundary = True
        self._take_model_step(lr_kwargs)
        self._force_grad_boundary = False
        self.micro_steps += self.micro_batches
        self.tput_timer.stop(global_step=True)

    # ------------------------------------------------------------------------------ checkpoints
    def module_state_dict(self, exclude_frozen_parameters=False):
        if self._curr_ckpt_path is None:
            raise AssertionError("PipelineEngine expects module_state_dict() to be called from save_checkpoint()")
        self.module.save_state_dict(self._curr_ckpt_path, checkpoint_engine=self.checkpoint_engine,
                                    exclude_frozen_params=exclude_frozen_parameters)
        return None       # layers are written as per-layer files, not into the model-states file

    def load_module_state_dict(self, checkpoint, strict=True, custom_load_fn=None, fetch_z3_params=False):
        if custom_load_fn is not None:
            raise AssertionError("custom_load_fn not supported w. pipeline parallelism")
        state = checkpoint if self.has_moe_layers else checkpoint.get("module")
        if state is not None and not isinstance(state, str):
            super().load_module_state_dict(state, strict)
        else:
            self.module.load_state_dir(load_dir=self._curr_ckpt_path, strict=strict, checkpoint_engine=self.checkpoint_engine)
