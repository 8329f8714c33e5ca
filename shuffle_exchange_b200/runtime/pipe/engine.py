"""PipelineEngine — executes ``PipeSchedule`` instruction streams over a ``PipelineModule``.

Reference: deepspeed/runtime/pipe/engine.py (1426 lines): ``train_batch`` :338, ``eval_batch`` :428,
p2p meta handshake :929-1018, instruction handlers :710-1240, ``_INSTRUCTION_MAP`` :1396-1407,
ZeRO stage < 2 assertion :77-78.  ``forward/backward/step`` are disabled in favour of ``train_batch``.
Transport stays NCCL send/recv (north star); tensors of a micro-batch are exchanged as tuples after a
one-time shape/dtype handshake (or every time with ``dynamic_shape``).
"""
from __future__ import annotations

from types import MethodType
from typing import Optional

import torch

from ... import comm as dist
from ...device import get_accelerator
from ...utils.logging import logger
from ...utils.timer import ThroughputTimer
from ..engine import MEMORY_OPT_ALLREDUCE_SIZE, DeepSpeedEngine
from ..zero.config import ZeroStageEnum
from . import p2p, schedule
from .module import PipelineError, PipelineModule

TARGET_ID = -2
LOG_STAGE = -2
DATA_PARALLEL_ID = -2


def is_even(n):
    return n % 2 == 0


_DTYPES = [torch.float32, torch.float16, torch.bfloat16, torch.int64, torch.int32, torch.bool, torch.uint8,
           torch.int8, torch.float64, torch.int16]


class PipelineEngine(DeepSpeedEngine):
    ID_TO_DTYPE = _DTYPES
    DTYPE_TO_ID = {d: i for i, d in enumerate(_DTYPES)}

    def __init__(self, has_bool_tensors=False, *super_args, **super_kwargs):
        super().__init__(*super_args, **super_kwargs)
        assert isinstance(self.module, PipelineModule), "model must base PipelineModule"
        assert self.zero_optimization_stage() < ZeroStageEnum.gradients, \
            "ZeRO-2 and ZeRO-3 are incompatible with pipeline parallelism"
        self.enable_backward_allreduce = False
        self.has_bool_tensors = has_bool_tensors
        self.eval_return_logits = False
        self.outputs = None
        self.pipeline_enable_backward_allreduce = True
        self.log_batch_step_id = -1
        self.micro_batch_size = self.train_micro_batch_size_per_gpu()
        self.micro_batches = self.gradient_accumulation_steps()
        self.grid = self.module._grid
        self.global_rank = self.grid.get_global_rank()
        assert self.dp_world_size == self.grid.data_parallel_size
        assert self.train_batch_size() == self.micro_batch_size * self.micro_batches * self.grid.data_parallel_size
        self.num_stages = self.grid.pipe_parallel_size
        self.stage_id = self.grid.get_stage_id()
        self.prev_stage = self.stage_id - 1
        self.next_stage = self.stage_id + 1
        self.data_iterator = None
        self.batch_fn = None
        self._force_grad_boundary = False
        self.batch_timer = ThroughputTimer(self._config.timers_config.throughput,
                                           batch_size=self.train_batch_size(), logging_fn=self.tput_log,
                                           monitor_memory=False, steps_per_output=self.steps_per_print())
        if self.training_data:
            self._build_data_iter(self.training_data)
        self.is_pipe_parallel = self.grid.pipe_parallel_size > 1
        self.is_data_parallel = self.grid.data_parallel_size > 1
        self.is_model_parallel = self.grid.model_parallel_size > 1
        pc = self._config.pipeline
        self.is_pipe_partitioned = self.is_model_parallel and pc["pipe_partitioned"]
        self.is_grad_partitioned = self.is_model_parallel and pc["grad_partitioned"]
        self.dynamic_shape = self.module.dynamic_shape
        if pc["activation_checkpoint_interval"]:
            self.module.activation_checkpoint_interval = pc["activation_checkpoint_interval"]
        self.module.checkpoint_parallel_write_pipeline = self._config.checkpoint_parallel_write_pipeline
        if self.is_pipe_parallel:
            p2p.init_process_groups(self.grid)
        self.pipe_buffers = {"inputs": [], "labels": [], "outputs": [], "output_tensors": []}
        self.pipe_recv_buf = None
        self.grad_layer = None
        self.meta_buffer = None
        self.first_output_send = True
        self.first_gradient_send = True
        self.pipe_partition_input_meta_cache = None
        self.pipe_partition_output_meta_cache = None
        self.pipe_partition_grad_meta_cache = None
        self.grad_partition_grad_layer_meta_cache = None
        self.loss = torch.tensor(0.0).to(self.device)
        self.total_loss = None
        self.total_additional_losses = None
        self.agg_loss = torch.tensor(0.0, requires_grad=False).to(self.device)
        self.dp_group_loss = torch.tensor(0.0, requires_grad=False).to(self.device)
        self.agg_train_loss = None
        if self._config.pipeline["activation_checkpoint_interval"] > 0:
            self.module.activation_checkpoint_interval = self._config.pipeline["activation_checkpoint_interval"]
        if self.is_last_stage():
            self.loss_model = self.module.loss_fn
        self.has_attention_mask = self.module.__class__.__name__ == "GPT2ModelPipe"
        # stages agree on tensor metadata once; even/odd ordering avoids send/send deadlock
        if self.is_pipe_parallel:
            probe = torch.zeros(1, device=self.device)
            if is_even(self.stage_id):
                if not self.is_last_stage():
                    p2p.send(probe, self.next_stage)
                if not self.is_first_stage():
                    p2p.recv(probe, self.prev_stage)
            else:
                if not self.is_first_stage():
                    p2p.recv(probe, self.prev_stage)
                if not self.is_last_stage():
                    p2p.send(probe, self.next_stage)

    # ----------------------------------------------------------------------------------- basics
    def set_has_attention_mask(self, value):
        assert isinstance(value, bool)
        self.has_attention_mask = value

    def _build_data_iter(self, dataset):
        sampler = torch.utils.data.distributed.DistributedSampler(dataset, num_replicas=self.dp_world_size,
                                                                  rank=self.mpu.get_data_parallel_rank(), shuffle=False)
        from ..dataloader import RepeatingLoader
        pipe_dataloader = self.deepspeed_io(dataset, data_sampler=sampler)
        self.set_dataloader(RepeatingLoader(pipe_dataloader))

    def log_for_device(self, *msg):
        """Print from the ranks of one device position (reference ``pipe/engine.py:1351-1362``): the first stage's
        first data-parallel replica unless ``SXB_PIPE_LOG_RANKS`` names other global ranks."""
        import os
        ranks = os.environ.get("SXB_PIPE_LOG_RANKS")
        wanted = {int(r) for r in ranks.split(",")} if ranks else {0}
        if self.global_rank in wanted:
            print(f"RANK={self.global_rank} PIPE-ID={self.stage_id} DATA-ID={self.grid.data_parallel_id} "
                  f"MBATCH-ID={getattr(self, 'microbatch_id', 0)} STEP-ID={getattr(self, 'log_batch_step_id', -1)} ::",
                  *msg, flush=True)

    def mem_status(self, msg, print_rank=-1, reset_max=False):
        """Allocated / cached device memory and their deltas since the previous call, printed on ``print_rank`` (-1:
        rank 0)."""
        acc = get_accelerator()
        if not acc.is_available():
            return
        rank = self.global_rank
        if print_rank != -1 and rank != print_rank:
            return
        if print_rank == -1 and rank != 0:
            return
        acc.synchronize()
        if reset_max:
            acc.reset_max_memory_cached(); acc.reset_max_memory_allocated()
        new_alloced, new_cached = acc.memory_allocated(), acc.memory_cached()
        prev = getattr(self, "_mem_status_prev", (0, 0))
        self._mem_status_prev = (new_alloced, new_cached)
        gb = 1024 ** 3
        print(f"RANK={rank} STAGE={self.stage_id} STEP={self.global_steps} MEMSTATS", msg,
              f"current alloc={new_alloced / gb:0.4f}GB (delta={(new_alloced - prev[0]) / gb:0.4f}GB max="
              f"{acc.max_memory_allocated() / gb:0.4f}GB) current cache={new_cached / gb:0.4f}GB "
              f"(delta={(new_cached - prev[1]) / gb:0.4f}GB max={acc.max_memory_cached() / gb:0.4f}GB)", flush=True)

    def is_first_stage(self):
        return self.stage_id == 0

    def is_last_stage(self):
        return self.stage_id == self.num_stages - 1

    def is_gradient_accumulation_boundary(self):
        return self._force_grad_boundary

    def set_dataloader(self, loader):
        if self.is_first_stage() or self.is_last_stage():
            self.training_dataloader = loader
            self.data_iterator = iter(self.training_dataloader)

    def set_dataiterator(self, iterator):
        if self.is_first_stage() or self.is_last_stage():
            self.training_dataloader = None
            self.data_iterator = iterator

    def set_batch_fn(self, fn):
        self.batch_fn = fn

    def get_additional_losses(self):
        return self.agg_additional_losses if hasattr(self, "agg_additional_losses") else None

    def tput_log(self, *msg):
        if self.global_rank == 0:
            print(*msg)

    def forward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def backward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def step(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def _reserve_pipe_buffers(self, num_buffers):
        if self.pipe_buffers["inputs"] and len(self.pipe_buffers["inputs"]) >= num_buffers:
            return
        have = len(self.pipe_buffers["inputs"])
        for key in self.pipe_buffers:
            self.pipe_buffers[key].extend([None] * (num_buffers - have))
        self.num_pipe_buffers = num_buffers

    def reset_activation_shape(self):
        self.first_output_send = True
        self.pipe_recv_buf = None
        self.grad_layer = None
        self.meta_buffer = None

    # ------------------------------------------------------------------------- train / eval
    def train_batch(self, data_iter=None):
        """One full batch: ``micro_batches`` forward/backward passes under the 1F1B schedule, gradient
        reduction, optimizer step.  Returns the batch loss (mean over micro-batches, averaged over DP)."""
        if not torch._C.is_grad_enabled():
            raise RuntimeError("train_batch() requires gradients enabled. Use eval_batch() instead.")
        if self.curriculum_enabled_legacy():
            self.curriculum_scheduler_legacy.update_difficulty(self.global_steps + 1)
        if data_iter is not None:
            self.set_dataiterator(data_iter)
        self.module.train()
        self.total_loss = None
        self.total_additional_losses = None
        self._compute_loss = True
        self.timers("train_batch").start()
        sched = schedule.TrainSchedule(micro_batches=self.micro_batches, stages=self.num_stages, stage_id=self.stage_id)
        self._exec_schedule(sched)
        with torch.no_grad():
            self.agg_train_loss = self._aggregate_total_loss()
        self.timers("train_batch").stop()
        if self.global_steps % self.steps_per_print() == 0 and self.global_rank == 0:
            elapsed = self.timers("train_batch").elapsed(reset=True) / 1000.0
            if elapsed > 0:
                print(f"steps: {self.global_steps} loss: {float(self.agg_train_loss):0.4f} iter time (s): "
                      f"{elapsed / self.steps_per_print():0.3f} samples/sec: "
                      f"{self.train_batch_size() * self.steps_per_print() / elapsed:0.3f}")
        if self.monitor.enabled and self.global_rank == 0:
            self.monitor.write_events([("Train/Samples/train_loss", float(self.agg_train_loss), self.global_samples)])
        return self.agg_train_loss

    def eval_batch(self, data_iter, return_logits=False, compute_loss=True, reduce_output="avg", bcast_loss=True,
                   num_micro_batches=None):
        self.eval_return_logits = return_logits
        self.module.eval()
        self._compute_loss = compute_loss
        train_iterator = self.data_iterator
        self.set_dataiterator(data_iter)
        micro_batches = self.micro_batches if num_micro_batches is None else num_micro_batches
        sched = schedule.InferenceSchedule(micro_batches=micro_batches, stages=self.num_stages, stage_id=self.stage_id)
        dist.barrier()
        self.fwd_outputs = []
        with torch.no_grad():
            self._exec_schedule(sched)
        eval_output = None
        if self.is_last_stage():
            eval_output = self._reduce_outputs(self.fwd_outputs, reduce=reduce_output, micro_batches=micro_batches)
        if compute_loss and (bcast_loss or self.monitor.enabled):
            eval_output = self._bcast_pipe_scalar(eval_output)
        if self.monitor.enabled and self.global_rank == 0 and eval_output is not None:
            self.monitor.write_events([("Train/Samples/eval_loss", float(eval_output), self.global_samples)])
        self.set_dataiterator(train_iterator)
        self.module.train()
        if return_logits:
            outputs = self.outputs
            self.outputs = None
            self.eval_return_logits = False
            return eval_output, outputs
        return eval_output

    def _reduce_outputs(self, outputs, reduce="avg", reduce_dp=True, micro_batches=None):
        if reduce is None:
            return outputs
        if reduce.lower() != "avg":
            raise NotImplementedError(f"reduction type {reduce} not supported.")
        if torch.is_tensor(outputs[0]):
            reduced = sum(outputs)
        else:
            assert isinstance(outputs, (list, tuple))
            reduced = [torch.zeros_like(o) for o in outputs[0]]
            for out in outputs:
                for i, o in enumerate(out):
                    reduced[i] += o
        n = micro_batches or self.micro_batches
        if torch.is_tensor(reduced):
            reduced = reduced / n
        else:
            reduced = [r / n for r in reduced]
        if reduce_dp and self.is_data_parallel:
            if torch.is_tensor(reduced):
                dist.all_reduce(reduced, group=self.mpu.get_data_parallel_group())
                reduced /= self.dp_world_size
            else:
                for i in range(len(reduced)):
                    dist.all_reduce(reduced[i], group=self.mpu.get_data_parallel_group())
                    reduced[i] /= self.dp_world_size
        return reduced

    def _bcast_pipe_scalar(self, data, src_rank=None, dtype=torch.float32):
        if src_rank is None:
            src_rank = self.grid.stage_to_global(self.num_stages - 1)
        assert src_rank in self.grid.pp_group
        if self.global_rank == src_rank:
            result = data.clone().detach().type(dtype).to(self.device)
        else:
            result = torch.zeros(1, dtype=dtype, device=self.device).squeeze()
        dist.broadcast(tensor=result, src=src_rank, group=self.mpu.get_pipe_parallel_group())
        return result

    def _aggregate_total_loss(self):
        if self.is_last_stage():
            loss = self._scale_loss_by_gas(self.total_loss)
            self.dp_group_loss = loss.clone().detach()
            agg = self.dp_group_loss.clone().detach().float()
            if self.is_data_parallel:
                dist.all_reduce(agg, group=self.mpu.get_data_parallel_group())
                agg /= self.dp_world_size
            assert self.global_rank in self.grid.pp_group
            losses = torch.stack([self.dp_group_loss.float(), agg])
            if self.is_pipe_parallel:
                dist.broadcast(tensor=losses, src=self.global_rank, group=self.mpu.get_pipe_parallel_group())
        else:
            src_rank = self.grid.stage_to_global(self.num_stages - 1)
            losses = torch.zeros(2, device=self.device)
            dist.broadcast(tensor=losses, src=src_rank, group=self.grid.get_pipe_parallel_group())
            self.dp_group_loss = losses[0].clone().detach()
            agg = losses[1].clone().detach()
        return agg

    def _next_batch(self):
        batch = next(self.data_iterator) if self.data_iterator is not None else None
        if self.batch_fn:
            batch = self.batch_fn(batch)
        return batch

    # ------------------------------------------------------------------- instruction handlers
    def _exec_load_micro_batch(self, buffer_id):
        batch = self._next_batch()
        if self.is_first_stage():
            inputs = batch[0]
            if torch.is_tensor(inputs):
                loaded = inputs.clone().to(self.device).detach()
                loaded.requires_grad = loaded.is_floating_point()
            else:
                loaded = []
                for x in inputs:
                    x = x.clone().detach().to(self.device)
                    x.requires_grad = x.is_floating_point()
                    loaded.append(x)
                loaded = tuple(loaded)
            self.pipe_buffers["inputs"][buffer_id] = loaded
        if self.is_last_stage():
            labels = batch[1]
            if torch.is_tensor(labels):
                labels = labels.to(self.device).detach()
            else:
                labels = tuple(x.to(self.device).detach() for x in labels)
            self.pipe_buffers["labels"][buffer_id] = labels

    def _exec_forward_pass(self, buffer_id):
        self.tput_timer.start()
        inputs = self.pipe_buffers["inputs"][buffer_id]
        if isinstance(inputs, tuple):
            inputs = tuple(t.clone() if torch.is_tensor(t) else t for t in inputs)
        else:
            inputs = inputs.clone()
        outputs = DeepSpeedEngine.forward(self, inputs)
        self.pipe_buffers["outputs"][buffer_id] = outputs
        if self.is_last_stage():
            if self._compute_loss and self.module.loss_fn is not None:
                labels = self.pipe_buffers["labels"][buffer_id]
                self.loss = self.module.loss_fn(outputs, labels)
            else:
                self.loss = outputs
            if self.eval_return_logits:
                self.outputs = outputs
            if isinstance(self.loss, torch.Tensor):
                self.fwd_outputs.append(self.loss.detach()) if hasattr(self, "fwd_outputs") and not self.module.training else None
                if self.total_loss is None:
                    self.total_loss = torch.zeros_like(self.loss)
                self.total_loss += self.loss.detach()
            else:
                if hasattr(self, "fwd_outputs") and not self.module.training:
                    self.fwd_outputs.append([l.detach() for l in self.loss])
                if self.total_loss is None:
                    self.total_loss = [torch.zeros_like(l) for l in self.loss]
                for i, l in enumerate(self.loss):
                    self.total_loss[i] += l.detach()

    def _exec_backward_pass(self, buffer_id):
        assert self.optimizer is not None, "must provide optimizer during init in order to use backward"
        if self.is_last_stage():
            loss = self._scale_loss_by_gas(self.loss) if self.micro_batches > 1 else self.loss
            if hasattr(self.optimizer, "backward") and not isinstance(self.optimizer, torch.optim.Optimizer):
                self.optimizer.set_gradient_accumulation_boundary(False) if hasattr(self.optimizer, "set_gradient_accumulation_boundary") else None
                self.optimizer.backward(loss)
            else:
                loss.backward()
            self.micro_steps += 0
            return
        outputs = self.pipe_buffers["outputs"][buffer_id]
        grad_tensors = self.grad_layer
        if isinstance(outputs, tuple):
            out_tensors = [t for t in outputs if torch.is_tensor(t) and t.is_floating_point()]
            assert len(out_tensors) == len(grad_tensors)
        else:
            out_tensors, grad_tensors = [outputs], [grad_tensors[0] if isinstance(grad_tensors, (list, tuple)) else grad_tensors]
        if hasattr(self.optimizer, "_pre_backward"):
            self.optimizer.micro_step_id += 1
            self.optimizer._pre_backward()
        torch.autograd.backward(tensors=out_tensors, grad_tensors=list(grad_tensors))
        if hasattr(self.optimizer, "_post_backward"):
            self.optimizer._post_backward()
        self.pipe_buffers["output_tensors"][buffer_id] = None
        self.pipe_buffers["outputs"][buffer_id] = None

    # ---- tensor tuple transport with a metadata handshake
    def _send_tensor_meta(self, tensors, recv_stage):
        """[count, (dtype_id, ndim, *shape)...] as int64."""
        items = [tensors] if torch.is_tensor(tensors) else list(tensors)
        meta = [1 if torch.is_tensor(tensors) else 2, len(items)]
        for t in items:
            meta += [self.DTYPE_TO_ID[t.dtype], t.dim(), *t.shape]
        n = torch.tensor([len(meta)], dtype=torch.long, device=self.device)
        p2p.send(n, recv_stage)
        p2p.send(torch.tensor(meta, dtype=torch.long, device=self.device), recv_stage)

    def _recv_tensor_meta(self, send_stage):
        n = torch.zeros(1, dtype=torch.long, device=self.device)
        p2p.recv(n, send_stage)
        meta = torch.zeros(int(n.item()), dtype=torch.long, device=self.device)
        p2p.recv(meta, send_stage)
        meta = meta.tolist()
        kind, count = meta[0], meta[1]
        pos, bufs = 2, []
        for _ in range(count):
            dt, nd = self.ID_TO_DTYPE[meta[pos]], meta[pos + 1]
            shape = meta[pos + 2: pos + 2 + nd]
            pos += 2 + nd
            bufs.append(torch.zeros(shape, dtype=dt, device=self.device))
        return bufs[0] if kind == 1 else tuple(bufs)

    def _exec_send_activations(self, buffer_id):
        outputs = self.pipe_buffers["outputs"][buffer_id]
        if self.first_output_send or self.dynamic_shape:
            self.first_output_send = False
            self._send_tensor_meta(outputs, self.next_stage)
        for t in ([outputs] if torch.is_tensor(outputs) else outputs):
            p2p.send(t.contiguous(), self.next_stage)

    def _exec_recv_activations(self, buffer_id):
        if self.pipe_recv_buf is None or self.dynamic_shape:
            self.pipe_recv_buf = self._recv_tensor_meta(self.prev_stage)
        bufs = [self.pipe_recv_buf] if torch.is_tensor(self.pipe_recv_buf) else list(self.pipe_recv_buf)
        recvd = []
        for b in bufs:
            p2p.recv(b, self.prev_stage)
            t = b.clone().detach()
            t.requires_grad = t.is_floating_point()
            recvd.append(t)
        self.pipe_buffers["inputs"][buffer_id] = recvd[0] if torch.is_tensor(self.pipe_recv_buf) else tuple(recvd)

    def _exec_send_grads(self, buffer_id):
        inputs = self.pipe_buffers["inputs"][buffer_id]
        items = [inputs] if torch.is_tensor(inputs) else list(inputs)
        for t in items:
            if t.is_floating_point():
                g = t.grad if t.grad is not None else torch.zeros_like(t)
                p2p.send(g.contiguous(), self.prev_stage)
        self.pipe_buffers["inputs"][buffer_id] = None

    def _exec_recv_grads(self, buffer_id):
        outputs = self.pipe_buffers["outputs"][buffer_id]
        items = [outputs] if torch.is_tensor(outputs) else [t for t in outputs if torch.is_tensor(t)]
        if self.grad_layer is None or self.dynamic_shape:
            self.grad_layer = [torch.zeros_like(t) for t in items if t.is_floating_point()]
        for g in self.grad_layer:
            p2p.recv(g, self.next_stage)

    def _exec_reduce_tied_grads(self):
        weight_group_list = self.module.get_tied_weights_and_groups()
        for weight, group in weight_group_list:
            grad = weight._sxb_main_grad if getattr(weight, "_sxb_main_grad", None) is not None else weight.grad
            if grad is not None:
                dist.all_reduce(grad, group=group)

    def _exec_reduce_grads(self):
        self._force_grad_boundary = True
        if self.pipeline_enable_backward_allreduce:
            if hasattr(self.optimizer, "set_gradient_accumulation_boundary"):
                self.optimizer.set_gradient_accumulation_boundary(True)
            self.allreduce_gradients(bucket_size=MEMORY_OPT_ALLREDUCE_SIZE)
        self._force_grad_boundary = False

    def _exec_optimizer_step(self, lr_kwargs=None):
        self._force_grad_boundary = True
        self._take_model_step(lr_kwargs)
        self._force_grad_boundary = False
        self.micro_steps += self.micro_batches
        self.tput_timer.stop(global_step=True)

    _INSTRUCTION_MAP = {
        schedule.OptimizerStep: _exec_optimizer_step,
        schedule.ReduceGrads: _exec_reduce_grads,
        schedule.ReduceTiedGrads: _exec_reduce_tied_grads,
        schedule.LoadMicroBatch: _exec_load_micro_batch,
        schedule.ForwardPass: _exec_forward_pass,
        schedule.BackwardPass: _exec_backward_pass,
        schedule.SendActivation: _exec_send_activations,
        schedule.RecvActivation: _exec_recv_activations,
        schedule.SendGrad: _exec_send_grads,
        schedule.RecvGrad: _exec_recv_grads,
    }

    def _exec_schedule(self, pipe_schedule):
        self._reserve_pipe_buffers(pipe_schedule.num_pipe_buffers())
        self.fwd_outputs = []
        for step_cmds in pipe_schedule:
            for cmd in step_cmds:
                if type(cmd) not in self._INSTRUCTION_MAP:
                    raise RuntimeError(f"{self.__class__.__name__} does not understand instruction {repr(cmd)}")
                self._exec_instr = MethodType(self._INSTRUCTION_MAP[type(cmd)], self)
                self._exec_instr(**cmd.kwargs)

    # ---------------------------------------------------------------------------- checkpoints
    def module_state_dict(self, exclude_frozen_parameters=False):
        assert isinstance(self.module, PipelineModule)
        assert self._curr_ckpt_path is not None, "PipelineEngine expects module_state_dict() to be called from save_checkpoint()"
        self.module.save_state_dict(self._curr_ckpt_path, checkpoint_engine=self.checkpoint_engine,
                                    exclude_frozen_params=exclude_frozen_parameters)
        return None

    def load_module_state_dict(self, checkpoint, strict=True, custom_load_fn=None, fetch_z3_params=False):
        assert custom_load_fn is None, "custom_load_fn not supported w. pipeline parallelism"
        state_dict = checkpoint if self.has_moe_layers else checkpoint.get("module")
        if state_dict is not None and not isinstance(state_dict, str):
            super().load_module_state_dict(state_dict, strict)
            return
        self.module.load_state_dir(load_dir=self._curr_ckpt_path, strict=strict, checkpoint_engine=self.checkpoint_engine)
