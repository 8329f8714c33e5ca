"""Pipeline schedules as instruction streams.

Reference: deepspeed/runtime/pipe/schedule.py — ``TrainSchedule`` (synchronous 1F1B, ``2*(M+S-1)`` steps,
even/odd stage interleave) :189-298, ``InferenceSchedule`` :135, ``DataParallelSchedule`` :301,
instruction classes :327-486, ``num_pipe_buffers = max(2, min(S - stage_id, M))`` :247-256.
"""
from __future__ import annotations

from abc import ABC, abstractmethod

from ..utils import call_to_str


class PipeInstruction:
    def __init__(self, **kwargs):
        self.name = self.__class__.__name__
        self.kwargs = kwargs
        for k, v in kwargs.items():
            setattr(self, k, v)

    def __repr__(self):
        return call_to_str(self.name, **self.kwargs)

    def __eq__(self, other):
        return type(self) is type(other) and self.kwargs == other.kwargs


class OptimizerStep(PipeInstruction): ...
class ReduceGrads(PipeInstruction): ...
class ReduceTiedGrads(PipeInstruction): ...


class BufferOpInstruction(PipeInstruction):
    def __init__(self, buffer_id, **kwargs):
        super().__init__(buffer_id=buffer_id, **kwargs)


class LoadMicroBatch(BufferOpInstruction): ...
class ForwardPass(BufferOpInstruction): ...
class BackwardPass(BufferOpInstruction): ...
class SendActivation(BufferOpInstruction): ...
class RecvActivation(BufferOpInstruction): ...
class SendGrad(BufferOpInstruction): ...
class RecvGrad(BufferOpInstruction): ...


class PipeSchedule(ABC):
    def __init__(self, micro_batches, stages, stage_id):
        self.micro_batches = micro_batches
        self.stages = stages
        self.stage_id = stage_id
        self.prev_stage = stage_id - 1
        self.next_stage = stage_id + 1

    @abstractmethod
    def steps(self):
        ...

    def num_pipe_buffers(self):
        return self.micro_batches

    def _valid_micro_batch(self, mb):
        return 0 <= mb < self.micro_batches

    def _valid_stage(self, s):
        return 0 <= s < self.stages

    @property
    def stage(self): return self.stage_id
    @property
    def num_stages(self): return self.stages
    @property
    def num_micro_batches(self): return self.micro_batches
    @property
    def is_first_stage(self): return self.stage_id == 0
    @property
    def is_last_stage(self): return self.stage_id == self.stages - 1

    def _buffer_idx(self, mb):
        assert self._valid_micro_batch(mb)
        return mb % self.num_pipe_buffers()

    def __iter__(self):
        self.it = None
        return self

    def __next__(self):
        if self.it is None:
            self.it = self.steps()
        return next(self.it)


class InferenceSchedule(PipeSchedule):
    """Forward-only pipelining with two alternating buffers."""

    def steps(self):
        total = self.micro_batches + self.stages - 1
        for step in range(total):
            cmds = []
            mb = step - self.stage_id
            recv_buf, send_buf = (step % 2, (step + 1) % 2) if self.stage_id % 2 == 0 else ((step + 1) % 2, step % 2)
            if (self.is_first_stage or self.is_last_stage) and self._valid_micro_batch(mb):
                cmds.append(LoadMicroBatch(recv_buf))
            if self.stage_id % 2 == 0:
                if self._valid_stage(self.next_stage) and self._valid_micro_batch(mb - 1):
                    cmds.append(SendActivation(send_buf))
                if self._valid_stage(self.prev_stage) and self._valid_micro_batch(mb):
                    cmds.append(RecvActivation(recv_buf))
            else:
                if self._valid_stage(self.prev_stage) and self._valid_micro_batch(mb):
                    cmds.append(RecvActivation(recv_buf))
                if self._valid_stage(self.next_stage) and self._valid_micro_batch(mb - 1):
                    cmds.append(SendActivation(send_buf))
            if self._valid_micro_batch(mb):
                cmds.append(ForwardPass(recv_buf))
            yield cmds

    def num_pipe_buffers(self):
        return 2


class TrainSchedule(PipeSchedule):
    """Synchronous 1F1B: each stage alternates forward and backward slots; even and odd stages are
    offset by one slot so every send meets a posted receive."""

    def steps(self):
        prev_mb = -1
        total = 2 * (self.micro_batches + self.stages - 1)
        for step in range(total):
            mb, is_forward = self._step_to_micro_batch(step)
            prev_buf = self._buffer_idx(prev_mb) if self._valid_micro_batch(prev_mb) else None
            cur_buf = self._buffer_idx(mb) if self._valid_micro_batch(mb) else None
            cmds = []
            if is_forward:
                if self._valid_micro_batch(prev_mb) and self._valid_stage(self.prev_stage):
                    cmds.append(SendGrad(prev_buf))
                if self._valid_micro_batch(mb) and self._valid_stage(self.prev_stage):
                    cmds.append(RecvActivation(cur_buf))
            else:
                if self._valid_micro_batch(mb) and self._valid_stage(self.next_stage):
                    cmds.append(RecvGrad(cur_buf))
                if self._valid_micro_batch(prev_mb) and self._valid_stage(self.next_stage):
                    cmds.append(SendActivation(prev_buf))
            if (self.stage_id == 0 or self.stage_id == self.stages - 1) and is_forward and self._valid_micro_batch(mb):
                cmds.append(LoadMicroBatch(cur_buf))
            if self._valid_micro_batch(mb):
                cmds.append(ForwardPass(cur_buf) if is_forward else BackwardPass(cur_buf))
            if step == total - 1:
                cmds += [ReduceTiedGrads(), ReduceGrads(), OptimizerStep()]
            prev_mb = mb
            yield cmds

    def num_pipe_buffers(self):
        return max(2, min(self.stages - self.stage_id, self.micro_batches))

    def _step_to_micro_batch(self, step):
        even_step, even_stage = step % 2 == 0, self.stage_id % 2 == 0
        if even_step == even_stage:      # forward slot
            base = step // 2 if even_step else (step - 1) // 2
            return base - self.stage_id // 2, True
        # backward slot
        base = step // 2 if even_step else (step - 1) // 2
        if even_step:   # odd stage
            return base - self.stages + (self.stage_id + 1) // 2, False
        return base - self.stages + 1 + self.stage_id // 2, False


class DataParallelSchedule(PipeSchedule):
    """Plain gradient accumulation expressed as a schedule (single stage)."""

    def steps(self):
        for step in range(self.micro_batches):
            cmds = [LoadMicroBatch(buffer_id=0), ForwardPass(buffer_id=0), BackwardPass(buffer_id=0)]
            if step == self.micro_batches - 1:
                cmds += [ReduceGrads(), OptimizerStep()]
            yield cmds

    def num_pipe_buffers(self):
        return 1
