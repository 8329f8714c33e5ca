"""Trampoline kept for API compatibility (reference: deepspeed/ops/adam/multi_tensor_apply.py)."""


class MultiTensorApply:
    def __init__(self, chunk_size):
        self.chunk_size = chunk_size

    def __call__(self, op, noop_flag_buffer, tensor_lists, *args):
        return op(self.chunk_size, noop_flag_buffer, tensor_lists, *args)
