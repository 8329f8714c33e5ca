"""Column-parallel linear whose backward overlaps the input-gradient all-reduce with the weight-gradient GEMM
(reference: ``runtime/domino/async_linear.py:18-93``).  The all-reduce handle is parked in ``handle_dic`` under the
micro-batch key so that the Domino layer can wait for it as late as possible."""
from __future__ import annotations

import torch
import torch.nn as nn

from ... import comm as dist
from ...ops import functional as F


class DominoAsyncColumnParallelLinearImpl(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inp, weight, bias, handle_dic, h_id, group):
        ctx.save_for_backward(inp, weight)
        ctx.use_bias = bias is not None
        ctx.handle_dic, ctx.h_id, ctx.group = handle_dic, h_id, group
        out = F.matmul_nt(inp, weight)
        return out + bias if bias is not None else out

    @staticmethod
    def backward(ctx, grad_output):
        inp, weight = ctx.saved_tensors
        grad_input = F.matmul_nn(grad_output, weight)
        handle = None
        if ctx.group is not None and dist.is_initialized() and dist.get_world_size(group=ctx.group) > 1:
            handle = dist.all_reduce(grad_input, group=ctx.group, async_op=True)
            if ctx.handle_dic is not None:
                ctx.handle_dic[ctx.h_id] = handle
        g2 = grad_output.reshape(-1, grad_output.shape[-1])
        grad_weight = F.matmul_tn(grad_output, inp)
        grad_bias = g2.sum(dim=0) if ctx.use_bias else None
        if handle is not None and ctx.handle_dic is None:
            handle.wait()
        return grad_input, grad_weight, grad_bias, None, None, None


class DominoAsyncColumnParallelLinear(nn.Module):
    def __init__(self, input_size, output_size, _tp_group, config=None, init_method=nn.init.xavier_normal_, bias=True):
        super().__init__()
        self.tp_group = _tp_group
        tp = dist.get_world_size(group=_tp_group) if (_tp_group is not None and dist.is_initialized()) else 1
        assert output_size % tp == 0
        self.weight = nn.Parameter(torch.empty(output_size // tp, input_size))
        init_method(self.weight)
        self.bias = nn.Parameter(torch.zeros(output_size // tp)) if bias else None

    def forward(self, inp, handle_dic=None, h_id=None):
        return DominoAsyncColumnParallelLinearImpl.apply(inp, self.weight, self.bias, handle_dic, h_id, self.tp_group)


class RowParallelLinearNoComm(nn.Module):
    """Row-parallel linear that does NOT reduce its output: ``forward`` returns ``(partial_output, bias)`` and the caller
    (the Domino schedule) all-reduces the partial result asynchronously and adds the bias after the wait (reference
    ``async_linear.py:95-140``).  Weight ``[output_size, input_size / tp]``; the bias is replicated."""

    def __init__(self, input_size: int, output_size: int, _tp_group=None, config=None,
                 init_method=nn.init.xavier_normal_, bias: bool = True, skip_bias_add: bool = True):
        super().__init__()
        tp = dist.get_world_size(group=_tp_group) if (_tp_group is not None and dist.is_initialized()) else 1
        assert input_size % tp == 0
        self.weight = nn.Parameter(torch.empty(output_size, input_size // tp))
        init_method(self.weight)
        self.bias = nn.Parameter(torch.zeros(output_size)) if bias else None
        self.skip_bias_add = skip_bias_add

    def forward(self, input_):
        out = _PartialLinear.apply(input_, self.weight)
        if self.skip_bias_add or self.bias is None:
            return out, self.bias
        return out + self.bias, None


class _PartialLinear(torch.autograd.Function):
    """``x W^T`` on this framework's GEMM with explicit dgrad / wgrad."""

    @staticmethod
    def forward(ctx, x, weight):
        ctx.save_for_backward(x, weight)
        return F.matmul_nt(x, weight)

    @staticmethod
    def backward(ctx, g):
        x, w = ctx.saved_tensors
        return F.matmul_nn(g, w), F.matmul_tn(g, x)
