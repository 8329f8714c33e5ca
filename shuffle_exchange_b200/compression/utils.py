"""Straight-through quantisers and top-k binarizer (reference: ``deepspeed/compression/utils.py:13-189``)."""
import torch
from torch import autograd


class TopKBinarizer(autograd.Function):
    """mask = 1 for the ``threshold`` fraction of largest scores (``sigmoid`` maps a learnable threshold to (0,1));
    gradients pass straight through to the scores and (summed) to the threshold."""

    @staticmethod
    def forward(ctx, inputs: torch.Tensor, threshold: float, sigmoid: bool):
        if sigmoid:
            threshold = torch.sigmoid(threshold).item()
        ctx.sigmoid = sigmoid
        mask = inputs.clone()
        _, idx = inputs.flatten().sort(descending=True)
        j = int(threshold * inputs.numel())
        flat = mask.flatten()
        flat[idx[j:]] = 0.0
        flat[idx[:j]] = 1.0
        ctx.save_for_backward(mask)
        return mask

    @staticmethod
    def backward(ctx, gradOutput):
        mask, = ctx.saved_tensors
        if ctx.sigmoid:
            return gradOutput.clone(), ((gradOutput * mask).sum()).view(-1), None
        return gradOutput.clone(), None, None


class SymQuantizer(torch.autograd.Function):
    """Per-group symmetric uniform quantisation with a straight-through gradient."""

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert (min_value is None and max_value is None) or (min_value is not None and max_value is not None and num_groups == 1)
        q_range = 2 ** num_bits
        shape = input.shape
        if min_value is None:
            x = input.reshape(num_groups, -1)
            max_in = torch.amax(torch.abs(x), dim=-1).view(num_groups, -1)
        else:
            x = input.reshape(1, -1)
            max_in = torch.max(min_value.abs(), max_value).view(-1)
        scale = 2 * max_in / q_range
        scale = torch.where(scale == 0, torch.ones_like(scale), scale)
        out = (x / scale).round().clamp(-q_range // 2, q_range // 2 - 1) * scale
        return out.reshape(shape).contiguous()

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output.clone(), None, None, None, None


class AsymQuantizer(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert (min_value is None and max_value is None) or (min_value is not None and max_value is not None and num_groups == 1)
        q_range = 2 ** num_bits
        shape = input.shape
        x = input.reshape(num_groups, -1)
        if min_value is None:
            min_value, max_value = x.amin(dim=-1, keepdim=True), x.amax(dim=-1, keepdim=True)
        scale = (max_value - min_value) / q_range
        scale = torch.where(scale == 0, torch.ones_like(scale), scale)
        zero_point = (min_value / scale).round() * scale
        out = ((x - zero_point) / scale).round().clamp(0, q_range - 1) * scale + zero_point
        return out.reshape(shape).contiguous()

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output.clone(), None, None, None, None


class TernaryQuantizer(torch.autograd.Function):
    """{-a, 0, +a} with threshold 0.7 * mean|x| per group (TWN)."""

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert min_value is None and max_value is None
        x = input.reshape(num_groups, -1)
        n = x.shape[1]
        m = x.norm(p=1, dim=1).div(n)
        thres = (0.7 * m).view(-1, 1)
        pos, neg = (x > thres).type(input.type()), (x < -thres).type(input.type())
        mask = (x.abs() > thres).type(input.type())
        alpha = ((mask * x).abs().sum(dim=1) / mask.sum(dim=1).clamp(min=1)).view(-1, 1)
        return (alpha * pos - alpha * neg).reshape(input.shape).contiguous()

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output.clone(), None, None, None, None


class BinaryQuantizer(torch.autograd.Function):
    """sign(x) * mean|x| per group (BWN)."""

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert min_value is None and max_value is None
        x = input.reshape(num_groups, -1)
        n = x.shape[1]
        m = x.norm(p=1, dim=1, keepdim=True).div(n)
        return x.sign().mul(m).reshape(input.shape).contiguous()

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output.clone(), None, None, None, None
