"""Fake-quantisation operators with straight-through gradients, and the top-k mask used by movement pruning.

All four quantisers (symmetric / asymmetric uniform, ternary, binary) share one autograd shell: reshape to
``[num_groups, -1]``, compute the per-group reconstruction in :meth:`_reconstruct`, reshape back; the backward pass
hands the incoming gradient to the input untouched.  Call signature as in the reference
(``deepspeed/compression/utils.py``): ``Quantizer.apply(input, num_bits, min_value, max_value, num_groups)``."""
import torch
from torch import autograd


class TopKBinarizer(autograd.Function):
    """``mask[i] = 1`` for the ``threshold`` fraction of entries with the largest score, else 0.  With ``sigmoid`` the
    threshold is a learnable logit.  Gradient: straight through to the scores; to the threshold, the masked sum."""

    @staticmethod
    def forward(ctx, inputs: torch.Tensor, threshold, sigmoid: bool):
        frac = torch.sigmoid(threshold).item() if sigmoid else float(threshold)
        keep = int(frac * inputs.numel())
        mask = torch.zeros_like(inputs)
        if keep > 0:
            top = torch.topk(inputs.flatten(), keep, sorted=False).indices
            mask.view(-1)[top] = 1.0
        ctx.learn_threshold = bool(sigmoid)
        ctx.save_for_backward(mask)
        return mask

    @staticmethod
    def backward(ctx, grad_mask):
        mask, = ctx.saved_tensors
        d_threshold = (grad_mask * mask).sum().view(-1) if ctx.learn_threshold else None
        return grad_mask.clone(), d_threshold, None


class _StraightThrough(autograd.Function):
    """forward = per-group reconstruction, backward = identity on the input."""

    @staticmethod
    def _reconstruct(x, num_bits, lo, hi):          # x: [groups, n]
        raise NotImplementedError

    @classmethod
    def _check_range(cls, lo, hi, num_groups):
        both_none, both_set = lo is None and hi is None, lo is not None and hi is not None
        if not (both_none or (both_set and num_groups == 1)):
            raise AssertionError("min_value / max_value go together and require num_groups == 1")

    @classmethod
    def forward(cls, ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        cls._check_range(min_value, max_value, num_groups)
        grouped = input.reshape(num_groups, -1)
        return cls._reconstruct(grouped, num_bits, min_value, max_value).reshape(input.shape).contiguous()

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output.clone(), None, None, None, None


def _safe(scale):
    return torch.where(scale == 0, torch.ones_like(scale), scale)


class SymQuantizer(_StraightThrough):
    """Uniform levels symmetric around zero: step = 2 * max|x| / 2^bits per group (or from the given range)."""

    @staticmethod
    def _reconstruct(x, num_bits, lo, hi):
        levels = 2 ** num_bits
        peak = x.abs().amax(dim=-1, keepdim=True) if lo is None else torch.max(lo.abs(), hi).view(-1)
        step = _safe(2 * peak / levels)
        return (x / step).round().clamp(-levels // 2, levels // 2 - 1) * step


class AsymQuantizer(_StraightThrough):
    """Uniform levels between the group's min and max (or the given range), zero point snapped to the grid."""

    @staticmethod
    def _reconstruct(x, num_bits, lo, hi):
        levels = 2 ** num_bits
        if lo is None:
            lo, hi = x.amin(dim=-1, keepdim=True), x.amax(dim=-1, keepdim=True)
        step = _safe((hi - lo) / levels)
        zero = (lo / step).round() * step
        return ((x - zero) / step).round().clamp(0, levels - 1) * step + zero


class TernaryQuantizer(_StraightThrough):
    """Ternary weight networks: {-a, 0, +a} with threshold 0.7 * mean|x| and a = mean of the surviving magnitudes."""

    @classmethod
    def _check_range(cls, lo, hi, num_groups):
        if lo is not None or hi is not None:
            raise AssertionError("the ternary quantiser takes no value range")

    @staticmethod
    def _reconstruct(x, num_bits, lo, hi):
        cut = 0.7 * x.abs().mean(dim=1, keepdim=True)
        alive = (x.abs() > cut).to(x.dtype)
        a = (alive * x.abs()).sum(dim=1, keepdim=True) / alive.sum(dim=1, keepdim=True).clamp(min=1)
        return a * torch.sign(x) * alive


class BinaryQuantizer(_StraightThrough):
    """Binary weight networks: sign(x) * mean|x| per group."""

    @classmethod
    def _check_range(cls, lo, hi, num_groups):
        if lo is not None or hi is not None:
            raise AssertionError("the binary quantiser takes no value range")

    @staticmethod
    def _reconstruct(x, num_bits, lo, hi):
        return torch.sign(x) * x.abs().mean(dim=1, keepdim=True)
