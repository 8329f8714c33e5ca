"""Static and dynamic loss scaling (reference: deepspeed/runtime/fp16/loss_scaler.py:93-209;
``update_scale`` :171, factory ``CreateLossScaler`` :211).

Dynamic policy: on overflow divide the scale by ``scale_factor`` (never below ``min_scale``) once the
hysteresis budget (``delayed_shift``) is used up; after ``scale_window`` consecutive clean steps
multiply by ``scale_factor`` and (unless ``consecutive_hysteresis``) refill the hysteresis budget.
"""
from __future__ import annotations

import torch

from ...utils.logging import logger

INITIAL_LOSS_SCALE = "init_scale"
SCALE_WINDOW = "scale_window"
DELAYED_SHIFT = "delayed_shift"
CONSECUTIVE_HYSTERESIS = "consecutive_hysteresis"
MIN_LOSS_SCALE = "min_scale"


def to_python_float(t):
    return t.item() if hasattr(t, "item") else t


class LossScalerBase:
    """What the optimizers need from a scaler: the current scale, ``backward`` of the scaled loss, an ``update_scale``
    verdict per step.  Attribute names are the reference's (they are pickled into optimizer checkpoints)."""
    dynamic = False

    def __init__(self, cur_scale):
        self.cur_scale = cur_scale

    @property
    def loss_scale(self):
        return self.cur_scale

    def scale_gradient(self, module, grad_in, grad_out):
        return tuple(g * self.cur_scale for g in grad_in)

    def update_scale(self, overflow):
        return None

    def backward(self, loss, retain_graph=False):
        (loss * self.cur_scale).backward(retain_graph=retain_graph)


class LossScaler(LossScalerBase):
    """Constant scale; never reports an overflow."""

    def has_overflow(self, params):
        return False

    @staticmethod
    def _has_inf_or_nan(x):
        return False


def _non_finite(x) -> bool:
    """One reduction + one host read: the sum of a tensor is non-finite iff some element is (or it overflowed)."""
    return not bool(torch.isfinite(x.float().sum()))


class DynamicLossScaler(LossScalerBase):
    """Backs off on overflow, grows after a clean window, with a hysteresis budget that absorbs isolated overflows."""
    dynamic = True

    def __init__(self, init_scale, scale_factor, scale_window, min_scale, delayed_shift, consecutive_hysteresis,
                 raise_error_at_min_scale, dtype):
        super().__init__(init_scale)
        self.scale_factor, self.scale_window, self.min_scale = scale_factor, scale_window, min_scale
        self.delayed_shift = self.cur_hysteresis = delayed_shift
        self.consecutive_hysteresis = consecutive_hysteresis
        self.raise_error_at_min_scale = raise_error_at_min_scale
        self.dtype = dtype
        self.cur_iter, self.last_overflow_iter = 0, -1

    def has_overflow_serial(self, params):
        return any(p.grad is not None and _non_finite(p.grad.data) for p in params)

    _has_inf_or_nan = staticmethod(_non_finite)

    # ---- the two transitions
    def _overflowed(self):
        budget_left = self.delayed_shift != 1 and self.cur_hysteresis != 1
        if budget_left:
            self.cur_hysteresis -= 1
            self._say(f"hysteresis left: {self.cur_hysteresis}, scale kept at {int(self.cur_scale)}")
        else:
            if self.raise_error_at_min_scale and self.cur_scale == self.min_scale:
                raise Exception("Current loss scale already at minimum - cannot decrease scale anymore. Exiting run.")
            lowered = max(self.cur_scale / self.scale_factor, self.min_scale)
            self._say(f"scale {int(self.cur_scale)} -> {int(lowered)}")
            self.cur_scale = lowered
        self.last_overflow_iter = self.cur_iter

    def _clean(self):
        window_full = (self.cur_iter - self.last_overflow_iter) % self.scale_window == 0
        if self.consecutive_hysteresis or window_full:
            self.cur_hysteresis = self.delayed_shift          # refilled every clean step, or once per clean window
        if window_full:
            self.cur_scale *= self.scale_factor

    def _say(self, what):
        if self.dtype == torch.half:
            logger.info(f"[sxb200] gradient overflow, step skipped; {what}")

    def update_scale(self, overflow):
        (self._overflowed if overflow else self._clean)()
        self.cur_iter += 1


def CreateLossScaler(dtype, static_loss_scale, dynamic_scaling, dynamic_loss_args):
    """fp16 + dynamic -> :class:`DynamicLossScaler` from the ``dynamic_loss_args`` dict; fp16 static -> that scale;
    any other dtype -> scale 1."""
    if dtype != torch.half:
        return LossScaler(cur_scale=1.0)
    if not dynamic_scaling:
        return LossScaler(cur_scale=static_loss_scale)
    if dynamic_loss_args is None:
        raise AssertionError("Dynamic loss scaling parameters must be defined.")
    pick = dynamic_loss_args.get
    return DynamicLossScaler(init_scale=pick(INITIAL_LOSS_SCALE, 2 ** 32), scale_factor=2.0,
                             scale_window=pick(SCALE_WINDOW, 1000), min_scale=pick(MIN_LOSS_SCALE, 1),
                             delayed_shift=pick(DELAYED_SHIFT, 1),
                             consecutive_hysteresis=pick(CONSECUTIVE_HYSTERESIS, False),
                             raise_error_at_min_scale=True, dtype=dtype)
