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
    def __init__(self, cur_scale):
        self.cur_scale = cur_scale
        self.dynamic = False

    @property
    def loss_scale(self):
        return self.cur_scale

    def scale_gradient(self, module, grad_in, grad_out):
        return tuple(self.loss_scale * g for g in grad_in)

    def update_scale(self, overflow):
        pass

    def backward(self, loss, retain_graph=False):
        (loss * self.loss_scale).backward(retain_graph=retain_graph)


class LossScaler(LossScalerBase):
    """Static scale."""

    def has_overflow(self, params):
        return False

    @staticmethod
    def _has_inf_or_nan(x):
        return False


class DynamicLossScaler(LossScalerBase):
    def __init__(self, init_scale, scale_factor, scale_window, min_scale, delayed_shift, consecutive_hysteresis,
                 raise_error_at_min_scale, dtype):
        super().__init__(init_scale)
        self.cur_iter = 0
        self.last_overflow_iter = -1
        self.scale_factor = scale_factor
        self.scale_window = scale_window
        self.min_scale = min_scale
        self.delayed_shift = delayed_shift
        self.cur_hysteresis = delayed_shift
        self.consecutive_hysteresis = consecutive_hysteresis
        self.raise_error_at_min_scale = raise_error_at_min_scale
        self.dynamic = True
        self.dtype = dtype

    def has_overflow_serial(self, params):
        return any(p.grad is not None and self._has_inf_or_nan(p.grad.data) for p in params)

    @staticmethod
    def _has_inf_or_nan(x):
        try:
            s = float(x.float().sum())
        except RuntimeError as e:
            if "value cannot be converted" not in e.args[0]:
                raise
            return True
        return s in (float("inf"), -float("inf")) or s != s

    def update_scale(self, overflow):
        if overflow:
            if self.delayed_shift == 1 or self.cur_hysteresis == 1:
                if self.cur_scale == self.min_scale and self.raise_error_at_min_scale:
                    raise Exception("Current loss scale already at minimum - cannot decrease scale anymore. Exiting run.")
                nxt = max(self.cur_scale / self.scale_factor, self.min_scale)
                if self.dtype == torch.half:
                    logger.info(f"[sxb200] OVERFLOW! Skipping step. Attempted loss scale: {int(self.cur_scale)}, "
                                f"reducing to {int(nxt)}")
                self.cur_scale = nxt
            else:
                if self.dtype == torch.half:
                    logger.info(f"[sxb200] OVERFLOW! Skipping step. Attempted loss scale: {int(self.cur_scale)}, but "
                                f"hysteresis is {self.cur_hysteresis}. Reducing hysteresis to {self.cur_hysteresis - 1}")
                self.cur_hysteresis -= 1
            self.last_overflow_iter = self.cur_iter
        else:
            if self.consecutive_hysteresis:
                self.cur_hysteresis = self.delayed_shift
            stable = (self.cur_iter - self.last_overflow_iter) % self.scale_window == 0
            if stable:
                if not self.consecutive_hysteresis:
                    self.cur_hysteresis = self.delayed_shift
                self.cur_scale *= self.scale_factor
        self.cur_iter += 1


def CreateLossScaler(dtype, static_loss_scale, dynamic_scaling, dynamic_loss_args):
    if dtype == torch.half and dynamic_scaling:
        assert dynamic_loss_args is not None, "Dynamic loss scaling parameters must be defined."
        return DynamicLossScaler(dtype=dtype, scale_factor=2.0, raise_error_at_min_scale=True,
                                 init_scale=dynamic_loss_args.get(INITIAL_LOSS_SCALE, 2 ** 32),
                                 scale_window=dynamic_loss_args.get(SCALE_WINDOW, 1000),
                                 min_scale=dynamic_loss_args.get(MIN_LOSS_SCALE, 1),
                                 delayed_shift=dynamic_loss_args.get(DELAYED_SHIFT, 1),
                                 consecutive_hysteresis=dynamic_loss_args.get(CONSECUTIVE_HYSTERESIS, False))
    loss_scale_value = static_loss_scale if dtype == torch.half else 1.0
    return LossScaler(cur_scale=loss_scale_value)
