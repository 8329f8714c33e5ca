"""Learning-rate schedules: LRRangeTest, OneCycle, WarmupLR, WarmupDecayLR, WarmupCosineLR.

Same names, constructor arguments and per-step values as the reference
(deepspeed/runtime/lr_schedules.py: LRRangeTest :273, OneCycle :371, WarmupLR :633,
WarmupDecayLR :726, WarmupCosineLR :777) built on one small base class
(:class:`_StepSchedule`) that owns the step counter, state dict and optimizer plumbing.
"""
from __future__ import annotations

import argparse
import math
from typing import List, Union

from torch.optim import Optimizer

from ..utils.logging import logger

LR_SCHEDULE = "lr_schedule"
LR_RANGE_TEST = "LRRangeTest"
ONE_CYCLE = "OneCycle"
WARMUP_LR = "WarmupLR"
WARMUP_DECAY_LR = "WarmupDecayLR"
WARMUP_COSINE_LR = "WarmupCosineLR"
VALID_LR_SCHEDULES = [LR_RANGE_TEST, ONE_CYCLE, WARMUP_LR, WARMUP_DECAY_LR, WARMUP_COSINE_LR]

LR_RANGE_TEST_MIN_LR = "lr_range_test_min_lr"
LR_RANGE_TEST_STEP_RATE = "lr_range_test_step_rate"
LR_RANGE_TEST_STEP_SIZE = "lr_range_test_step_size"
LR_RANGE_TEST_STAIRCASE = "lr_range_test_staircase"
EDGE_VALUE = "edge_value"
MID_VALUE = "mid_value"
CYCLE_FIRST_STEP_SIZE = "cycle_first_step_size"
CYCLE_FIRST_STAIR_COUNT = "cycle_first_stair_count"
CYCLE_SECOND_STEP_SIZE = "cycle_second_step_size"
CYCLE_SECOND_STAIR_COUNT = "cycle_second_stair_count"
DECAY_STEP_SIZE = "decay_step_size"
CYCLE_MIN_LR = "cycle_min_lr"
CYCLE_MAX_LR = "cycle_max_lr"
DECAY_LR_RATE = "decay_lr_rate"
CYCLE_MIN_MOM = "cycle_min_mom"
CYCLE_MAX_MOM = "cycle_max_mom"
DECAY_MOM_RATE = "decay_mom_rate"
WARMUP_MIN_LR = "warmup_min_lr"
WARMUP_MAX_LR = "warmup_max_lr"
WARMUP_NUM_STEPS = "warmup_num_steps"
WARMUP_TYPE = "warmup_type"
WARMUP_LOG_RATE = "log"
WARMUP_LINEAR_RATE = "linear"
WARMUP_MIN_RATIO = "warmup_min_ratio"
COS_MIN_RATIO = "cos_min_ratio"
TOTAL_NUM_STEPS = "total_num_steps"


def add_tuning_arguments(parser: argparse.ArgumentParser):
    g = parser.add_argument_group("Convergence Tuning", "Convergence tuning configurations")
    g.add_argument("--lr_schedule", type=str, default=None, help="LR schedule for training.")
    g.add_argument("--lr_range_test_min_lr", type=float, default=0.001)
    g.add_argument("--lr_range_test_step_rate", type=float, default=1.0)
    g.add_argument("--lr_range_test_step_size", type=int, default=1000)
    g.add_argument("--lr_range_test_staircase", type=bool, default=False)
    g.add_argument("--cycle_first_step_size", type=int, default=1000)
    g.add_argument("--cycle_first_stair_count", type=int, default=-1)
    g.add_argument("--cycle_second_step_size", type=int, default=-1)
    g.add_argument("--cycle_second_stair_count", type=int, default=-1)
    g.add_argument("--decay_step_size", type=int, default=1000)
    g.add_argument("--cycle_min_lr", type=float, default=0.01)
    g.add_argument("--cycle_max_lr", type=float, default=0.1)
    g.add_argument("--decay_lr_rate", type=float, default=0.0)
    g.add_argument("--cycle_momentum", default=False, action="store_true")
    g.add_argument("--cycle_min_mom", type=float, default=0.8)
    g.add_argument("--cycle_max_mom", type=float, default=0.9)
    g.add_argument("--decay_mom_rate", type=float, default=0.0)
    g.add_argument("--warmup_min_lr", type=float, default=0)
    g.add_argument("--warmup_max_lr", type=float, default=0.001)
    g.add_argument("--warmup_num_steps", type=int, default=1000)
    g.add_argument("--warmup_type", type=str, default=WARMUP_LOG_RATE)
    return parser


def parse_arguments():
    parser = add_tuning_arguments(argparse.ArgumentParser())
    return parser.parse_known_args()


def get_config_from_args(args):
    """Build ``{"type": ..., "params": {...}}`` from CLI flags, or ``(None, reason)``."""
    name = getattr(args, LR_SCHEDULE, None)
    if name is None:
        return None, "--lr_schedule not specified on command line"
    if name not in VALID_LR_SCHEDULES:
        return None, f"{name} is not supported LR schedule"
    keys = {
        LR_RANGE_TEST: [LR_RANGE_TEST_MIN_LR, LR_RANGE_TEST_STEP_RATE, LR_RANGE_TEST_STEP_SIZE, LR_RANGE_TEST_STAIRCASE],
        ONE_CYCLE: [CYCLE_FIRST_STEP_SIZE, CYCLE_FIRST_STAIR_COUNT, CYCLE_SECOND_STEP_SIZE, CYCLE_SECOND_STAIR_COUNT,
                    DECAY_STEP_SIZE, CYCLE_MIN_LR, CYCLE_MAX_LR, DECAY_LR_RATE, CYCLE_MIN_MOM, CYCLE_MAX_MOM,
                    DECAY_MOM_RATE],
        WARMUP_LR: [WARMUP_MIN_LR, WARMUP_MAX_LR, WARMUP_NUM_STEPS, WARMUP_TYPE],
        WARMUP_DECAY_LR: [WARMUP_MIN_LR, WARMUP_MAX_LR, WARMUP_NUM_STEPS, WARMUP_TYPE],
        WARMUP_COSINE_LR: [WARMUP_NUM_STEPS, WARMUP_TYPE],
    }[name]
    return {"type": name, "params": {k: getattr(args, k) for k in keys if hasattr(args, k)}}, None


def get_lr_from_config(config):
    if "type" not in config:
        return None, "LR schedule type not defined in config"
    if "params" not in config:
        return None, "LR schedule params not defined in config"
    t, p = config["type"], config["params"]
    if t not in VALID_LR_SCHEDULES:
        return None, f"{t} is not a valid LR schedule"
    if t == LR_RANGE_TEST:
        return p[LR_RANGE_TEST_MIN_LR], ""
    if t == ONE_CYCLE:
        return p[CYCLE_MAX_LR], ""
    return p[WARMUP_MAX_LR], ""


def get_torch_optimizer(optimizer):
    if isinstance(optimizer, Optimizer):
        return optimizer
    if hasattr(optimizer, "optimizer") and isinstance(optimizer.optimizer, Optimizer):
        return optimizer.optimizer
    raise TypeError(f"{type(optimizer).__name__} is not a subclass of torch.optim.Optimizer")


def _per_group(value, optimizer, name):
    n = len(optimizer.param_groups)
    if isinstance(value, (list, tuple)):
        if len(value) != n:
            raise ValueError(f"Expected {n} value for {name}, got {len(value)}")
        return list(value)
    return [value] * n


class _StepSchedule:
    """Step counter + optimizer write-back shared by all schedules."""

    def __init__(self, optimizer, last_batch_iteration=-1):
        self.optimizer = get_torch_optimizer(optimizer)
        self.last_batch_iteration = last_batch_iteration
        self._last_lr = None

    def get_lr(self) -> List[float]:  # pragma: no cover - abstract
        raise NotImplementedError

    def get_last_lr(self):
        assert self._last_lr is not None, "need to call step() first"
        return self._last_lr

    def _apply(self):
        lrs = self.get_lr()
        for g, lr in zip(self.optimizer.param_groups, lrs):
            g["lr"] = lr
        self._last_lr = [g["lr"] for g in self.optimizer.param_groups]

    def step(self, last_batch_iteration=None):
        if last_batch_iteration is None:
            last_batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = last_batch_iteration
        self._apply()

    def state_dict(self):
        return {"last_batch_iteration": self.last_batch_iteration}

    def load_state_dict(self, sd):
        self.last_batch_iteration = sd["last_batch_iteration"]


class LRRangeTest(_StepSchedule):
    """lr = min_lr * (1 + step_rate * interval) with interval continuous or staircase."""

    def __init__(self, optimizer, lr_range_test_min_lr: Union[float, list] = 1e-3, lr_range_test_step_size: int = 2000,
                 lr_range_test_step_rate: float = 1.0, lr_range_test_staircase: bool = False,
                 last_batch_iteration: int = -1):
        super().__init__(optimizer, last_batch_iteration)
        self.min_lr = _per_group(lr_range_test_min_lr, self.optimizer, "lr_range_test_min_lr")
        self.step_size = lr_range_test_step_size
        self.step_rate = lr_range_test_step_rate
        self.staircase = lr_range_test_staircase
        self.interval_fn = self._staircase_interval if lr_range_test_staircase else self._continuous_interval
        if last_batch_iteration == -1:
            for g, lr in zip(self.optimizer.param_groups, self.min_lr):
                g["lr"] = lr

    def _staircase_interval(self):
        return math.floor(float(self.last_batch_iteration + 1) / self.step_size)

    def _continuous_interval(self):
        return float(self.last_batch_iteration + 1) / self.step_size

    def _get_increase(self):
        return 1 + self.step_rate * self.interval_fn()

    def get_lr(self):
        inc = self._get_increase()
        return [m * inc for m in self.min_lr]


class OneCycle(_StepSchedule):
    """Triangular 1-cycle policy on lr (and inversely on momentum) followed by a decay phase."""

    def __init__(self, optimizer, cycle_min_lr, cycle_max_lr, decay_lr_rate=0.0, cycle_first_step_size=2000,
                 cycle_second_step_size=None, cycle_first_stair_count=0, cycle_second_stair_count=None,
                 decay_step_size=0, cycle_momentum=True, cycle_min_mom=0.8, cycle_max_mom=0.9, decay_mom_rate=0.0,
                 last_batch_iteration=-1):
        super().__init__(optimizer, last_batch_iteration)
        first = float(cycle_first_step_size)
        second = float(cycle_second_step_size) if cycle_second_step_size is not None else first
        self.total_size = first + second
        self.step_ratio = first / self.total_size
        self.first_stair_count = cycle_first_stair_count
        self.second_stair_count = cycle_first_stair_count if cycle_second_stair_count is None else cycle_second_stair_count
        self.decay_step_size = decay_step_size
        if math.isclose(self.decay_step_size, 0):
            self.skip_lr_decay = self.skip_mom_decay = True
        else:
            self.skip_lr_decay = math.isclose(decay_lr_rate, 0)
            self.skip_mom_decay = math.isclose(decay_mom_rate, 0)
        n = len(self.optimizer.param_groups)
        self.min_lrs, self.max_lrs = [cycle_min_lr] * n, [cycle_max_lr] * n
        self.decay_lr_rate = decay_lr_rate
        if last_batch_iteration == -1:
            for lr, g in zip(self.min_lrs, self.optimizer.param_groups):
                g["lr"] = lr
        self.cycle_momentum = cycle_momentum
        if cycle_momentum:
            if "betas" not in self.optimizer.defaults:
                logger.warning(f"cycle_momentum is disabled because optimizer {type(self.optimizer).__name__} does "
                               "not support momentum, no betas attribute in defaults")
                self.cycle_momentum = False
            else:
                self.decay_mom_rate = decay_mom_rate
                self.min_moms = [(cycle_min_mom, 0.99)] * n
                self.max_moms = [(cycle_max_mom, 0.99)] * n
                if last_batch_iteration == -1:
                    for m, g in zip(self.min_moms, self.optimizer.param_groups):
                        g["betas"] = m

    def _scale_factor(self):
        it = self.last_batch_iteration + 1
        cycle = math.floor(1 + it / self.total_size)
        x = 1.0 + it / self.total_size - cycle
        return x / self.step_ratio if x <= self.step_ratio else (x - 1) / (self.step_ratio - 1)

    def _cycle_lr(self):
        s = self._scale_factor()
        return [lo + (hi - lo) * s for lo, hi in zip(self.min_lrs, self.max_lrs)]

    def _cycle_mom(self):
        s = self._scale_factor()
        return [(hi[0] - (hi[0] - lo[0]) * s, lo[1]) for lo, hi in zip(self.min_moms, self.max_moms)]

    def _decay_interval(self):
        return (self.last_batch_iteration - self.total_size + 1) / self.decay_step_size

    def get_lr(self):
        if self.last_batch_iteration < self.total_size:
            return self._cycle_lr()
        if self.skip_lr_decay:
            return list(self.min_lrs)
        f = 1 + self.decay_lr_rate * self._decay_interval()
        return [lr / f for lr in self.min_lrs]

    def get_mom(self):
        if not self.cycle_momentum:
            return None
        if self.last_batch_iteration < self.total_size:
            return self._cycle_mom()
        if self.skip_mom_decay:
            return list(self.max_moms)
        f = 1 + self.decay_mom_rate * self._decay_interval()
        return [(b0 * f, b1) for b0, b1 in self.max_moms]

    def step(self, batch_iteration=None):
        if batch_iteration is None:
            batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = batch_iteration
        self._apply()
        if self.cycle_momentum:
            for g, m in zip(self.optimizer.param_groups, self.get_mom()):
                g["betas"] = m


class WarmupLR(_StepSchedule):
    """min_lr -> max_lr over ``warmup_num_steps`` (log or linear ramp), then constant."""

    def __init__(self, optimizer, warmup_min_lr: Union[float, list] = 0.0, warmup_max_lr: Union[float, list] = 0.001,
                 warmup_num_steps: int = 1000, warmup_type: str = WARMUP_LOG_RATE, last_batch_iteration: int = -1):
        super().__init__(optimizer, last_batch_iteration)
        self.min_lrs = _per_group(warmup_min_lr, self.optimizer, "min_lr")
        self.max_lrs = _per_group(warmup_max_lr, self.optimizer, "max_lr")
        self.delta_lrs = [hi - lo for hi, lo in zip(self.max_lrs, self.min_lrs)]
        self.warmup_num_steps = max(2, warmup_num_steps)
        if warmup_type not in (WARMUP_LOG_RATE, WARMUP_LINEAR_RATE):
            logger.warning(f"Using unknown warmup_type: {warmup_type}. The increasing function is set to default (log)")
            warmup_type = WARMUP_LOG_RATE
        self.warmup_type = warmup_type
        self.inverse_log_warm_up = 1.0 / math.log(self.warmup_num_steps)
        if last_batch_iteration == -1:
            self._last_lr = self._update_optimizer_initial()

    def _update_optimizer_initial(self):
        self._apply()
        return self._last_lr

    def _get_gamma(self):
        if self.last_batch_iteration < self.warmup_num_steps:
            if self.warmup_type == WARMUP_LOG_RATE:
                return self.inverse_log_warm_up * math.log(self.last_batch_iteration + 1)
            return self.last_batch_iteration / self.warmup_num_steps
        return 1.0

    def get_lr(self):
        if self.last_batch_iteration < 0:
            logger.warning("Attempting to get learning rate from scheduler before it has started")
            return list(self.min_lrs)
        g = self._get_gamma()
        return [lo + d * g for lo, d in zip(self.min_lrs, self.delta_lrs)]


class WarmupDecayLR(WarmupLR):
    """Warm-up as :class:`WarmupLR`, then linear decay to 0 at ``total_num_steps``."""

    def __init__(self, optimizer, total_num_steps: int, warmup_min_lr=0.0, warmup_max_lr=0.001, warmup_num_steps=1000,
                 warmup_type=WARMUP_LOG_RATE, last_batch_iteration=-1):
        self.total_num_steps = total_num_steps
        super().__init__(optimizer, warmup_min_lr, warmup_max_lr, warmup_num_steps, warmup_type, last_batch_iteration)
        if self.total_num_steps < self.warmup_num_steps:
            logger.warning(f"total_num_steps {total_num_steps} is less than warmup_num_steps {warmup_num_steps}")

    def _get_gamma(self):
        if self.last_batch_iteration < self.warmup_num_steps:
            return super()._get_gamma()
        return max(0.0, float(self.total_num_steps - self.last_batch_iteration) /
                   float(max(1.0, self.total_num_steps - self.warmup_num_steps)))


class WarmupCosineLR(_StepSchedule):
    """Ratio schedule on the optimizer's initial lr: warm-up from ``warmup_min_ratio`` to 1, then cosine
    down to ``cos_min_ratio``."""

    def __init__(self, optimizer, total_num_steps: int, warmup_min_ratio: float = 0.0, warmup_num_steps: int = 1000,
                 cos_min_ratio: float = 0.0001, warmup_type: str = WARMUP_LOG_RATE, last_batch_iteration: int = -1):
        super().__init__(optimizer, last_batch_iteration)
        self.total_num_steps = total_num_steps
        self.cos_min_ratio = cos_min_ratio
        self.warmup_type = warmup_type if warmup_type in (WARMUP_LOG_RATE, WARMUP_LINEAR_RATE) else WARMUP_LOG_RATE
        self.warmup_min_ratio = warmup_min_ratio
        self.warmup_num_steps = max(2, warmup_num_steps)
        self.inverse_log_warm_up = 1.0 / math.log(self.warmup_num_steps)
        if self.total_num_steps < self.warmup_num_steps:
            logger.warning(f"total_num_steps {total_num_steps} is less than warmup_num_steps {warmup_num_steps}")
        self.org_lrs = [g["lr"] for g in self.optimizer.param_groups]
        if last_batch_iteration == -1:
            self._apply()

    def get_lr_ratio(self):
        it = self.last_batch_iteration
        if it < 0:
            logger.warning("Attempting to get learning rate from scheduler before it has started")
            return 0.0
        if it < self.warmup_num_steps:
            if self.warmup_type == WARMUP_LOG_RATE:
                ratio = self.inverse_log_warm_up * math.log(it + 1)
            else:
                ratio = it / self.warmup_num_steps
            return self.warmup_min_ratio + (1.0 - self.warmup_min_ratio) * ratio
        real_last = it - self.warmup_num_steps + 1
        real_total = self.total_num_steps - self.warmup_num_steps
        cos = 0.5 * (1.0 + math.cos(math.pi * min(1.0, real_last / max(1, real_total))))
        return max(0.0, self.cos_min_ratio + (1.0 - self.cos_min_ratio) * cos)

    def get_lr(self):
        if self.last_batch_iteration < 0:
            return [0.0] * len(self.org_lrs)
        r = self.get_lr_ratio()
        return [lr * r for lr in self.org_lrs]


# ---- command-line overrides of a config's scheduler block (reference ``lr_schedules.py:130-206``)
def override_lr_range_test_params(args, params):
    for key in (LR_RANGE_TEST_MIN_LR, LR_RANGE_TEST_STEP_RATE, LR_RANGE_TEST_STEP_SIZE, LR_RANGE_TEST_STAIRCASE):
        if getattr(args, key, None) is not None:
            params[key] = getattr(args, key)


def override_1cycle_params(args, params):
    for key in (CYCLE_FIRST_STEP_SIZE, CYCLE_FIRST_STAIR_COUNT, CYCLE_SECOND_STEP_SIZE, CYCLE_SECOND_STAIR_COUNT,
                DECAY_STEP_SIZE, CYCLE_MIN_LR, CYCLE_MAX_LR, DECAY_LR_RATE, CYCLE_MIN_MOM, CYCLE_MAX_MOM, DECAY_MOM_RATE):
        if getattr(args, key, None) is not None:
            params[key] = getattr(args, key)


def override_warmupLR_params(args, params):
    for key in (WARMUP_MIN_LR, WARMUP_MAX_LR, WARMUP_NUM_STEPS, WARMUP_TYPE):
        if getattr(args, key, None) is not None:
            params[key] = getattr(args, key)


def override_params(args, params):
    """Flags given on the command line win over the values in the config's ``scheduler.params``."""
    override_lr_range_test_params(args, params)
    override_1cycle_params(args, params)
    override_warmupLR_params(args, params)


def update_lr(param_groups, lrs):
    for group, lr in zip(param_groups, lrs):
        group["lr"] = lr
    return [group["lr"] for group in param_groups]
