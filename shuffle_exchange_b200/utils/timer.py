"""Timers (reference: deepspeed/utils/timer.py — ``SynchronizedWallClockTimer`` :44-161,
``NoopTimer`` :164, ``ThroughputTimer`` :199-294; timer name constants :11-20).

CUDA-event pairs are recorded on the current stream and only resolved when a value is read, so
enabling ``wall_clock_breakdown`` does not serialise the step.
"""
from __future__ import annotations

import time

from ..device import get_accelerator
from .logging import log_dist

FORWARD_MICRO_TIMER = "fwd_microstep"
FORWARD_GLOBAL_TIMER = "fwd"
BACKWARD_MICRO_TIMER = "bwd_microstep"
BACKWARD_GLOBAL_TIMER = "bwd"
BACKWARD_INNER_MICRO_TIMER = "bwd_inner_microstep"
BACKWARD_INNER_GLOBAL_TIMER = "bwd_inner"
BACKWARD_REDUCE_MICRO_TIMER = "bwd_allreduce_microstep"
BACKWARD_REDUCE_GLOBAL_TIMER = "bwd_allreduce"
STEP_MICRO_TIMER = "step_microstep"
STEP_GLOBAL_TIMER = "step"
TIME_EPSILON = 1e-6

try:
    import psutil
    _PSUTIL = True
except ImportError:  # pragma: no cover
    _PSUTIL = False


class CudaEventTimer:
    """One (start, end) event pair; the elapsed time is resolved when asked for (reference ``timer.py:32-41``)."""

    def __init__(self, start_event, end_event):
        self.start_event, self.end_event = start_event, end_event

    def get_elapsed_msec(self) -> float:
        self.end_event.synchronize()
        return self.start_event.elapsed_time(self.end_event)


class _EventTimer:
    def __init__(self, name):
        self.name_ = name
        self.started_ = False
        self.pairs = []       # unresolved (start, end) CUDA events
        self.elapsed_ms = 0.0
        self.records = []
        self._t0 = None
        self._use_events = get_accelerator().is_available()

    def start(self):
        assert not self.started_, f"{self.name_} timer has already been started"
        if self._use_events:
            self._ev0 = get_accelerator().Event(enable_timing=True)
            self._ev0.record()
        else:
            self._t0 = time.perf_counter()
        self.started_ = True

    def stop(self, reset=False, record=False):
        assert self.started_, "timer is not started"
        if self._use_events:
            ev1 = get_accelerator().Event(enable_timing=True)
            ev1.record()
            self.pairs.append(CudaEventTimer(self._ev0, ev1))
        else:
            self.elapsed_ms += (time.perf_counter() - self._t0) * 1e3
        self.started_ = False
        if record:
            self.records.append(self._resolve())

    def _resolve(self) -> float:
        if self.pairs:
            self.elapsed_ms += sum(pair.get_elapsed_msec() for pair in reversed(self.pairs))
            self.pairs = []
        return self.elapsed_ms

    def reset(self):
        self.started_ = False
        self.pairs = []
        self.elapsed_ms = 0.0

    def elapsed(self, reset=True) -> float:
        was = self.started_
        if was:
            self.stop()
        ms = self._resolve()
        if reset:
            self.reset()
        if was:
            self.start()
        return ms

    def mean(self) -> float:
        n = max(1, len(self.records))
        return sum(self.records) / n


class SynchronizedWallClockTimer:
    Timer = _EventTimer

    def __init__(self):
        self.timers = {}

    def get_timers(self):
        return self.timers

    def __call__(self, name):
        if name not in self.timers:
            self.timers[name] = _EventTimer(name)
        return self.timers[name]

    @staticmethod
    def memory_usage() -> str:
        acc = get_accelerator()
        gb = 1024 ** 3
        return (f" | mem_allocated: {acc.memory_allocated() / gb:.4f} GB | max_mem_allocated: "
                f"{acc.max_memory_allocated() / gb:.4f} GB | cache_allocated: {acc.memory_reserved() / gb:.4f} GB"
                f" | max_cache_allocated: {acc.max_memory_reserved() / gb:.4f} GB")

    def log(self, names, normalizer=1.0, reset=True, memory_breakdown=False, ranks=None):
        assert normalizer > 0.0
        s = "time (ms)"
        for n in names:
            if n in self.timers:
                s += f" | {n}: {self.timers[n].elapsed(reset=reset) / normalizer:.2f}"
        if memory_breakdown:
            s += self.memory_usage()
        log_dist(s, ranks=ranks or [0])

    def get_mean(self, names, normalizer=1.0, reset=True):
        out = {}
        for n in names:
            if n in self.timers:
                out[n] = self.timers[n].mean() * 1000.0 / normalizer
                if reset:
                    self.timers[n].records = []
        return out


class NoopTimer:
    class Timer:
        def start(self): ...
        def reset(self): ...
        def stop(self, **kw): ...
        def elapsed(self, **kw): return 0.0
        def mean(self): return 0.0

    def __init__(self):
        self.timer = self.Timer()

    def __call__(self, name):
        return self.timer

    def get_timers(self):
        return {}

    def log(self, *a, **k): ...
    def get_mean(self, *a, **k): return {}


class ThroughputTimer:
    """samples/sec after ``start_step`` warm-up steps (reference: timer.py:199-294)."""

    def __init__(self, config=None, batch_size=1, start_step=2, steps_per_output=None, monitor_memory=False,
                 logging_fn=None):
        self.config = config
        self.enabled = True if config is None else getattr(config, "enabled", True)
        self.synchronized = True if config is None else getattr(config, "synchronized", True)
        self.start_time = 0.0
        self.end_time = 0.0
        self.started = False
        self.batch_size = max(1, batch_size or 1)
        self.start_step = start_step
        self.epoch_count = 0
        self.micro_step_count = 0
        self.global_step_count = 0
        self.total_elapsed_time = 0.0
        self.step_elapsed_time = 0.0
        self.steps_per_output = steps_per_output
        self.monitor_memory = monitor_memory
        self.logging = logging_fn or (lambda m: log_dist(m, ranks=[0]))
        self.initialized = False

    def update_epoch_count(self):
        self.epoch_count += 1
        self.micro_step_count = 0

    def _init_timer(self):
        self.initialized = True

    def start(self):
        if not self.enabled:
            return
        self._init_timer()
        self.started = True
        if self.global_step_count >= self.start_step:
            if self.synchronized:
                get_accelerator().synchronize()
            self.start_time = time.time()

    def stop(self, global_step=False, report_speed=True):
        if not self.enabled or not self.started:
            return
        self.started = False
        self.micro_step_count += 1
        if global_step:
            self.global_step_count += 1
        if self.start_time > 0:
            if self.synchronized:
                get_accelerator().synchronize()
            self.end_time = time.time()
            d = self.end_time - self.start_time
            self.total_elapsed_time += d
            self.step_elapsed_time += d
            if global_step:
                if report_speed and self.steps_per_output and self.global_step_count % self.steps_per_output == 0:
                    self.logging(
                        f"epoch={self.epoch_count}/micro_step={self.micro_step_count}/global_step="
                        f"{self.global_step_count}, RunningAvgSamplesPerSec={self.avg_samples_per_sec():.4f}, "
                        f"CurrSamplesPerSec={self.batch_size / max(self.step_elapsed_time, TIME_EPSILON):.4f}, "
                        f"MemAllocated={get_accelerator().memory_allocated() / 1024 ** 3:.2f}GB, "
                        f"MaxMemAllocated={get_accelerator().max_memory_allocated() / 1024 ** 3:.2f}GB")
                    if self.monitor_memory and _PSUTIL:
                        vm = psutil.virtual_memory()
                        self.logging(f"virtual memory used={vm.used / 1024 ** 3:.2f}GB ({vm.percent}%)")
                self.step_elapsed_time = 0.0

    def avg_samples_per_sec(self) -> float:
        if self.global_step_count > self.start_step and self.total_elapsed_time > 0:
            steps = self.global_step_count - self.start_step
            return self.batch_size / (self.total_elapsed_time / steps)
        return float("-inf")


def trim_mean(data, trim_percent):
    assert 0.0 <= trim_percent <= 1.0
    n = len(data)
    if n == 0:
        return 0
    data = sorted(data)
    k = int(round(n * trim_percent))
    return sum(data[k:n - k]) / max(1, n - 2 * k)
