"""How often, and how many elements, each parameter-movement event of ZeRO-3 touched, with wall-clock timers.

Reference: ``runtime/zero/partitioned_param_profiler.py:13-63`` (``start_event`` / ``stop_event`` bracket a fetch,
release, or swap; ``log_events`` prints per-event counts and the timers).  Here one record per event holds its tally and
the profiler degrades to a no-op when the engine runs without wall-clock timers."""
from __future__ import annotations

from typing import Dict


class _Tally:
    """Occurrences and elements moved for one event name."""
    __slots__ = ("name", "count", "num_elem")

    def __init__(self, name: str):
        self.name, self.count, self.num_elem = name, 0, 0

    def increment(self, numel: int) -> None:
        self.count, self.num_elem = self.count + 1, self.num_elem + int(numel)

    def reset(self) -> None:
        self.count = self.num_elem = 0

    def __repr__(self):
        return f"{self.name}: count = {self.count}, numel = {self.num_elem}"


class PartitionedParameterProfiler:
    EventCounter = _Tally                      # reference name of the record type

    def __init__(self, timers):
        self.timers = timers
        self.event_counters: Dict[str, _Tally] = {}

    @property
    def enabled(self) -> bool:
        return self.timers is not None

    def start_event(self, name: str) -> None:
        if self.enabled:
            self.event_counters.setdefault(name, _Tally(name))
            self.timers(name).start()

    def stop_event(self, name: str, num_elem: int) -> None:
        if self.enabled:
            tally = self.event_counters.get(name)
            if tally is None:
                raise KeyError(f"stop_event({name!r}) without a matching start_event")
            tally.increment(num_elem)
            self.timers(name).stop()

    def reset_events(self) -> None:
        for tally in self.event_counters.values():
            tally.reset()

    def log_events(self) -> None:
        from ...utils.logging import log_dist
        for tally in self.event_counters.values():
            log_dist(repr(tally), ranks=[0])
        if self.enabled and self.event_counters:
            self.timers.log(names=list(self.event_counters))
