"""Fetch / prefetch / release policy of ZeRO-3, exposed with the reference's coordinator vocabulary.

Reference: ``runtime/zero/partitioned_param_coordinator.py:66-585`` — records the order in which sub-modules run on
the first iteration, then prefetches ahead of the compute stream up to ``prefetch_bucket_size`` elements (bounded by
``max_live_parameters``), releases a module's parameters once they are not reused within ``max_reuse_distance``,
and invalidates the trace when the execution order changes.

In this framework the unit of gathering is a *unit* (one transformer block = one flat buffer, see
``runtime/zero/stage3.py``), so a fetch is ONE all-gather (or one pull kernel over NVLink) instead of a coalesced
gather of many parameter shards; the bookkeeping lives in :class:`DeepSpeedZeroOptimizer_Stage3` and this class is
the policy view over it (used by the engine, by tests and by tools that want to inspect or steer prefetching).
"""
from __future__ import annotations

import collections
from typing import Deque, Iterable, List, Set

from ...utils.logging import logger


class ZeRoTraceMode:
    RECORD, COMPLETE, INVALID = 1, 2, 3


class InflightParamRegistry(dict):
    """unit index -> fetch event of all-gathers that were issued and not yet waited on."""

    def __setitem__(self, unit_index, event):
        if unit_index in self:
            raise RuntimeError(f"unit {unit_index} already has an all-gather in flight")
        super().__setitem__(unit_index, event)


class PartitionedParameterCoordinator:
    def __init__(self, optimizer):
        self.opt = optimizer

    # ------------------------------------------------------------------ limits (live, from config)
    @property
    def prefetch_bucket_sz(self) -> int:
        return self.opt.prefetch_bucket_size

    @property
    def max_available_parameters(self) -> int:
        return self.opt.max_live_parameters

    @property
    def max_reuse_distance(self) -> int:
        return self.opt.max_reuse_distance

    # ------------------------------------------------------------------------------------ trace
    def trace_mode(self) -> int:
        return ZeRoTraceMode.COMPLETE if self.opt._trace_complete else ZeRoTraceMode.RECORD

    def is_complete_trace(self) -> bool:
        return self.opt._trace_complete

    def is_record_trace(self) -> bool:
        return not self.opt._trace_complete

    def is_invalid_trace(self) -> bool:
        return False

    def trace(self) -> List[str]:
        return [self.opt.units[i].name for i in self.opt._trace]

    def reset_step(self) -> None:
        """Called at the start of every forward: on the first pass the order is still being recorded."""
        self.opt._trace_pos = 0

    def _invalidate_trace(self) -> None:
        logger.info("ZeRO-3 unit trace invalidated; it will be re-recorded on the next iteration")
        self.opt._trace, self.opt._trace_complete, self.opt._trace_pos = [], False, 0

    def trace_prologue(self, sub_module) -> None:
        """Before a sub-module runs: in COMPLETE mode check that execution still follows the recorded order and fall
        back to re-recording when it does not (reference ``coordinator.py:179-213``)."""
        u = self._unit_of_module(sub_module)
        if u is None or not self.opt._trace_complete:
            return
        pos = getattr(self.opt, "_trace_pos", 0)
        if pos < len(self.opt._trace) and self.opt._trace[pos] != u.index and not self.opt._in_backward:
            logger.info(f"ZeRO-3 trace mismatch at step {pos}: expected unit {self.opt._trace[pos]}, running {u.index}")
            self._invalidate_trace()

    def record_module(self, sub_module) -> None:
        """RECORD mode: append the sub-module's unit to the trace."""
        u = self._unit_of_module(sub_module)
        if u is None:
            return
        if self.opt._trace_complete:
            raise RuntimeError("attempted to record trace when status = COMPLETE")
        self.opt._trace.append(u.index)

    def record_parameters(self, sub_module) -> None:
        """Parameters travel with their unit here, so recording the module records its parameters."""
        if self.opt._trace_complete:
            raise RuntimeError("attempted to record trace when status = COMPLETE")

    def construct_parameter_trace_from_module_trace(self):
        self.opt._trace_complete = len(self.opt._trace) > 0

    # ---------------------------------------------------------------------------- fetch/release
    def inflight(self) -> InflightParamRegistry:
        reg = InflightParamRegistry()
        for u in self.opt.units:
            if u.fetch_event is not None:
                reg[u.index] = u.fetch_event
        return reg

    def n_available_params(self) -> int:
        return self.opt._live_numel

    def fetch_sub_module(self, module, forward: bool = True) -> None:
        u = self._unit_of_module(module)
        if u is None:
            return
        self.opt._fetch(u)
        self.opt._prefetch_after(u, forward)

    def release_sub_module(self, module, forward: bool = True) -> None:
        u = self._unit_of_module(module)
        if u is not None and self._reuse_distance(u, forward) > self.max_reuse_distance:
            self.opt._release(u)

    def release_and_reset_all(self, module=None) -> None:
        self.opt.release_all()
        self._invalidate_trace()

    def prefetch(self, after_module, forward: bool = True) -> None:
        u = self._unit_of_module(after_module)
        if u is not None:
            self.opt._prefetch_after(u, forward)

    # --------------------------------------------------------------------------------- helpers
    def _unit_of_module(self, module):
        for u in self.opt.units:
            if u.module is module:
                return u
        return None

    def _reuse_distance(self, u, forward: bool) -> int:
        """Elements gathered between now and the next use of ``u`` in the recorded order (inf if never)."""
        if not self.opt._trace_complete:
            return self.max_reuse_distance + 1
        order = self.opt._trace if forward else list(reversed(self.opt._trace))
        try:
            pos = order.index(u.index)
        except ValueError:
            return self.max_reuse_distance + 1
        dist_ = 0
        for nxt in order[pos + 1:]:
            if nxt == u.index:
                return dist_
            dist_ += self.opt.units[nxt].numel
        return self.max_reuse_distance + 1
