"""Per-unit compute-time and activation-size profile (reference ``compile/profilers/graph_profile.py`` profiles FX
nodes by interpreting the graph; here forward/backward hooks on the ZeRO units bracket each unit with CUDA events)."""
from __future__ import annotations

import time
from typing import Dict, List

import torch


class UnitProfiler:
    def __init__(self, engine):
        self.engine = engine
        self._handles: List = []
        self._ev: Dict[int, Dict[str, List]] = {}
        self._mem: Dict[int, int] = {}
        self.cuda = torch.cuda.is_available()

    def _now(self):
        if self.cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            return e
        return time.perf_counter()

    def __enter__(self):
        opt = self.engine.optimizer
        for u in getattr(opt, "units", []):
            if u.module is None:
                continue
            idx = u.index
            self._ev[idx] = {"fs": [], "fe": [], "bs": [], "be": []}

            def pre(m, a, idx=idx):
                self._ev[idx]["fs"].append(self._now())
                if self.cuda:
                    self._mem[idx] = -torch.cuda.memory_allocated()

            def post(m, a, o, idx=idx):
                self._ev[idx]["fe"].append(self._now())
                if self.cuda:
                    self._mem[idx] += torch.cuda.memory_allocated()

            def bpre(m, g, idx=idx):
                self._ev[idx]["bs"].append(self._now())

            def bpost(m, gi, go, idx=idx):
                self._ev[idx]["be"].append(self._now())

            self._handles += [u.module.register_forward_pre_hook(pre), u.module.register_forward_hook(post),
                              u.module.register_full_backward_pre_hook(bpre),
                              u.module.register_full_backward_hook(bpost)]
        return self

    def __exit__(self, *exc):
        for h in self._handles:
            h.remove()
        self._handles = []
        return False

    def _span(self, a, b) -> float:
        if self.cuda:
            return a.elapsed_time(b) * 1e3
        return (b - a) * 1e6

    def result(self) -> Dict[str, Dict[int, float]]:
        if self.cuda:
            torch.cuda.synchronize()
        fwd, bwd = {}, {}
        for idx, ev in self._ev.items():
            if ev["fs"] and ev["fe"]:
                fwd[idx] = sum(self._span(a, b) for a, b in zip(ev["fs"], ev["fe"]))
            if ev["bs"] and ev["be"]:
                bwd[idx] = sum(self._span(a, b) for a, b in zip(ev["bs"], ev["be"]))
        return {"fwd_us": fwd, "bwd_us": bwd, "activation_bytes": {k: max(0, v) for k, v in self._mem.items()}}
