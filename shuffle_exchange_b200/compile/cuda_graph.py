"""Static-shape CUDA-graph capture of a callable — the "compiled function" of this framework.

Launch-bound inner loops (decode steps, small-batch fine-tuning micro-steps, optimizer sweeps) are captured once and
replayed; inputs are copied into static buffers, outputs are returned as views of static buffers.  The reference gets
its launch-overhead reduction from inductor's CUDA-graph trees (``compile/inductor.py``)."""
from __future__ import annotations

from typing import Callable, Dict, Tuple

import torch


def _key(args) -> Tuple:
    return tuple((tuple(a.shape), a.dtype, a.device) if torch.is_tensor(a) else ("py", a) for a in args)


class GraphedCallable:
    def __init__(self, fn: Callable, warmup: int = 3, pool=None):
        self.fn, self.warmup, self.pool = fn, warmup, pool
        self._graphs: Dict[Tuple, Tuple] = {}

    def __call__(self, *args):
        if not args or not any(torch.is_tensor(a) and a.is_cuda for a in args):
            return self.fn(*args)
        k = _key(args)
        ent = self._graphs.get(k)
        if ent is None:
            ent = self._capture(args)
            self._graphs[k] = ent
        graph, static_in, static_out = ent
        for s, a in zip(static_in, args):
            if torch.is_tensor(a):
                s.copy_(a, non_blocking=True)
        graph.replay()
        return static_out

    def _capture(self, args):
        static_in = [a.clone() if torch.is_tensor(a) else a for a in args]
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(self.warmup):
                self.fn(*static_in)
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, pool=self.pool):
            static_out = self.fn(*static_in)
        return graph, static_in, static_out

    @property
    def num_graphs(self) -> int:
        return len(self._graphs)
