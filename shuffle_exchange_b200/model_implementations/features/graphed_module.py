"""Shared capture/replay logic for the graphed wrappers (UNet, VAE, CLIP encoder): one graph per named entry point,
static input/output buffers, 3 warm-up iterations on a side stream before capture."""
from __future__ import annotations

from typing import Any, Callable, Dict

import torch

from .cuda_graph import CUDAGraph


class GraphedModule(CUDAGraph, torch.nn.Module):

    def __init__(self, enable_cuda_graph: bool = True):
        torch.nn.Module.__init__(self)
        self.enable_cuda_graph = bool(enable_cuda_graph and torch.cuda.is_available())
        self._graphs: Dict[str, Any] = {}
        self._static_in: Dict[str, Any] = {}
        self._static_out: Dict[str, Any] = {}

    def _copy_in(self, name, args, kwargs):
        sa, sk = self._static_in[name]
        for s, a in zip(sa, args):
            if torch.is_tensor(a):
                s.copy_(a)
        for k, a in kwargs.items():
            if torch.is_tensor(a):
                sk[k].copy_(a)

    def _graph_replay(self, name: str, *args, **kwargs):
        self._copy_in(name, args, kwargs)
        self._graphs[name].replay()
        return self._static_out[name]

    def _create_cuda_graph(self, name: str, fn: Callable, *args, **kwargs):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                fn(*args, **kwargs)
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        self._static_in[name] = ([a.clone() if torch.is_tensor(a) else a for a in args],
                                 {k: (a.clone() if torch.is_tensor(a) else a) for k, a in kwargs.items()})
        sa, sk = self._static_in[name]
        with torch.cuda.graph(g):
            self._static_out[name] = fn(*sa, **sk)
        self._graphs[name] = g

    def _run(self, name: str, fn: Callable, *args, **kwargs):
        if not self.enable_cuda_graph:
            return fn(*args, **kwargs)
        if name not in self._graphs:
            self._create_cuda_graph(name, fn, *args, **kwargs)
        return self._graph_replay(name, *args, **kwargs)
