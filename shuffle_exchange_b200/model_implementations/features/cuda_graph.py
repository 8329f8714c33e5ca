"""Mixin for modules that replay their forward from a captured CUDA graph (reference
``model_implementations/features/cuda_graph.py``): subclasses implement ``_create_cuda_graph`` and
``_graph_replay``."""
from abc import ABC, abstractmethod


class CUDAGraph(ABC):

    def __init__(self, enable_cuda_graph=False):
        super().__init__()
        self.enable_cuda_graph = enable_cuda_graph

    @abstractmethod
    def _create_cuda_graph(self):
        raise NotImplementedError

    @abstractmethod
    def _graph_replay(self):
        raise NotImplementedError
