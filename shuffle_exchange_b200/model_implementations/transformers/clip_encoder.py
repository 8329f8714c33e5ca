"""CUDA-graphed CLIP text encoder wrapper (reference ``model_implementations/transformers/clip_encoder.py``): the
prompt and the negative prompt alternate, so two graphs are kept and used round-robin."""
from __future__ import annotations

import torch

from ..features.graphed_module import GraphedModule


class DSClipEncoder(GraphedModule):

    def __init__(self, enc, enable_cuda_graph=False):
        super().__init__(enable_cuda_graph=enable_cuda_graph)
        if hasattr(enc, "text_model") and hasattr(enc.text_model, "_build_causal_attention_mask"):
            enc.text_model._build_causal_attention_mask = self._build_causal_attention_mask
        self.enc = enc
        self.device = getattr(enc, "device", None)
        self.dtype = getattr(enc, "dtype", None)
        self.cuda_graph_created = [False, False]
        self.iter = 0
        self.config = getattr(enc, "config", None)

    def _build_causal_attention_mask(self, bsz, seq_len, dtype):
        dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else "cpu"
        mask = torch.empty(bsz, seq_len, seq_len, dtype=dtype, device=dev)
        mask.fill_(torch.finfo(dtype).min)
        mask.triu_(1)
        return mask.unsqueeze(1)

    def forward(self, *inputs, **kwargs):
        name = f"enc{self.iter}"
        out = self._run(name, self._forward, *inputs, **kwargs)
        self.cuda_graph_created[self.iter] = name in self._graphs
        self.iter = (self.iter + 1) % 2
        return out

    def _forward(self, *inputs, **kwargs):
        return self.enc(*inputs, **kwargs)
