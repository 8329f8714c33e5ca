"""Keep a fused inference layer consistent with a client layer that continues to train — the RLHF generate/train
alternation (reference: ``features/hybrid_engine.py``)."""
from __future__ import annotations

from typing import List, Tuple

import torch

from ...policy import maybe_get_lora


class HybridEngineContainer:
    """Mix-in for :class:`BaseTransformerContainer`.

    * ``refresh()`` re-reads the client's (updated) weights into the existing fused module — cheap, no reallocation;
    * ``set_lora_params()`` collects the LoRA triplets of the client's projections; ``fuse_lora()`` /
      ``unfuse_lora()`` fold ``scaling * (right @ left)^T`` into / out of the CLIENT weights, so a following
      ``refresh()`` hands the fused layer the merged weights for generation;
    * ``release_memory()`` drops the fused module's tensors between generation phases."""

    lora_params: List[Tuple[torch.nn.Parameter, list]] = ()

    def initialize_tensors(self, enable_training: bool = False):
        super().initialize_tensors()
        self.set_lora_params()

    def _lora_owners(self):
        for m in self.child.modules():
            w = getattr(m, "weight", None)
            lora = maybe_get_lora(m)
            if w is not None and lora:
                yield w, lora

    def set_lora_params(self):
        self.lora_params = list(self._lora_owners())

    def get_lora_params(self):
        return [p for _, l in self.lora_params for p in l[:2]]

    @torch.no_grad()
    def fuse_lora(self):
        for w, (right, left, scaling) in self.lora_params:
            w.data += float(scaling) * torch.matmul(left.t(), right.t()).to(w.dtype)

    @torch.no_grad()
    def unfuse_lora(self):
        for w, (right, left, scaling) in self.lora_params:
            w.data -= float(scaling) * torch.matmul(left.t(), right.t()).to(w.dtype)

    @torch.no_grad()
    def refresh(self):
        """Client weights -> fused module (after optimizer steps or LoRA folding)."""
        super().initialize_tensors()
        self.apply_tensor_parallelism()
        self.copy_data_to_new_module()

    def release_memory(self):
        self.attn = self.mlp_t = self.norms = None

    # reference names
    def transform_for_training(self):
        self.release_memory()

    def transform_for_inference(self):
        self.refresh()
