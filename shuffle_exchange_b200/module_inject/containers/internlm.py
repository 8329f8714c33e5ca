"""InternLM (reference: ``containers/internlm.py``): LLaMA layout with biases on the attention projections."""
from __future__ import annotations

from ...utils.types import ActivationFuncType, NormType
from .base import AttentionTensors, BaseTransformerContainer, MLPTensors, NormTensors, TransformerPolicy

from .llama import LLAMALayerPolicy


class InternLMLayerPolicy(LLAMALayerPolicy):
    _orig_layer_class_name = "InternLMDecoderLayer"


class DS_InternLMContainer(BaseTransformerContainer):
    pass
