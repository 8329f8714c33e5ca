"""Transformer-layer ops: the fused BERT-style training layer (reference ``ops/transformer/__init__.py``)."""
from .transformer import DeepSpeedTransformerConfig, DeepSpeedTransformerLayer, TransformerConfig

__all__ = ["DeepSpeedTransformerConfig", "DeepSpeedTransformerLayer", "TransformerConfig"]
