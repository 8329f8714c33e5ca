"""MoE container base under the reference's module path (``containers/base_moe.py``)."""
from .megatron_gpt_moe import BaseTransformerMoEContainer  # noqa: F401
