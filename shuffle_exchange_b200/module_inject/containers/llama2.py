"""LLaMA-2 (Meta checkpoint format) policy and container under the reference's module path (``containers/llama2.py``)."""
from .llama import DS_LLAMA2Container, LLAMA2LayerPolicy  # noqa: F401
