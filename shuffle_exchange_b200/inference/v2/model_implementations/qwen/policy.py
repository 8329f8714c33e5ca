"""``QwenPolicy`` (reference ``model_implementations/qwen/policy.py``): spec + name map live in ``..policies``."""
from ..policies import QwenPolicy  # noqa: F401
