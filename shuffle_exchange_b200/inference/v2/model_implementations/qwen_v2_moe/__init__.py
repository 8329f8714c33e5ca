"""``Qwen2MoePolicy`` (reference ``model_implementations/qwen_v2_moe/policy.py``): spec + name map live in ``..policies``."""
from ..policies import Qwen2MoePolicy  # noqa: F401
