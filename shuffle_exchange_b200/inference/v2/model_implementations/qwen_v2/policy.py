"""``Qwen2Policy`` (reference ``model_implementations/qwen_v2/policy.py``): spec + name map live in ``..policies``."""
from ..policies import Qwen2Policy  # noqa: F401
