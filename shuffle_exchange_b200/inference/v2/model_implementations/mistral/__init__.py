"""``MistralPolicy`` (reference ``model_implementations/mistral/policy.py``): spec + name map live in ``..policies``."""
from ..policies import MistralPolicy  # noqa: F401
