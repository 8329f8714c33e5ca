"""``PhiPolicy`` (reference ``model_implementations/phi/policy.py``): spec + name map live in ``..policies``."""
from ..policies import PhiPolicy  # noqa: F401
