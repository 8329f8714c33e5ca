"""``MixtralPolicy`` (reference ``model_implementations/mixtral/policy.py``): spec + name map live in ``..policies``."""
from ..policies import MixtralPolicy  # noqa: F401
