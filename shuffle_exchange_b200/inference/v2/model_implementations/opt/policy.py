"""``OPTPolicy`` (reference ``model_implementations/opt/policy.py``): spec + name map live in ``..policies``."""
from ..policies import OPTPolicy  # noqa: F401
