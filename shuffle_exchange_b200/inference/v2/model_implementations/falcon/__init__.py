"""``FalconPolicy`` (reference ``model_implementations/falcon/policy.py``): spec + name map live in ``..policies``."""
from ..policies import FalconPolicy  # noqa: F401
