"""``Phi3Policy`` (reference ``model_implementations/phi3/policy.py``): spec + name map live in ``..policies``."""
from ..policies import Phi3Policy  # noqa: F401
