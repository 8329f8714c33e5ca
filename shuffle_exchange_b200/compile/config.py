"""``compile`` section of the config (reference ``deepspeed/compile/config.py:9-60``, same field names)."""
from ..runtime.config import CompileConfig  # noqa: F401
