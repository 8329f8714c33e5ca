"""``shuffle_exchange_b200.pipe``: the user-facing pipeline names (reference: ``deepspeed/pipe/__init__.py``)."""
from ..runtime.pipe import LayerSpec, PipelineModule, ProcessTopology, TiedLayerSpec  # noqa: F401
