"""``DeepSpeedFlopsProfilerConfig`` built from the whole config dictionary (reference: ``profiling/config.py``)."""
from ..runtime.config import FlopsProfilerConfig
from .constants import FLOPS_PROFILER


class DeepSpeedFlopsProfilerConfig(FlopsProfilerConfig):

    def __init__(self, param_dict=None, **kw):
        section = dict((param_dict or {}).get(FLOPS_PROFILER, {})) if param_dict is not None else {}
        section.update(kw)
        super().__init__(**section)
