from .graph_profile import UnitProfiler
from .comm_profile import create_predictor, estimate_us_per_byte, run_all_gather, run_reduce_scatter
