from .comm import *  # noqa: F401,F403
from .comm import (Noop, ReduceOp, all_gather, all_gather_into_tensor, all_reduce, all_to_all_single, barrier,  # noqa: F401
                   broadcast, configure, comms_logger, get_rank, get_world_size, init_distributed, is_initialized,
                   new_group, recv, reduce_scatter_tensor, send, timed_op)
