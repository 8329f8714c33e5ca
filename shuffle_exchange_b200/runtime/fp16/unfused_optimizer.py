"""FP16_UnfusedOptimizer (reference: deepspeed/runtime/fp16/unfused_optimizer.py) — per-parameter
fp32 masters for optimizers that need tensor-shaped state (LAMB).  Provided as an alias of the
flat implementation with one flat group per parameter group."""
from .fused_optimizer import FP16_Optimizer as FP16_UnfusedOptimizer  # noqa: F401
