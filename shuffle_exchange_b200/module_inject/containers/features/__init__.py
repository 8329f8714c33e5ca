"""Container capabilities as mix-ins (reference: ``module_inject/containers/features``).  The base container here
already handles split q/k/v, gated MLPs and Megatron row orders generically; these classes add the behaviours that
are genuinely optional: loading from meta tensors, and keeping a fused inference layer in sync with a module that is
still being trained (hybrid engine), including LoRA folding."""
from .gated_mlp import HybridGatedMLPContainer
from .hybrid_engine import HybridEngineContainer
from .hybrid_megatron import HybridMegatronContainer
from .megatron import MegatronContainer
from .meta_tensor import MetaTensorContainer
from .split_qkv import HybridSplitQKVContainer
