from .config import LoRAConfig, QuantizationConfig
from .context_manager import Init, init_lora
from .optimized_linear import LoRAOptimizedLinear, OptimizedLinear
from .quantization import QuantizedLinear, QuantizedParameter
