"""Post-init group-wise INT4/INT8 weight quantisation for inference (reference ``deepspeed/inference/quantization``)."""
from .quantization import _init_group_wise_weight_quantization
from .layers import QuantizedEmbedding, QuantizedLinear, QUANTIZATION_LAYER_MAPPINGS
from .utils import DeQuantizer, Quantizer
