from .quantizer import (Quantizer, ds_quantizer, quantize, dequantize, swizzle_quant, quantized_reduction,
                        CUDAQuantizer)
