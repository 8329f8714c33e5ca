from .quantize import FP_Quantize, Quantizer, fp_quantize, fp_dequantize
from .fp8_gemm import matmul_fp8
