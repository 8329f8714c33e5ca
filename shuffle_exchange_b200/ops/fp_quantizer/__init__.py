from .quantize import FP_Quantize, Quantizer, fp_quantize, fp_dequantize, fp_gemv, fp_gemv_ok
from .fp8_gemm import matmul_fp8
