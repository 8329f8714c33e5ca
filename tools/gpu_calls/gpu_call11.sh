#!/bin/bash
# 1 GPU: new attention modes (varlen, atoms), regression of the dense / scatter attention paths and the AutoTP / FPDT
# GEMM routing.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_attention_modes_gpu.py tests/test_gemm_routing_gpu.py -x -q -m gpu 2>&1 | tail -25
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_fused_ulysses_gpu.py tests/test_kernels2_gpu.py -x -q -m gpu 2>&1 | tail -8
timeout 300 python tools/gpu_selftest.py attn_perf 2>&1 | tail -12
