#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_moe_ep_gpu.py tests/test_kernels_gpu.py tests/test_attention_modes_gpu.py tests/test_fused_ulysses_gpu.py -x -q -m gpu 2>&1 | tail -5
timeout 120 python tools/attn_debug.py 2>&1 | tail -1
SXB_ATTN_MC=0 timeout 120 python tools/attn_debug.py 2>&1 | tail -1
timeout 200 python tools/attn_stress.py 10 2>&1 | tail -3
