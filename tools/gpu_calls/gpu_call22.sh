#!/bin/bash
mkdir -p gpurun_out
export SXB_ATTN_FWD128=2
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_attention_modes_gpu.py tests/test_fused_ulysses_gpu.py -x -q -m gpu -k "attn or atom or varlen or packed or ulysses or llama" 2>&1 | tail -6
timeout 120 python tools/attn_debug.py 2>&1 | tail -1
SXB_ATTN_MC=0 timeout 120 python tools/attn_debug.py 2>&1 | tail -1
timeout 200 python tools/paged_attn_bench.py 2>&1 | tail -5 | cut -c1-330
