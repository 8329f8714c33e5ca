#!/bin/bash
SXB_ATTN_FWD128=2 timeout 120 python tools/attn_prof_fwd.py 2>&1 | tail -1
export SXB_ATTN_FWD128=2
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_attention_modes_gpu.py -x -q -m gpu -k "attn or atom or varlen or packed" 2>&1 | tail -2
timeout 120 python tools/attn_debug.py 2>&1 | tail -1
SXB_ATTN_DEBUG=4 timeout 120 python tools/attn_debug.py 2>&1 | tail -1
