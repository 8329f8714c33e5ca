#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/attn_debug.py 2>&1 | tail -1
SXB_LIB_PATH=build/libsxb200_spin.so timeout 120 python tools/attn_debug.py 2>&1 | tail -1
SXB_LIB_PATH=build/libsxb200_spin.so SXB_ATTN_DEBUG=2 timeout 120 python tools/attn_debug.py 2>&1 | tail -1
