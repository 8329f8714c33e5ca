#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_attention_modes_gpu.py tests/test_symm_emulated_gpu.py tests/test_kernels_gpu.py -x -q -m gpu 2>&1 | tail -6
timeout 300 python tools/paged_attn_bench.py 2>&1 | tail -6
