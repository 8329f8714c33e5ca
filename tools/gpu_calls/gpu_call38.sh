#!/bin/bash
timeout 600 python -m pytest tests/test_engine_gpu.py tests/test_attention_modes_gpu.py tests/test_kernels_gpu.py -x -q -m gpu -k "fp16 or head_sizes or native_library" 2>&1 | tail -12 | cut -c1-400
