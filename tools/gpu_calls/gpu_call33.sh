#!/bin/bash
timeout 600 python -m pytest tests/test_attention_modes_gpu.py tests/test_kernels_gpu.py tests/test_inference_v1_ops.py tests/test_kernel_inject.py -x -q -m gpu 2>&1 | tail -6
