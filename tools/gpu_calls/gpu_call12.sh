#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gemm_routing_gpu.py -x -q -m gpu 2>&1 | tail -8
timeout 300 python tools/paged_attn_bench.py 2>&1 | tail -8
timeout 600 python tools/step_profile.py --model llama3-8b --layers 32 --steps 2 --trace-kernel FillFunctor,elementwise_kernel,CUDAFunctor_add --out gpurun_out/step_profile_1gpu_trace.json 2>&1 | tail -30
