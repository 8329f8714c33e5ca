#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_moe_ep_gpu.py -q > gpurun_out/pytest_moe.log 2>&1; echo "moe tests exit $?"; tail -12 gpurun_out/pytest_moe.log | cut -c1-300
timeout 600 python tools/gemm_table.py > gpurun_out/gemm_table.log 2>&1; echo "gemm table exit $?"; tail -16 gpurun_out/gemm_table.log
