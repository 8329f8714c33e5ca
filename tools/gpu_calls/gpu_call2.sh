#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_ext_gpu.py tests/test_symm_emulated_gpu.py -q -x > gpurun_out/pytest_new.log 2>&1; echo "new tests exit $?"
tail -25 gpurun_out/pytest_new.log
timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_gemm_ext_gpu.py --deselect tests/test_symm_emulated_gpu.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -8 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_ref_1gpu.json 2> gpurun_out/bench_ref_1gpu.err; echo "ref exit $?"
tail -c 1200 gpurun_out/bench_ref_1gpu.json
