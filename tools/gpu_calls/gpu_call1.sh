#!/bin/bash
# first GPU call of the round: GPU test tier, then both bench arms (short)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/bench_ours_1gpu.json 2> gpurun_out/bench_ours_1gpu.err; echo "ours exit $?"
tail -c 1500 gpurun_out/bench_ours_1gpu.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_ref_1gpu.json 2> gpurun_out/bench_ref_1gpu.err; echo "ref exit $?"
tail -c 1500 gpurun_out/bench_ref_1gpu.json; tail -5 gpurun_out/bench_ref_1gpu.err
