#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -12
timeout 600 python bench.py --steps 5 --warmup 3 --no-extras > gpurun_out/bench1_lazyzero.json 2> gpurun_out/bench1_lazyzero.err; echo "bench exit $?"; tail -c 1500 gpurun_out/bench1_lazyzero.json
