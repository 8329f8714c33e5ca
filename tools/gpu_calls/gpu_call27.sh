#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels2_gpu.py tests/test_moe_ep_gpu.py -x -q -m gpu 2>&1 | tail -3
timeout 200 python tools/w6a16_bench.py 2>&1 | tail -9 | cut -c1-260
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/bench1_final.json 2> gpurun_out/bench1_final.err; echo "bench exit $?"; tail -c 2200 gpurun_out/bench1_final.json
