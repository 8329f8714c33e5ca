#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu 2>&1 | tail -4
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 600 python bench.py --no-extras > gpurun_out/bench1_final.json 2> gpurun_out/bench1_final.err; echo "bench exit $?"; tail -c 600 gpurun_out/bench1_final.json
