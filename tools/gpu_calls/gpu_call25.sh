#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 200 python tools/w6a16_bench.py 2>&1 | tail -9 | cut -c1-260
timeout 120 python tools/sanitize_targets.py 2>&1 | tail -1 | cut -c1-200
timeout 420 compute-sanitizer --tool racecheck --print-limit 30 python tools/sanitize_targets.py > gpurun_out/racecheck.log 2>&1; echo "racecheck exit $?"; grep -E "RACECHECK SUMMARY|ERROR SUMMARY|sanitize targets ok|Error|hazard" gpurun_out/racecheck.log | head -12 | cut -c1-300
