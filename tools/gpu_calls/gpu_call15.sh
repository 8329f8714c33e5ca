#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu 2>&1 | grep -E "FAILED|passed|failed|AssertionError: \(" | cut -c1-600 | head -40
