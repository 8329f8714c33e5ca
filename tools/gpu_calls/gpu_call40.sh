#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/attn_debug.py 2>&1 | tail -1 | cut -c1-300
timeout 600 python bench.py --no-extras > gpurun_out/bench1_final.json 2> gpurun_out/bench1_final.err; echo "bench exit $?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench1_final.json').read().strip().split('\n')[-1])
print(d['value'], d['ms_per_step'], d['clocks'], d['e2e']['value'])
PY
