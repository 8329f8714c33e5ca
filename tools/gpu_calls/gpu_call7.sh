#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 420 $TR --master-port 29561 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench8_default.json 2> gpurun_out/bench8_default.err; echo "bench8 exit $?"; tail -c 2600 gpurun_out/bench8_default.json; grep -i "watchdog\|error\|Traceback" gpurun_out/bench8_default.err | head -5
timeout 300 $TR --master-port 29571 tools/step_profile.py --model llama3-8b --layers 32 --steps 2 --gaps --out gpurun_out/prof8.json > gpurun_out/prof8.log 2>&1; echo "prof8 exit $?"
python - <<'PY'
import json
try:
    d=json.load(open("gpurun_out/prof8.json"))
    print("wall", round(d["wall_ms_per_step"],1), "kernel", round(d["gpu_kernel_ms_per_step"],1), "idle", d.get("main_stream_idle_ms_per_step"), d.get("streams_busy_ms_per_step"))
    for t in d["top"][:12]: print("   ", t['kernel'][:60], t['calls'], t['ms'])
    for g in d.get("gaps_by_neighbours",[])[:8]: print("   gap", g)
except Exception as e: print("failed", e)
PY
