#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_moe_ep_gpu.py tests/test_gemm_ext_gpu.py -q > gpurun_out/pytest_moe.log 2>&1; echo "moe+gemm tests exit $?"; tail -4 gpurun_out/pytest_moe.log | cut -c1-300
timeout 600 python tools/gemm_table.py > gpurun_out/gemm_table.log 2>&1; echo "gemm table exit $?"; grep -E "dgrad" gpurun_out/gemm_table.log
timeout 500 $TR --master-port 29571 tools/step_profile.py --model llama3-8b --layers 32 --steps 2 --gaps --out gpurun_out/prof2_gated.json > gpurun_out/prof2_gated.log 2>&1; echo "prof gated exit $?"
SXB_Z3_GATED=0 timeout 500 $TR --master-port 29572 tools/step_profile.py --model llama3-8b --layers 32 --steps 2 --gaps --out gpurun_out/prof2_nogate.json > gpurun_out/prof2_nogate.log 2>&1; echo "prof nogate exit $?"
python - <<'PY'
import json
for n in ("gated","nogate"):
    try:
        d=json.load(open(f"gpurun_out/prof2_{n}.json"))
        print(n, "wall", round(d["wall_ms_per_step"],1), "kernel", round(d["gpu_kernel_ms_per_step"],1), "idle", d.get("main_stream_idle_ms_per_step"), "streams", d.get("streams_busy_ms_per_step"))
        for t in d["top"][:8]: print("   ", t)
        for g in d.get("gaps_by_neighbours",[])[:6]: print("   gap", g)
    except Exception as e: print(n, "failed", e)
PY
