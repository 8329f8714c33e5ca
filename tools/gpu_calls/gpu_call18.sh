#!/bin/bash
# 2 GPUs: new tests, ZeRO-3 fused path with lazily zeroed gradient arenas, replica averaging two-phase at world 2
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_moe_ep_gpu.py tests/test_symm_emulated_gpu.py -x -q -m gpu 2>&1 | tail -4
timeout 120 python tools/sanitize_targets.py 2>&1 | tail -3
for m in 0 1 2; do SXB_ATTN_DEBUG=$m timeout 120 python tools/attn_debug.py 2>&1 | tail -1; done
timeout 300 $TR --master-port 29571 bench.py --gpus 2 --steps 5 --warmup 3 --no-extras > gpurun_out/bench2_lazyzero.json 2> gpurun_out/bench2_lazyzero.err; echo "bench2 exit $?"; tail -c 900 gpurun_out/bench2_lazyzero.json; grep -E "Error|Traceback" gpurun_out/bench2_lazyzero.err | head -3
timeout 120 $TR --master-port 29572 tools/replica_avg_bench.py > gpurun_out/replica_avg_bench2.log 2>&1; echo "replica exit $?"; tail -1 gpurun_out/replica_avg_bench2.log | cut -c1-900
