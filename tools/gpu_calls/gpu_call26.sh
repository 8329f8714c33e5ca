#!/bin/bash
# 8 GPUs: Mixtral-8x7B EP=8 with the fused exchange (capacity factor 2.0, recompute sized from a probe step)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 230 $TR --master-port 29561 bench.py --gpus 8 --model mixtral-8x7b --steps 4 --warmup 3 --no-e2e --moe-fused > gpurun_out/mixtral8_fused.json 2> gpurun_out/mixtral8_fused.err; echo "mixtral fused exit $?"; tail -c 1500 gpurun_out/mixtral8_fused.json; grep -E "Error|error|Traceback|overflow" gpurun_out/mixtral8_fused.err | head -4 | cut -c1-400
