#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 150 $TR --master-port 29565 tools/replica_avg_bench.py > gpurun_out/replica_avg_bench.log 2>&1; echo "replica bench exit $?"; tail -2 gpurun_out/replica_avg_bench.log | cut -c1-1800
timeout 240 $TR --master-port 29561 bench.py --gpus 8 --model mixtral-8x7b --steps 4 --warmup 3 --no-e2e --moe-fused > gpurun_out/mixtral8_fused.json 2> gpurun_out/mixtral8_fused.err; echo "mixtral fused exit $?"; tail -c 1300 gpurun_out/mixtral8_fused.json; grep -E "Error|error|Traceback|overflow" gpurun_out/mixtral8_fused.err | head -6 | cut -c1-400
