#!/bin/bash
# 8 GPUs: replica averaging (two-phase kernel), Mixtral-8x7B EP=8 fused exchange, Llama-3-8B ZeRO-3 default bench
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 100 $TR --master-port 29565 tools/replica_avg_bench.py > gpurun_out/replica_avg_bench.log 2>&1; echo "replica bench exit $?"; tail -1 gpurun_out/replica_avg_bench.log | cut -c1-1500
timeout 200 $TR --master-port 29561 bench.py --gpus 8 --model mixtral-8x7b --steps 4 --warmup 3 --no-e2e --moe-fused > gpurun_out/mixtral8_fused.json 2> gpurun_out/mixtral8_fused.err; echo "mixtral fused exit $?"; tail -c 1200 gpurun_out/mixtral8_fused.json; grep -E "Error|error|Traceback|overflow" gpurun_out/mixtral8_fused.err | head -4 | cut -c1-300
timeout 200 $TR --master-port 29562 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench8_final.json 2> gpurun_out/bench8_final.err; echo "llama8 exit $?"; tail -c 2500 gpurun_out/bench8_final.json; grep -E "Error|Traceback" gpurun_out/bench8_final.err | head -3
