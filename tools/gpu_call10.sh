#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29561 bench.py --gpus 8 --model mixtral-8x7b --steps 4 --warmup 3 --no-e2e > gpurun_out/mixtral8_fused.json 2> gpurun_out/mixtral8_fused.err; echo "mixtral fused exit $?"; tail -c 1300 gpurun_out/mixtral8_fused.json; grep -E "Error|error|Traceback|overflow" gpurun_out/mixtral8_fused.err | head -6 | cut -c1-400
timeout 600 $TR --master-port 29562 bench.py --gpus 8 --model mixtral-8x7b --steps 4 --warmup 3 --no-e2e --fp8-experts > gpurun_out/mixtral8_fp8.json 2> gpurun_out/mixtral8_fp8.err; echo "mixtral fp8 exit $?"; tail -c 700 gpurun_out/mixtral8_fp8.json; grep -E "Error|error|Traceback|overflow" gpurun_out/mixtral8_fp8.err | head -6 | cut -c1-400
