#!/bin/bash
# 8 GPUs: Mixtral-8x7B EP=8 (fused vs NCCL exchange), then Llama-3-8B Ulysses SP=8 at 128K
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29561 bench.py --gpus 8 --model mixtral-8x7b --steps 4 --warmup 3 --no-e2e > gpurun_out/mixtral8_fused.json 2> gpurun_out/mixtral8_fused.err; echo "mixtral fused exit $?"; tail -c 1500 gpurun_out/mixtral8_fused.json; grep -E "Error|error|Traceback|overflow" gpurun_out/mixtral8_fused.err | head -8
timeout 600 $TR --master-port 29562 bench.py --gpus 8 --model mixtral-8x7b --steps 4 --warmup 3 --no-e2e --moe-nccl > gpurun_out/mixtral8_nccl.json 2> gpurun_out/mixtral8_nccl.err; echo "mixtral nccl exit $?"; tail -c 600 gpurun_out/mixtral8_nccl.json
timeout 600 $TR --master-port 29563 bench.py --gpus 8 --sp 8 --seq 131072 --steps 3 --warmup 3 --no-e2e > gpurun_out/ulysses8_128k.json 2> gpurun_out/ulysses8_128k.err; echo "ulysses exit $?"; tail -c 1500 gpurun_out/ulysses8_128k.json; grep -E "Error|error|Traceback" gpurun_out/ulysses8_128k.err | head -8
