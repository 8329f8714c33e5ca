#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29583 bench.py --impl reference --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_ref_2gpu.json 2> gpurun_out/bench_ref_2gpu.err; echo "ref exit $?"; tail -c 1500 gpurun_out/bench_ref_2gpu.json; grep -E "Error|Traceback" gpurun_out/bench_ref_2gpu.err | head -5 | cut -c1-300
