#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gemm_ext_gpu.py -q > gpurun_out/pytest_gemm.log 2>&1; echo "gemm tests exit $?"; tail -3 gpurun_out/pytest_gemm.log | cut -c1-300
timeout 600 python tools/gemm_table.py > gpurun_out/gemm_table.log 2>&1; echo "gemm table exit $?"; tail -13 gpurun_out/gemm_table.log
timeout 400 $TR --master-port 29561 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench2_default.json 2> gpurun_out/bench2_default.err; echo "bench gated exit $?"; tail -c 500 gpurun_out/bench2_default.json
SXB_Z3_GATED=0 timeout 400 $TR --master-port 29562 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench2_nogate.json 2> gpurun_out/bench2_nogate.err; echo "bench nogate exit $?"; tail -c 500 gpurun_out/bench2_nogate.json
