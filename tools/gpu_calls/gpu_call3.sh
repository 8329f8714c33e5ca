#!/bin/bash
# 2-GPU call: MoE EP tests (1 GPU), fused-vs-NCCL-vs-oracle training check, symm kernel test, Llama-3-8B bench A/B
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_moe_ep_gpu.py -q -x > gpurun_out/pytest_moe.log 2>&1; echo "moe tests exit $?"; tail -15 gpurun_out/pytest_moe.log
timeout 600 $TR --master-port 29558 tools/z3_fused_check.py > gpurun_out/z3_check.log 2>&1; echo "z3 check exit $?"; tail -3 gpurun_out/z3_check.log | cut -c1-1200
timeout 300 $TR --master-port 29555 tools/symm_test.py > gpurun_out/symm_test.log 2>&1; echo "symm test exit $?"; tail -3 gpurun_out/symm_test.log | cut -c1-600
timeout 400 $TR --master-port 29561 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench2_default.json 2> gpurun_out/bench2_default.err; echo "bench default exit $?"; tail -c 900 gpurun_out/bench2_default.json
SXB_Z3_GATED=0 timeout 400 $TR --master-port 29562 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench2_nogate.json 2> gpurun_out/bench2_nogate.err; echo "bench nogate exit $?"; tail -c 400 gpurun_out/bench2_nogate.json
SXB_Z3_GATED=0 SXB_Z3_LOOKAHEAD=0 timeout 400 $TR --master-port 29563 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e > gpurun_out/bench2_old.json 2> gpurun_out/bench2_old.err; echo "bench old exit $?"; tail -c 400 gpurun_out/bench2_old.json
