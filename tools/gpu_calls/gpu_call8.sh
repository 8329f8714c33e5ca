#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_fused_ulysses_gpu.py tests/test_moe_ep_gpu.py tests/test_kernels_gpu.py -q -x > gpurun_out/pytest_new.log 2>&1; echo "new tests exit $?"; tail -12 gpurun_out/pytest_new.log | cut -c1-400
timeout 600 $TR --master-port 29559 tools/moe_ep_check.py > gpurun_out/moe_ep_check.log 2>&1; echo "moe ep check exit $?"; tail -4 gpurun_out/moe_ep_check.log | cut -c1-900
