#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/ncu_targets.py 2>&1 | tail -3 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:'gemm_pair_x_kernel|gemm_pair_kernel|rs_adam_emu|push_emu|pull_emu|attn_fwd_kernel|attn_bwd|all_gather_emu' \
  -s 30 -c 15 -f -o gpurun_out/prof_r2 python tools/ncu_targets.py > gpurun_out/ncu_r2.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_r2.log; ls -la gpurun_out/prof_r2.ncu-rep
