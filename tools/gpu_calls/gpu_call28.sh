#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/ncu_targets.py 2>&1 | tail -2 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:'gemm_pair_x_kernel|gemm_pair_kernel|rs_adam_emu|push_emu|pull_emu|attn_fwd_kernel|attn_bwd|all_gather_emu|fp_gemv' \
  -s 36 -c 18 -f -o gpurun_out/prof_r2b python tools/ncu_targets.py > gpurun_out/ncu_r2b.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/ncu_r2b.log; ls -la gpurun_out/prof_r2b.ncu-rep
