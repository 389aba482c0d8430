#!/bin/bash
# ncu --set full of the padded-batch GEMM kernels (weight-stationary q|k|v and up-projection, one row-LayerNorm GEMM) at nb = 64
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
export CBX_CFM_GRAPH=0 CFM_STEPS=1 CFM_REPS=1
timeout 300 python tools/cfm_batch_time.py 64 > gpurun_out/ws_plain.log 2>&1 || { tail -3 gpurun_out/ws_plain.log; exit 1; }
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"s3_gemm_ws_kernel|s3_gemm_kernel" --launch-skip 60 --launch-count 6 \
  -o gpurun_out/prof_ws_r02 -f python tools/cfm_batch_time.py 64 > gpurun_out/ncu_ws.log 2>&1
tail -2 gpurun_out/ncu_ws.log
ncu -i gpurun_out/prof_ws_r02.ncu-rep --page raw --csv > gpurun_out/prof_ws_r02_raw.csv 2>/dev/null; wc -l gpurun_out/prof_ws_r02_raw.csv
