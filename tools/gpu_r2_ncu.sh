#!/bin/bash
# round-2 ncu pass: --set full of (a) the T3 decode launch (256 tokens), (b) one launch each of the estimator kernels of the final tree
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 300 python tools/t3_once.py 256 > gpurun_out/t3_once.log 2>&1 || { echo "t3_once failed"; tail -5 gpurun_out/t3_once.log; exit 1; }
timeout 1500 ncu --set full --import-source on --clock-control none -k regex:t3_megakernel --launch-skip 0 --launch-count 1 \
  -o gpurun_out/prof_t3_r02 -f python tools/t3_once.py 256 > gpurun_out/ncu_t3.log 2>&1
tail -2 gpurun_out/ncu_t3.log
ncu -i gpurun_out/prof_t3_r02.ncu-rep --page raw --csv > gpurun_out/prof_t3_r02_raw.csv 2>/dev/null
CFM_REPS=1 timeout 300 python tools/cfm_once.py 812 1 > gpurun_out/cfm_once.log 2>&1 || { echo "cfm_once failed"; tail -5 gpurun_out/cfm_once.log; exit 1; }
CFM_REPS=1 timeout 900 ncu --set full --import-source on --clock-control none -k regex:"s3_gemm_kernel|s3_attn_tc_kernel|s3_ffn_kernel" --launch-skip 700 --launch-count 16 \
  -o gpurun_out/prof_cfm_r02 -f python tools/cfm_once.py 812 1 > gpurun_out/ncu_cfm.log 2>&1
tail -2 gpurun_out/ncu_cfm.log
ncu -i gpurun_out/prof_cfm_r02.ncu-rep --page raw --csv > gpurun_out/prof_cfm_r02_raw.csv 2>/dev/null
ls -la gpurun_out/prof_*
