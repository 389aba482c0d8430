#!/bin/bash
# BASELINE configs 3, 4 (parts), 5 on the round-2 final tree
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
{ timeout 300 python tools/vc_time.py 750 150 2>&1 | tail -1
  timeout 300 python tools/t3_time.py 2000 2>&1 | tail -1
  timeout 300 python tools/t3_time.py 256 2>&1 | tail -1
  timeout 300 python tools/t3b_time.py 64 2>&1 | tail -1
  timeout 300 python tools/t3b_time.py 1 2>&1 | tail -1
  timeout 300 python tools/cfm_batch_time.py 64 2>&1 | tail -2 | head -1
  timeout 300 python tools/cfm_batch_time.py 16 2>&1 | tail -2 | head -1
  timeout 300 python tools/cfm_batch_time.py 1 2>&1 | tail -2 | head -1
  CBX_ATTN_TC=0 timeout 300 python tools/cfm_batch_time.py 64 2>&1 | tail -2 | head -1
} | tee gpurun_out/r02_z_configs.log
