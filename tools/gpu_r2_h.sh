#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 200 python tools/cfm_once.py 812 10 2>/dev/null | tail -1
timeout 300 python tools/vc_time.py 750 150 2>&1 | tail -1
CBX_ATTN_QT=0 timeout 300 python tools/cfm_batch_time.py 64 2>/dev/null | tail -2 | head -1
timeout 900 python -m pytest tests/test_s3_attn_gpu.py tests/test_cfm_switches_gpu.py tests/test_cfm_gpu.py -m gpu -q 2>&1 | tail -2
