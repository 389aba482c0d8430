#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_s3_attn_gpu.py -q -s -x 2>&1 | grep -E "attention B|passed|failed|rror|assert" | tail -24
CBX_ATTN_TC=0 timeout 300 python tools/cfm_once.py 812 10 2>&1 | tail -1
CBX_ATTN_TC=1 timeout 300 python tools/cfm_once.py 812 10 2>&1 | tail -1
timeout 900 python -m pytest tests/test_cfm_gpu.py tests/test_flow_frontend.py tests/test_boundary.py -m gpu -x -q 2>&1 | tail -4
