#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_s3_attn_gpu.py -q -s -x 2>&1 | grep -E "tcgen05 attention B|passed|failed|rror|assert" | tail -12
timeout 300 python tools/attn_stamps.py 2>&1 | tail -3
CBX_ATTN_TC=0 timeout 300 python tools/cfm_once.py 812 10 2>&1 | tail -1
CBX_ATTN_TC=1 timeout 300 python tools/cfm_once.py 812 10 2>&1 | tail -1
