#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_s3_attn_gpu.py -m gpu -x -q 2>&1 | tail -1
CBX_ATTN_QT=0 timeout 100 python tools/attn_batch_stamps.py 32 812 2>/dev/null | head -1
timeout 100 python tools/attn_batch_stamps.py 32 812 2>/dev/null | head -1
CBX_ATTN_QT=0 timeout 100 python tools/attn_batch_stamps.py 128 812 2>/dev/null | head -1
timeout 100 python tools/attn_batch_stamps.py 128 812 2>/dev/null | head -1
CBX_ATTN_QT=0 timeout 300 python tools/cfm_batch_time.py 64 2>/dev/null | tail -2
timeout 300 python tools/cfm_batch_time.py 64 2>/dev/null | tail -2
