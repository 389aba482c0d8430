#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
echo "=== default build (one-pass for NG=1), shape-chosen"; timeout 600 python -m pytest tests/test_s3_attn_gpu.py -m gpu -q -s 2>&1 | grep -E "tcgen05 attention|passed|failed|Error" | head -40
echo "=== default build, NG=1 forced"; CBX_ATTN_GROUPS=1 timeout 600 python -m pytest tests/test_s3_attn_gpu.py -m gpu -q -s -k "fp32_torch and 1]" 2>&1 | grep -E "tcgen05 attention|passed|failed|Error" | head -40
timeout 300 python tools/cfm_batch_time.py 64 2>&1 | tail -2
timeout 300 python tools/cfm_batch_time.py 16 2>&1 | tail -2
