#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_t3_align_gpu.py -q -s > gpurun_out/pytest_align.log 2>&1; echo "align exit $?"
grep -E "max-abs|decoded|passed|failed|rror|assert" gpurun_out/pytest_align.log | tail -20
timeout 900 python -m pytest tests/test_t3_gpu.py tests/test_t3_fp32w_gpu.py -m gpu -q > gpurun_out/pytest_t3.log 2>&1; echo "t3 exit $?"; tail -3 gpurun_out/pytest_t3.log
timeout 300 python tools/t3_time.py 2>&1 | tail -1
timeout 300 python tools/t3_time.py 2000 2>&1 | tail -1
