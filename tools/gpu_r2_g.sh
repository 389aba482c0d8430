#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "suite exit $?"; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python tools/t3_time.py 2>&1 | tail -1
for u in 1 64; do timeout 300 python tools/t3b_time.py $u 2>&1 | tail -1; done
timeout 300 python tools/cfm_once.py 812 10 2>&1 | tail -1
timeout 300 python tools/cfm_batch_time.py 16 2>&1 | tail -1
