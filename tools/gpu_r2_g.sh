#!/bin/bash
# cheaper GELU: timing (single + batch), then the whole GPU suite
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 200 python tools/cfm_once.py 812 10 2>/dev/null | tail -1
timeout 300 python tools/cfm_batch_time.py 64 2>/dev/null | tail -2 | head -1
timeout 2400 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|Error" gpurun_out/pytest_gpu.log | tail -5
grep -E "rel-L2|max-abs" gpurun_out/pytest_gpu.log | grep -i "cfm\|solve\|estimator\|mel" | head -12
