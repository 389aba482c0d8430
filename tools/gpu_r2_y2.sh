#!/bin/bash
# one-pass softmax in the two-group (single-utterance) attention too: A/B, then the whole GPU suite
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
CBX_LIBCBX=$PWD/_variants/libcbx_op1.so timeout 200 python tools/cfm_once.py 812 10 2>/dev/null | tail -1
timeout 200 python tools/cfm_once.py 812 10 2>/dev/null | tail -1
CBX_LIBCBX=$PWD/_variants/libcbx_op1.so timeout 200 python tools/cfm_once.py 1800 10 2>/dev/null | tail -1
timeout 200 python tools/cfm_once.py 1800 10 2>/dev/null | tail -1
timeout 2400 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
grep -E "passed|failed|Error" gpurun_out/pytest_gpu.log | tail -5
