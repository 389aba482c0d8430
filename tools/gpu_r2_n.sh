#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
CBX_GEMM_NSPLIT=1 CFM_SAVE=/tmp/o0.pt timeout 300 python tools/cfm_once.py 812 10 2>&1 | tail -1
CBX_GEMM_NSPLIT=4 CFM_SAVE=/tmp/o1.pt timeout 300 python tools/cfm_once.py 812 10 2>&1 | tail -1
CBX_GEMM_NSPLIT=2 timeout 300 python tools/cfm_once.py 812 10 2>&1 | tail -1
python - <<'PY'
import torch
a, b = torch.load("/tmp/o0.pt"), torch.load("/tmp/o1.pt")
print("nsplit 4 vs 1: rel-L2 %.3e max-abs %.3e" % (float((a - b).norm() / a.norm()), float((a - b).abs().max())))
PY
timeout 900 python -m pytest tests/test_cfm_gpu.py tests/test_hift_gpu.py tests/test_flow_frontend.py tests/test_boundary.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/cfm_timeline.py 2>&1 | tail -22
