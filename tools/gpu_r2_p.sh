#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
CBX_CFM_OUTFFN=0 CFM_SAVE=/tmp/o0.pt timeout 300 python tools/cfm_once.py 812 10 2>&1 | tail -1
CBX_CFM_OUTFFN=1 CFM_SAVE=/tmp/o1.pt timeout 300 python tools/cfm_once.py 812 10 2>&1 | tail -1
python - <<'PY'
import torch
a, b = torch.load("/tmp/o0.pt"), torch.load("/tmp/o1.pt")
print("out-proj fused vs separate: rel-L2 %.3e max-abs %.3e finite %s" % (float((a - b).norm() / a.norm()), float((a - b).abs().max()), bool(torch.isfinite(b).all())))
PY
CBX_CFM_OUTFFN=1 timeout 900 python -m pytest tests/test_cfm_gpu.py tests/test_flow_frontend.py tests/test_boundary.py -m gpu -x -q 2>&1 | tail -3
