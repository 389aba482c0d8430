#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
CBX_FFN_RED_L2=0 CFM_SAVE=/tmp/o0.pt timeout 300 python tools/cfm_once.py 812 10 2>&1 | tail -1
CBX_FFN_RED_L2=1 CFM_SAVE=/tmp/o1.pt timeout 300 python tools/cfm_once.py 812 10 2>&1 | tail -1
python - <<'PY'
import torch
a, b = torch.load("/tmp/o0.pt"), torch.load("/tmp/o1.pt")
print("reduce through L2 vs DSMEM: equal %s max-abs %.3e" % (bool(torch.equal(a, b)), float((a - b).abs().max())))
PY
timeout 300 python tools/ffn_stamps.py 2>&1 | tail -1
timeout 900 python -m pytest tests/test_cfm_gpu.py tests/test_flow_frontend.py tests/test_boundary.py -m gpu -x -q 2>&1 | tail -2
