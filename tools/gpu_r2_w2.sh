#!/bin/bash
# weight k-blocks multicast across clusters of row tiles (CBX_GEMM_WMC): A/B on the padded-batch solve
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
CBX_GEMM_WMC=1 CFM_SAVE=/tmp/b1.pt timeout 300 python tools/cfm_batch_time.py 16 2>/dev/null | tail -2
CBX_GEMM_WMC=2 CFM_SAVE=/tmp/b2.pt timeout 300 python tools/cfm_batch_time.py 16 2>/dev/null | tail -2
CBX_GEMM_WMC=4 CFM_SAVE=/tmp/b4.pt timeout 300 python tools/cfm_batch_time.py 16 2>/dev/null | tail -2
python - <<'PY'
import torch
a, b, c = torch.load("/tmp/b1.pt"), torch.load("/tmp/b2.pt"), torch.load("/tmp/b4.pt")
print("multicast 2 / 4 vs none: equal %s %s" % (bool(torch.equal(a, b)), bool(torch.equal(a, c))))
PY
CBX_GEMM_WMC=1 timeout 300 python tools/cfm_batch_time.py 64 2>/dev/null | tail -2 | head -1
CBX_GEMM_WMC=2 timeout 300 python tools/cfm_batch_time.py 64 2>/dev/null | tail -2 | head -1
CBX_GEMM_WMC=4 timeout 300 python tools/cfm_batch_time.py 64 2>/dev/null | tail -2 | head -1
