#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
CBX_GEMM_WS=0 CFM_SAVE=/tmp/b0.pt timeout 300 python tools/cfm_batch_time.py 16 2>&1 | tail -2
CBX_GEMM_WS=1 CFM_SAVE=/tmp/b1.pt timeout 300 python tools/cfm_batch_time.py 16 2>&1 | tail -2
python - <<'PY'
import torch
a, b = torch.load("/tmp/b0.pt"), torch.load("/tmp/b1.pt")
print("weight-stationary vs tile-per-CTA GEMM: rel-L2 %.3e max-abs %.3e equal %s" % (float((a - b).norm() / a.norm()), float((a - b).abs().max()), bool(torch.equal(a, b))))
PY
CBX_GEMM_WS=1 timeout 300 python tools/cfm_batch_time.py 64 2>&1 | tail -2
