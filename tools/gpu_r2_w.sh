#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_s3_attn_gpu.py -m gpu -x -q 2>&1 | tail -1
timeout 100 python tools/attn_batch_stamps.py 32 812 2>/dev/null | head -1
timeout 100 python tools/attn_batch_stamps.py 128 812 2>/dev/null | head -1
CBX_ATTN_QT=0 CFM_SAVE=/tmp/b0.pt timeout 300 python tools/cfm_batch_time.py 16 2>/dev/null | tail -2
CFM_SAVE=/tmp/b1.pt timeout 300 python tools/cfm_batch_time.py 16 2>/dev/null | tail -2
python - <<'PY'
import torch
a, b = torch.load("/tmp/b0.pt"), torch.load("/tmp/b1.pt")
print("persistent two-tile vs one-tile attention, nb=16 solve: equal %s rel %.2e" % (bool(torch.equal(a, b)), float((a-b).norm()/a.norm())))
PY
timeout 300 python tools/cfm_batch_time.py 64 2>/dev/null | tail -2
