#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_s3_attn_gpu.py -m gpu -q -s 2>&1 | grep -E "tcgen05 attention B=16|passed|failed|Error" | head
timeout 100 python tools/attn_batch_stamps.py 32 812 | head -1
CBX_ATTN_QT=0 timeout 100 python tools/attn_batch_stamps.py 32 812 | head -1
timeout 100 python tools/attn_batch_stamps.py 128 812 | head -1
CFM_SAVE=/tmp/b1.pt timeout 300 python tools/cfm_batch_time.py 16 2>&1 | tail -2
CBX_ATTN_QT=0 CFM_SAVE=/tmp/b0.pt timeout 300 python tools/cfm_batch_time.py 16 2>&1 | tail -2
python - <<'PY'
import torch
a, b = torch.load("/tmp/b0.pt"), torch.load("/tmp/b1.pt")
print("two-tile vs one-tile attention, nb=16 solve: rel-L2 %.3e max-abs %.3e equal %s" % (float((a - b).norm() / a.norm()), float((a - b).abs().max()), bool(torch.equal(a, b))))
PY
timeout 300 python tools/cfm_batch_time.py 64 2>&1 | tail -2
