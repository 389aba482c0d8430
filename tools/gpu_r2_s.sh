#!/bin/bash
# one-pass softmax (S tile read once, O accumulated in TMEM, lazy rescale) vs the two-pass form
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_s3_attn_gpu.py -m gpu -x -q 2>&1 | tail -3
CBX_LIBCBX=$PWD/_variants/libcbx_op0.so CFM_SAVE=/tmp/b0.pt timeout 300 python tools/cfm_batch_time.py 16 2>&1 | tail -2
CFM_SAVE=/tmp/b1.pt timeout 300 python tools/cfm_batch_time.py 16 2>&1 | tail -2
python - <<'PY'
import torch
a, b = torch.load("/tmp/b0.pt"), torch.load("/tmp/b1.pt")
print("one-pass vs two-pass softmax, nb=16 solve: rel-L2 %.3e max-abs %.3e" % (float((a - b).norm() / a.norm()), float((a - b).abs().max())))
PY
timeout 300 python tools/cfm_batch_time.py 64 2>&1 | tail -2
echo "--- single utterance: two-pass / one-pass for NG=1 only / one-pass both"
CBX_LIBCBX=$PWD/_variants/libcbx_op0.so timeout 200 python tools/cfm_once.py 812 10 2>&1 | tail -1
timeout 200 python tools/cfm_once.py 812 10 2>&1 | tail -1
CFM_SAVE=/tmp/s3.pt timeout 200 python tools/cfm_once.py 812 10 $PWD/_variants/libcbx_op3.so 2>&1 | tail -1
CFM_SAVE=/tmp/s1.pt timeout 200 python tools/cfm_once.py 812 10 2>&1 | tail -1
python - <<'PY'
import torch
a, b = torch.load("/tmp/s1.pt"), torch.load("/tmp/s3.pt")
print("single utterance, one-pass NG=2 vs two-pass: rel-L2 %.3e" % float((a - b).norm() / a.norm()))
PY
CBX_LIBCBX=$PWD/_variants/libcbx_op3.so timeout 600 python -m pytest tests/test_s3_attn_gpu.py -m gpu -x -q 2>&1 | tail -3
