#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_flow_frontend.py tests/test_boundary.py -m gpu -q -s -x 2>&1 | grep -E "rel-L2|passed|failed|rror|assert" | tail -12
for v in 0 1; do CBX_RELATTN_TC=$v python - <<'PY' 2>&1 | tail -2
import torch, sys, os
sys.path.insert(0, '.')
from chatterbox_b200 import synthetic as S
from chatterbox_b200.flow_frontend import FlowFrontend
def tm(f, n=20):
    for _ in range(5): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
fe = FlowFrontend(S.make_state_dict("flow"), "cuda", max_tokens=1024)
ref = S.s3gen_ref(seed=4, prompt_len=150)
ref = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in ref.items()}
for n in (256, 750):
    tok = S.speech_tokens(n, seed=5).cuda()
    f = lambda: fe.flow_inputs(tok, ref["prompt_token"], ref["prompt_feat"], ref["embedding"])
    print(f"CBX_RELATTN_TC={os.environ.get('CBX_RELATTN_TC')}: flow front-end T = {n + 150} tokens: {tm(f):.0f} us")
PY
done
