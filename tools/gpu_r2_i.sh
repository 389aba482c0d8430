#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_t3_front_gpu.py -q -s > gpurun_out/pytest_front.log 2>&1; echo "front exit $?"
grep -E "max-abs|passed|failed|rror" gpurun_out/pytest_front.log | tail -12
timeout 1500 python -m pytest tests/test_t3_gpu.py tests/test_t3_batch_gpu.py tests/test_t3_align_gpu.py tests/test_t3_fp32w_gpu.py tests/test_boundary.py tests/test_integration_snippets.py -m gpu -x -q > gpurun_out/pytest_t3all.log 2>&1; echo "t3 exit $?"; tail -3 gpurun_out/pytest_t3all.log
python - <<'PY'
import torch, sys
sys.path.insert(0, '.')
from chatterbox_b200 import synthetic as S
from chatterbox_b200.t3 import T3, T3Cond
t3 = T3(S.make_state_dict("t3", n_layers=1), "cuda", max_pos=256, n_layers=1)
c = T3Cond(**S.t3_cond()).to(device="cuda")
tt = torch.cat([S.text_tokens(64)] * 2, 0).cuda()
for _ in range(5): t3._prefill_embeds(c, tt, 0.5, 2)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): t3._prefill_embeds(c, tt, 0.5, 2)
e1.record(); torch.cuda.synchronize()
print(f"T3 conditioning front-end (150 prompt tokens, 66 text tokens, CFG pair): {e0.elapsed_time(e1) / 20 * 1e3:.0f} us per utterance, 16 launches")
PY
timeout 300 python tools/t3_time.py 2>&1 | tail -1
