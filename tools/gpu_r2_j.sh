#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_flow_frontend.py tests/test_t3_front_gpu.py -m gpu -q -s > gpurun_out/pytest_fronts.log 2>&1; echo "fronts exit $?"
grep -E "rel-L2|max-abs|passed|failed|rror|assert" gpurun_out/pytest_fronts.log | tail -30
python - <<'PY' 2>&1 | tail -8
import torch, sys
sys.path.insert(0, '.')
from chatterbox_b200 import synthetic as S
from chatterbox_b200.t3 import T3, T3Cond
from chatterbox_b200.flow_frontend import FlowFrontend
def tm(f, n=20):
    for _ in range(5): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
t3 = T3(S.make_state_dict("t3", n_layers=1), "cuda", max_pos=256, n_layers=1)
c = T3Cond(**S.t3_cond()).to(device="cuda")
tt = torch.cat([S.text_tokens(64)] * 2, 0).cuda()
print(f"T3 conditioning front-end (150 prompt tokens, 66 text tokens, CFG pair): {tm(lambda: t3._prefill_embeds(c, tt, 0.5, 2)):.0f} us per utterance")
fe = FlowFrontend(S.make_state_dict("flow"), "cuda", max_tokens=1024)
ref = S.s3gen_ref(seed=4, prompt_len=150)
ref = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in ref.items()}
for n in (256, 750):
    tok = S.speech_tokens(n, seed=5).cuda()
    f = lambda: fe.flow_inputs(tok, ref["prompt_token"], ref["prompt_feat"], ref["embedding"])
    print(f"flow front-end T = {n + 150} tokens: {tm(f):.0f} us, {fe.launches()} launches")
PY
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_j.log 2>&1; echo "gpu suite exit $?"; tail -4 gpurun_out/pytest_gpu_j.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_j.json 2> gpurun_out/bench_j.err; echo "bench exit $?"; cat gpurun_out/bench_j.json | cut -c1-1500
