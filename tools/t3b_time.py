"""Batched decode timing: python tools/t3b_time.py U [n_tokens] -> ms per decode step, tokens/s, HBM GB/s of the step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import synthetic as S
from chatterbox_b200.t3 import T3, T3Cond
from chatterbox_b200.t3_batch import T3Batch, T3Request
U = int(sys.argv[1]) if len(sys.argv) > 1 else 64
N = int(sys.argv[2]) if len(sys.argv) > 2 else 256
dev = torch.device("cuda:0")
max_pos = 34 + 66 + 2 + N + 8
t3 = T3(S.make_state_dict("t3"), dev, max_pos=max_pos)
eng = T3Batch(t3, U, max_pos=max_pos, max_new_cap=N)
P = dict(cfg_weight=0.5, temperature=0.8, repetition_penalty=1.2, min_p=0.05, top_p=1.0)
reqs = [T3Request(t3_cond=T3Cond(**S.t3_cond(seed=100 + k)).to(device=dev), text_tokens=torch.cat([S.text_tokens(64, seed=100 + k)] * 2, 0),
                  max_new_tokens=N, do_sample=False, stop_on_eos=False, **P) for k in range(U)]
e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
for rep in range(3):
    e[0].record()
    for k in range(U):
        eng.admit(k, reqs[k])
    e[1].record()
    eng.step(N)
    e[2].record()
    torch.cuda.synchronize()
    n_gen, done, _ = eng.poll()
    assert n_gen == [N] * U, n_gen
    for k in range(U):
        eng.release(k)
pf, dm = e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2])
Pl = eng.prefill_len[0]
by = sum(t3.packed.weight_bytes_per_step() + U * 245760 * (Pl + i + 1) for i in range(N - 1))
print(f"U={U} N={N} kv_splits={eng.info()['kv_splits']}: prefill {pf:.1f} ms ({pf / U:.2f}/utt), decode {dm:.1f} ms = {dm / N * 1e3:.0f} us/step, "
      f"{U * N / dm * 1e3:.0f} tokens/s, {by / dm / 1e6:.0f} GB/s")
