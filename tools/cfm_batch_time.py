"""Batched CFM solve timing: python tools/cfm_batch_time.py nb [L] -> ms per 10-step solve, ms per utterance, TFLOP/s.
CFM_STEPS / CFM_REPS / CFM_SAVE: Euler steps, timed repetitions, file for the result (tests/test_cfm_switches_gpu.py)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import synthetic as S
from chatterbox_b200.cfm import CausalConditionalCFM, ConditionalDecoder
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 16
L = int(sys.argv[2]) if len(sys.argv) > 2 else 812
est = ConditionalDecoder(S.make_state_dict("flow"), "cuda", prefix="decoder.estimator.", max_len=L, max_batch=nb)
cfm = CausalConditionalCFM(est, rand_noise=S.cfm_noise())
i = S.cfm_inputs(L)
mu, cond, spks = (torch.cat([i[k]] * nb, 0).cuda() for k in ("mu", "cond", "spks"))
steps, reps = int(os.environ.get("CFM_STEPS", "10")), int(os.environ.get("CFM_REPS", "3"))
for _ in range(reps):
    cfm.forward_batch(mu, None, steps, spks=spks, cond=cond)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    out = cfm.forward_batch(mu, None, steps, spks=spks, cond=cond)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
flop = nb * steps * 2 * L * (132.18e6 + 114688.0 * L)
print(f"nb={nb} L={L}: {ms:.2f} ms per solve, {ms / nb:.2f} ms per utterance, {flop / ms / 1e9:.1f} TFLOP/s")
o = out[0] if isinstance(out, (tuple, list)) else out
print(f"checksum {float(o.double().sum()):.6f} absmax {float(o.abs().max()):.4f} finite {bool(torch.isfinite(o).all())}")
if os.environ.get("CFM_SAVE"):
    torch.save(o.cpu(), os.environ["CFM_SAVE"])
