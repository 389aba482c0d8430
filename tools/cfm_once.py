"""One CFM solve at the bench shape (L = 812) — a short target for ncu captures of the GEMM / attention kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import synthetic as S, _lib
if len(sys.argv) > 3:      # A/B: python tools/cfm_once.py 812 10 _variants/libcbx_x.so
    _lib.LIB_PATH = os.path.abspath(sys.argv[3])
from chatterbox_b200.cfm import CausalConditionalCFM, ConditionalDecoder

L = int(sys.argv[1]) if len(sys.argv) > 1 else 812
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
est = ConditionalDecoder(S.make_state_dict("flow"), "cuda", prefix="decoder.estimator.", max_len=1024)
cfm = CausalConditionalCFM(est, rand_noise=S.cfm_noise())
inp = {k: v.cuda() for k, v in S.cfm_inputs(L).items()}
reps = int(os.environ.get("CFM_REPS", "5"))
for _ in range(3):
    out = cfm(mu=inp["mu"], mask=None, n_timesteps=steps, spks=inp["spks"], cond=inp["cond"])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    out = cfm(mu=inp["mu"], mask=None, n_timesteps=steps, spks=inp["spks"], cond=inp["cond"])
e1.record()
torch.cuda.synchronize()
o = out[0] if isinstance(out, (tuple, list)) else out
print(f"cfm solve L={L} steps={steps}: {e0.elapsed_time(e1) / reps:.2f} ms, launches so far {est.launches()}, "
      f"checksum {float(o.double().sum()):.6f} absmax {float(o.abs().max()):.4f}")
if os.environ.get("CFM_SAVE"):
    torch.save(o.cpu(), os.environ["CFM_SAVE"])
