"""Decode time per token of the T3 megakernel (second call, CUDA events)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import _lib
NTOK = 256
for a in sys.argv[1:]:     # A/B: another build of the library (python tools/t3_time.py _variants/libcbx_x.so); a number = tokens to decode
    if a.isdigit():
        NTOK = int(a)
    else:
        _lib.LIB_PATH = os.path.abspath(a)
from chatterbox_b200 import synthetic as S
from chatterbox_b200.t3 import T3, T3Cond
dev = torch.device("cuda:0")
t3 = T3(S.make_state_dict("t3"), dev, max_pos=max(512, NTOK + 160))
cond = T3Cond(**{k: v.to(dev) for k, v in S.t3_cond().items()})
text = torch.cat([S.text_tokens(64)] * 2, 0).to(dev)
kw = dict(max_new_tokens=NTOK, do_sample=False, stop_on_eos=False, cfg_weight=0.5, temperature=0.8,
          repetition_penalty=1.2, min_p=0.05, top_p=1.0)
t3.record_events = True
for _ in range(3):
    t3.inference(t3_cond=cond, text_tokens=text, **kw)
torch.cuda.synchronize()
ev = t3.events
print("prefill %.2f ms, decode %.2f ms = %.1f us/token" % (ev["prefill0"].elapsed_time(ev["decode0"]), ev["decode0"].elapsed_time(ev["decode1"]), ev["decode0"].elapsed_time(ev["decode1"]) * 1e3 / NTOK))
