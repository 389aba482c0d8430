"""One prefill + one decode launch of the T3 megakernel (ncu target)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import synthetic as S
from chatterbox_b200.t3 import T3, T3Cond

ntok = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda:0")
t3 = T3(S.make_state_dict("t3"), dev, max_pos=512)
cond = T3Cond(**{k: v.to(dev) for k, v in S.t3_cond().items()})
text = torch.cat([S.text_tokens(64)] * 2, 0).to(dev)
kw = dict(max_new_tokens=ntok, do_sample=False, stop_on_eos=False, cfg_weight=0.5, temperature=0.8,
          repetition_penalty=1.2, min_p=0.05, top_p=1.0)
torch.cuda.synchronize(); t0 = time.time()
t3.inference(t3_cond=cond, text_tokens=text, **kw)
torch.cuda.synchronize()
print(f"prefill + {ntok} tokens: {(time.time() - t0) * 1e3:.1f} ms")
