"""Per-phase timing of one decode iteration of the T3 megakernel (CTA 0 %globaltimer stamps)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import synthetic as S, _lib
from chatterbox_b200.t3 import T3, T3Cond

dev = torch.device("cuda:0")
t3 = T3(S.make_state_dict("t3"), dev, max_pos=512)
cond = T3Cond(**{k: v.to(dev) for k, v in S.t3_cond().items()})
text = torch.cat([S.text_tokens(64)] * 2, 0).to(dev)
kw = dict(max_new_tokens=256, do_sample=False, stop_on_eos=False, cfg_weight=0.5, temperature=0.8,
          repetition_penalty=1.2, min_p=0.05, top_p=1.0)
for _ in range(2):
    t3.inference(t3_cond=cond, text_tokens=text, **kw)
for it in (8, 200):
    st = torch.zeros(256, dtype=torch.int64, device=dev)
    _lib.check(t3._lib.cbx_t3_set_profile(t3._h, st.data_ptr(), it))
    t3.inference(t3_cond=cond, text_tokens=text, **kw)
    s = st.cpu().tolist()
    n = 2 + 150 + 1
    d = [s[i + 1] - s[i] for i in range(n - 1)]
    names = ["sample"] + ["qkv", "attn", "oproj", "gateup", "down"] * 30 + ["head"]
    agg = {}
    for nm, v in zip(names, d):
        agg.setdefault(nm, []).append(v)
    print(f"iteration {it}: total {sum(d)/1e3:.1f} us")
    for nm, v in agg.items():
        v2 = sorted(v)
        print(f"  {nm:7s} n={len(v):3d} mean {sum(v)/len(v)/1e3:6.2f} us  min {v2[0]/1e3:6.2f}  max {v2[-1]/1e3:6.2f}  sum {sum(v)/1e3:7.1f}")
_lib.check(t3._lib.cbx_t3_set_profile(t3._h, None, -1))
