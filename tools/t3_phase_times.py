"""Per-phase timing of one decode iteration of the T3 megakernel (CTA 0 %globaltimer stamps).
Needs a profiling build of the library: CBX_NVCC_EXTRA=-DT3_PROFILE python -m chatterbox_b200.build --force"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import synthetic as S, _lib
if len(sys.argv) > 1:      # python tools/t3_phase_times.py _variants/libcbx_prof.so
    _lib.LIB_PATH = os.path.abspath(sys.argv[1])
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
    st = torch.zeros(512, dtype=torch.int64, device=dev)
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
    sp = [v / 1e3 for v in s[300:304]]
    print('  sample phase (CTA 0): wait+process logits %.2f, argmax %.2f, token+embedding publish %.2f us' % tuple(sp[i + 1] - sp[i] for i in range(3)))
    sub = s[200:215]
    lab = ["A:stage", "A:rmsnorm+frags", "A:consume", "A:cbar", "A:epilogue", "A:cbar2", "-", "-", "D:stage", "D:rmsnorm+frags",
           "D:consume", "D:cbar", "D:epilogue", "D:cbar2"]
    print("  sub-phase (layer 5, CTA 0 thread 0):", ", ".join(f"{lab[i]} {(sub[i+1]-sub[i])/1e3:.2f}" for i in range(14) if lab[i] != "-"))
    dd = s[216:221]
    print("  D rmsnorm detail: sumsq+shfl+sts %.2f, cp.async.wait %.2f, cbar %.2f, scale+ln %.2f; stage_end->rms_start %.2f, rms_end->frags_end %.2f" % (
        (dd[1]-dd[0])/1e3, (dd[2]-dd[1])/1e3, (dd[3]-dd[2])/1e3, (dd[4]-dd[3])/1e3, (dd[0]-s[209])/1e3, (s[210]-dd[4])/1e3))
    print(f"  ring (CTA 0 warp 0, whole iteration): waiting on full barriers {s[230]/1e3:.1f} us, load+mma+arrive {s[231]/1e3:.1f} us, units so far {s[232]}")
_lib.check(t3._lib.cbx_t3_set_profile(t3._h, None, -1))
