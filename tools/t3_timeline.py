"""All-CTA timeline of one decode iteration of the T3 megakernel: for every step, when each CTA finished its
prologue (inputs staged), its weight units and its epilogue — separates hand-over latency from CTA skew.
Needs a timeline build of the library: CBX_NVCC_EXTRA=-DT3_TIMELINE python -m chatterbox_b200.build --force"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import synthetic as S, _lib
if len(sys.argv) > 1:      # time another build of the library (python tools/t3_timeline.py _variants/libcbx_tl.so)
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
G = _lib.lib().cbx_t3_grid(0)
it = 200
st = torch.zeros(1024 + G * 1024, dtype=torch.int64, device=dev)
_lib.check(t3._lib.cbx_t3_set_profile(t3._h, st.data_ptr(), it | 0x40000000))
t3.inference(t3_cond=cond, text_tokens=text, **kw)
_lib.check(t3._lib.cbx_t3_set_profile(t3._h, None, -1))
tl = st[1024:].view(G, 256, 4).cpu().double() / 1e3          # [cta][step][prologue, consume, epilogue, attn] in us
names = ["qkv", "oproj", "gateup", "down"]
import statistics as stt
rows = {n: [] for n in names}
for step in range(4, 120):
    ph = step & 3
    pro, con, epi, att = (tl[:, step, k] for k in range(4))
    prev_epi = tl[:, step - 1, 2] if ph != 1 else tl[:, step - 1, 3]     # producer finish (attention end before oproj)
    prev_epi = prev_epi[prev_epi > 0]
    act = epi > 0
    rows[names[ph]].append(dict(
        prod_skew=float(prev_epi.max() - prev_epi.median()),                 # slowest producer vs median producer
        handover=float(pro[act].min() - prev_epi.max()),                     # last producer done -> first consumer staged
        stage_spread=float(pro[act].max() - pro[act].min()),
        frags_consume=float((con - pro)[act].median()),
        consume_max=float((con - pro)[act].max()),
        epilogue=float((epi - con)[act].median()),
        step_total=float(epi[act].max() - prev_epi.max())))
for n in names:
    r = rows[n]
    print(n, {k: round(stt.mean(x[k] for x in r), 2) for k in r[0]})
e0 = tl[:, 4, 2]; e1 = tl[:, 116, 2]
print("per-layer (steps 4..116):", float((e1.max() - e0.max()) / 28))
# attention sub-stamps (CTAs that own a (pair, split)): q staged / barrier / keys done / published, relative to the
# last qkv epilogue of the layer
at = st[1024:].view(G, 1024)[:, 512:512 + 120].view(G, 30, 4).cpu().double() / 1e3
acc = {k: [] for k in ("q_polled_med", "q_polled_max", "keys_done_max", "published_max", "keys_dur_med")}
for l in range(1, 29):
    qkv_done = tl[:, 4 * l, 2]; qkv_done = qkv_done[qkv_done > 0].max()
    a = at[:, l, :]; m = a[:, 0] > 0
    acc["q_polled_med"].append(float((a[m, 0] - qkv_done).median()))
    acc["q_polled_max"].append(float((a[m, 0] - qkv_done).max()))
    acc["keys_done_max"].append(float((a[m, 2] - qkv_done).max()))
    acc["keys_dur_med"].append(float((a[m, 2] - a[m, 1]).median()))
    acc["published_max"].append(float((tl[m, 4 * l, 3] - qkv_done).max()))
print("attention vs last qkv epilogue:", {k: round(stt.mean(v), 2) for k, v in acc.items()})
# head step (step 4 L): when do the CTAs stage / finish, relative to the last down epilogue of the last layer
hs = 120
d_end = tl[:, hs - 1, 2]; d_end = d_end[d_end > 0].max()
for nm, k in (("staged", 0), ("units done", 1), ("published", 2)):
    v = tl[:, hs, k]; v = v[v > 0] - d_end
    print(f"head {nm}: min {float(v.min()):.2f} median {float(v.median()):.2f} max {float(v.max()):.2f} us after the last down epilogue")
# attention phase, per CTA (medians / maxima over the CTAs that own a (pair, split), mean over layers 1..28):
# qkv epilogue of THIS CTA -> q staged -> barrier -> keys done -> published
seg = {k: [] for k in ("own_epi_to_q_staged med", "own_epi_to_q_staged max", "wait_all+bar med", "scores med", "softmax+pv+publish med", "softmax+pv+publish max",
                       "q_staged_to_published med", "q_staged_to_published max")}
for l in range(1, 29):
    a = at[:, l, :]; m = a[:, 0] > 0
    own = tl[m, 4 * l, 2]; pub = tl[m, 4 * l, 3]
    seg["own_epi_to_q_staged med"].append(float((a[m, 0] - own).median())); seg["own_epi_to_q_staged max"].append(float((a[m, 0] - own).max()))
    seg["wait_all+bar med"].append(float((a[m, 1] - a[m, 0]).median()))
    seg["scores med"].append(float((a[m, 2] - a[m, 1]).median()))
    seg["softmax+pv+publish med"].append(float((pub - a[m, 2]).median())); seg["softmax+pv+publish max"].append(float((pub - a[m, 2]).max()))
    seg["q_staged_to_published med"].append(float((pub - a[m, 0]).median())); seg["q_staged_to_published max"].append(float((pub - a[m, 0]).max()))
print("attention segments (us):", {k: round(stt.mean(v), 2) for k, v in seg.items()})
