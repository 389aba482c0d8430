"""Kernel-by-kernel timeline of a CFM solve under graph replay: %globaltimer stamps of CTA 0 of every GEMM / attention /
fused feed-forward launch (start, programmatic-launch wait returned, exit) -> how long each kernel runs after its inputs are
ready and how long the hand-over between consecutive kernels takes."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import synthetic as S, _lib
from chatterbox_b200.cfm import CausalConditionalCFM, ConditionalDecoder
L = int(sys.argv[1]) if len(sys.argv) > 1 else 812
est = ConditionalDecoder(S.make_state_dict("flow"), "cuda", prefix="decoder.estimator.", max_len=2048)
cfm = CausalConditionalCFM(est, rand_noise=S.cfm_noise())
inp = {k: v.cuda() for k, v in S.cfm_inputs(L).items()}
lib = C.CDLL(_lib.LIB_PATH)
N = 8192
bufs = {k: torch.zeros(N * 16, dtype=torch.int64, device="cuda") for k in ("gemm", "attn", "ffn")}
def arm():
    lib.cbx_dbg_gemm_stamps(C.c_void_p(bufs["gemm"].data_ptr()))
    lib.cbx_dbg_attn_stamps(C.c_void_p(bufs["attn"].data_ptr()))
    lib.cbx_dbg_ffn_stamps(C.c_void_p(bufs["ffn"].data_ptr()))
for _ in range(3):          # eager, capture, replay: every pass re-arms, so the captured launches carry the same slots
    arm()
    cfm(mu=inp["mu"], mask=None, n_timesteps=10, spks=inp["spks"], cond=inp["cond"])
torch.cuda.synchronize()
idx = {"gemm": (0, 11, 7), "attn": (0, 1, 15), "ffn": (0, 2, 11)}       # start, wait returned, exit
rec = []
for k, b in bufs.items():
    t = b.cpu().view(-1, 16)
    s, w, e = idx[k]
    for r in t[t[:, 0] > 0]:
        rec.append((int(r[s]), int(r[w]), int(r[e]), k))
rec.sort()
print(f"{len(rec)} stamped launches")
mid = len(rec) // 2
t0 = rec[mid][0]
print("kernel   start   inputs-ready   exit   | busy after inputs | hand-over from previous exit   (us, CTA 0 of each launch)")
for i in range(mid, mid + 16):
    s, w, e, k = rec[i]
    print(f"{k:5s} {(s - t0) / 1e3:8.2f} {(w - t0) / 1e3:8.2f} {(e - t0) / 1e3:8.2f} | {(e - w) / 1e3:6.2f} | {(w - rec[i - 1][2]) / 1e3:6.2f}")
import statistics as st
for k in ("gemm", "attn", "ffn"):
    busy = [(e - w) / 1e3 for s, w, e, kk in rec[100:] if kk == k and w > 0]
    hand = [(rec[i][1] - rec[i - 1][2]) / 1e3 for i in range(100, len(rec)) if rec[i][3] == k and rec[i][1] > 0]
    print(f"{k}: n = {len(busy)}, busy after inputs median {st.median(busy):.2f} us, hand-over median {st.median(hand):.2f} us")
# inside the GEMM: stamps of CTA 0 relative to "inputs ready" (pdl_wait returned in the producer lane)
t = bufs["gemm"].cpu().view(-1, 16).double() / 1e3
t = t[t[:, 0] > 0]
lab = [("producer issued all", 2), ("mma issued all", 3), ("accumulator ready", 4), ("drained", 5), ("barrier", 8), ("N-split: moments", 9), ("pushed", 10), ("stored", 12), ("all moments in", 13), ("epilogue done", 6), ("exit", 7)]
dur = (t[:, 7] - t[:, 11])
for name, lo, hi in (("q|k|v-like (busy < 7 us)", 0, 7), ("out-proj-like (8.5 - 10.5 us)", 8.5, 10.5), ("resnet-like (> 10.5 us)", 10.5, 99)):
    m = (dur > lo) & (dur < hi)
    if int(m.sum()) == 0: continue
    r = t[m]
    print(f"{name}: n = {int(m.sum())}: " + ", ".join(f"{nm} {float((r[:, k] - r[:, 11]).median()):.2f}" for nm, k in lab))
