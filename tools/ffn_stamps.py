"""Per-phase %globaltimer stamps of the fused feed-forward kernel (csrc/s3_ffn.cu) for every launch of one estimator call."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import synthetic as S, _lib
from chatterbox_b200.cfm import ConditionalDecoder
L = int(sys.argv[1]) if len(sys.argv) > 1 else 812
est = ConditionalDecoder(S.make_state_dict("flow"), "cuda", prefix="decoder.estimator.", max_len=2048)
inp = {k: v.cuda() for k, v in S.cfm_inputs(L).items()}
x = torch.cat([inp["z"], inp["z"]], 0); mu = torch.cat([inp["mu"], inp["mu"] * 0], 0)
sp = torch.cat([inp["spks"], inp["spks"] * 0], 0); cd = torch.cat([inp["cond"], inp["cond"] * 0], 0)
t = torch.tensor([0.3, 0.3]); m = torch.ones(2, 1, L)
for _ in range(2): est(x, m, mu, t, sp, cd)
buf = torch.zeros(16 * 400, dtype=torch.int64, device="cuda")
lib = C.CDLL(_lib.LIB_PATH)
lib.cbx_dbg_ffn_stamps(C.c_void_p(buf.data_ptr()))
est(x, m, mu, t, sp, cd)
lib.cbx_dbg_ffn_stamps(None)
torch.cuda.synchronize()
b = buf.cpu().view(-1, 16)
n = int((b[:, 0] > 0).sum())
names = ["init+tmem", "pdl_wait(tma warp)", "first k-block landed", "last k-block landed", "D1 ready (all warps)", "epi1 done+sync",
         "D2 ready (all warps)", "drained", "cluster sync", "reduce+epilogue", "cluster sync 2",
         "OUT: D0 ready", "OUT: moments in", "OUT: x slice stored", "OUT: cluster barrier"]
d = b[:n].double() / 1e3
print(f"{n} fused feed-forward launches; stamps relative to kernel start (us), launches 8..13 and the mean")
for i in list(range(8, 14)) + [-1]:
    r = d[i] if i >= 0 else d[4:].mean(0)
    print("mean" if i < 0 else i, " ".join(f"{nm} {float(r[k + 1] - r[0]):.2f} |" for k, nm in enumerate(names) if float(r[k + 1]) > 0))
