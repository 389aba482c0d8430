"""Per-phase %globaltimer stamps of the tcgen05 GEMM kernel for every launch of one estimator call."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import synthetic as S, _lib
from chatterbox_b200.cfm import ConditionalDecoder
L = 812
est = ConditionalDecoder(S.make_state_dict("flow"), "cuda", prefix="decoder.estimator.", max_len=1024)
inp = {k: v.cuda() for k, v in S.cfm_inputs(L).items()}
x = torch.cat([inp["z"], inp["z"]], 0); mu = torch.cat([inp["mu"], inp["mu"] * 0], 0)
sp = torch.cat([inp["spks"], inp["spks"] * 0], 0); cd = torch.cat([inp["cond"], inp["cond"] * 0], 0)
t = torch.tensor([0.3, 0.3]); m = torch.ones(2, 1, L)
for _ in range(2): est(x, m, mu, t, sp, cd)
buf = torch.zeros(16 * 400, dtype=torch.int64, device="cuda")
lib = C.CDLL(_lib.LIB_PATH)
lib.cbx_dbg_gemm_stamps(C.c_void_p(buf.data_ptr()))
est(x, m, mu, t, sp, cd)
lib.cbx_dbg_gemm_stamps(None)
torch.cuda.synchronize()
b = buf.cpu().view(-1, 16)
n = int((b[:, 0] > 0).sum())
names = ["alloc+init", "tma issue (from start)", "mma commit (from start)", "accum ready (from start)", "epi phase1", "epi phase2", "tail"]
print(f"{n} gemm launches; first 14 and a summary (us):")
for i in list(range(14)):
    r = b[i].tolist()
    print(i, f"init {(r[1]-r[0])/1e3:.1f} tma_done {(r[2]-r[0])/1e3:.1f} mma_done {(r[3]-r[0])/1e3:.1f} accum {(r[4]-r[0])/1e3:.1f} "
             f"ph1 {(r[5]-r[4])/1e3:.1f} bar {(r[8]-r[5])/1e3:.1f} ld {(r[9]-r[8])/1e3:.1f} math {(r[10]-r[9])/1e3:.1f} st {(r[6]-r[10])/1e3:.1f} ph2 {(r[6]-r[5])/1e3:.1f} end {(r[7]-r[0])/1e3:.1f}  gap_to_next_start {(b[i+1][0]-r[7]).item()/1e3:.1f}")
d = b[:n].double()
print("mean: init %.1f accum %.1f ph1 %.1f ph2 %.1f total %.1f gap %.1f" % (
    ((d[:, 1] - d[:, 0]).mean() / 1e3), ((d[:, 4] - d[:, 0]).mean() / 1e3), ((d[:, 5] - d[:, 4]).mean() / 1e3),
    ((d[:, 6] - d[:, 5]).mean() / 1e3), ((d[:, 7] - d[:, 0]).mean() / 1e3), ((d[1:n, 0] - d[:n - 1, 7]).mean() / 1e3)))
