"""Per-phase %globaltimer stamps of the tcgen05 attention kernel (csrc/s3_attn_tc.cu) for every launch of one estimator call."""
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
buf = torch.zeros(32 * 400, dtype=torch.int64, device="cuda")
lib = C.CDLL(_lib.LIB_PATH)
lib.cbx_dbg_attn_stamps(C.c_void_p(buf.data_ptr()))
est(x, m, mu, t, sp, cd)
lib.cbx_dbg_attn_stamps(None)
torch.cuda.synchronize()
b = buf.cpu().view(-1, 32)[:, :16]
n = int((b[:, 0] > 0).sum())
names = ["pdl_wait (issuer)", "blk2 S ready", "blk2 max pass done", "blk2 o_full seen", "blk2 fold done", "blk2 exp pass done", "blk2 proxy fence done",
         "blk2 arrived", "-", "blk4 S ready", "blk6 S ready", "-", "-", "loop end", "exit"]
d = b[:n].double() / 1.9e3      # SM cycles -> us at the nominal 1.9 GHz
print(f"{n} attention launches; stamps of thread 0 (softmax group 0) relative to kernel start (us), mean over launches")
r = d[4:].mean(0)
print(" | ".join(f"{nm} {float(r[k + 1] - r[0]):.2f}" for k, nm in enumerate(names) if nm != "-"))
