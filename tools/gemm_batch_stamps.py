"""%globaltimer stamps of CTA (0, 0, 0) of every tile-per-CTA GEMM launch of one 64-utterance estimator call (eager): where a
128 x 256 tile's time goes at 896 row tiles.  python tools/gemm_batch_stamps.py [nb]"""
import ctypes as C, os, sys
os.environ["CBX_CFM_GRAPH"] = "0"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import synthetic as S, _lib
from chatterbox_b200.cfm import CausalConditionalCFM, ConditionalDecoder
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 64
L = 812
est = ConditionalDecoder(S.make_state_dict("flow"), "cuda", prefix="decoder.estimator.", max_len=L, max_batch=nb)
cfm = CausalConditionalCFM(est, rand_noise=S.cfm_noise())
i = S.cfm_inputs(L)
mu, cond, spks = (torch.cat([i[k]] * nb, 0).cuda() for k in ("mu", "cond", "spks"))
cfm.forward_batch(mu, None, 1, spks=spks, cond=cond)
torch.cuda.synchronize()
buf = torch.zeros(16 * 400, dtype=torch.int64, device="cuda")
lib = C.CDLL(_lib.LIB_PATH)
lib.cbx_dbg_gemm_stamps(C.c_void_p(buf.data_ptr()))
cfm.forward_batch(mu, None, 1, spks=spks, cond=cond)
lib.cbx_dbg_gemm_stamps(None)
torch.cuda.synchronize()
b = buf.cpu().view(-1, 16).double() / 1e3
n = int((b[:, 0] > 0).sum())
print(f"{n} stamped GEMM launches (CTA 0 = first wave; us from kernel start)")
for k in range(20, 34):
    r = b[k]
    print(k, f"init {r[1]-r[0]:.1f} | pdl_wait done {r[11]-r[0]:.1f} | loads issued {r[2]-r[0]:.1f} | mma issued {r[3]-r[0]:.1f} | accum ready {r[4]-r[0]:.1f} | "
             f"drained {r[5]-r[0]:.1f} | rows: first loaded {r[9]-r[0]:.1f} first done {r[10]-r[0]:.1f} | epilogue end {r[6]-r[0]:.1f} | exit {r[7]-r[0]:.1f}")
