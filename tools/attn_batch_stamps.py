"""%globaltimer stamps of CTA (0, 0, 0) of the one-group attention kernel on a padded batch (B x 8 heads x q-tiles > 296 CTAs),
plus the launch's duration.  python tools/attn_batch_stamps.py [B] [L]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import _lib
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
L = int(sys.argv[2]) if len(sys.argv) > 2 else 812
lib = C.CDLL(_lib.LIB_PATH)
qkv = (torch.randn(B, L, 1536, generator=torch.Generator().manual_seed(1)) * 0.8).to(torch.bfloat16).cuda()
out = torch.empty(B, L, 512, dtype=torch.bfloat16, device="cuda")
run = lambda: lib.cbx_dbg_s3_attn(C.c_void_p(qkv.data_ptr()), C.c_void_p(out.data_ptr()), B, L, None, 1, None)
for _ in range(3): assert run() == 0
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): run()
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 100
fl = 4.0 * B * 8 * L * L * 64
print(f"attention B={B} L={L}: {us:.1f} us per launch, {fl / us / 1e6:.1f} TFLOP/s")
buf = torch.zeros(32 * 8, dtype=torch.int64, device="cuda")
lib.cbx_dbg_attn_stamps(C.c_void_p(buf.data_ptr()))
for _ in range(4): run()
lib.cbx_dbg_attn_stamps(None)
torch.cuda.synchronize()
b = buf.cpu().view(-1, 32).double()
names = {1: "issuer past pdl_wait", 2: "S: blk2 S ready", 4: "S: row in registers", 3: "S: max", 5: "S: rescale checked", 26: "S: chunk 0 exps done",
         27: "S: o_full(1) seen", 6: "S: exps issued", 7: "S: P stored", 8: "S: arrived p_full(2)", 10: "S: blk3 S ready", 14: "loop end", 15: "exit",
         20: "I: s_free(2) seen", 21: "I: S(3) issued", 22: "I: p_full(2) seen", 23: "I: V(2) ready", 24: "I: PV(2) issued", 25: "I: s_free(3) seen"}
print("SM cycles relative to kernel start (CTA 0,0,0; S = softmax thread 0, I = issuer)")
for r in b[1:4]:
    print(" | ".join(f"{nm} {int(r[k] - r[0])}" for k, nm in sorted(names.items(), key=lambda kv: float(r[kv[0]])) if r[k] > 0))
