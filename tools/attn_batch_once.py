"""A few launches of the batched attention kernel alone (target for ncu): python tools/attn_batch_once.py [B] [L]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import _lib
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
L = int(sys.argv[2]) if len(sys.argv) > 2 else 812
lib = C.CDLL(_lib.LIB_PATH)
qkv = (torch.randn(B, L, 1536, generator=torch.Generator().manual_seed(1)) * 0.8).to(torch.bfloat16).cuda()
out = torch.empty(B, L, 512, dtype=torch.bfloat16, device="cuda")
for _ in range(3):
    assert lib.cbx_dbg_s3_attn(C.c_void_p(qkv.data_ptr()), C.c_void_p(out.data_ptr()), B, L, None, 1, None) == 0
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
