"""(The tcgen05 attention has two forms chosen by the CTA count: two softmax groups per CTA for one utterance (B = 2 cases), and
for padded batches — B x 8 heads x q-tiles > 296 — the persistent two-query-tile kernel, whose CTAs each walk several work items
with different key counts in the ragged cases; the one-tile-per-CTA form behind CBX_ATTN_QT=0 is covered by
tests/test_cfm_switches_gpu.py.)

The estimator's attention kernels alone (csrc/s3_attn_tc.cu: tcgen05; csrc/s3_attn.cu: mma.sync) against a plain PyTorch
fp32 attention on the same bf16 q|k|v (reference: matcha/transformer.py:196-204 -> diffusers AttnProcessor2_0; 8 heads x 64,
scale 1/8, key-padding mask of utils/mask.py:167-193).  Tolerance: the kernels round the probabilities to bf16 before the
P V product and the output to bf16: 2^-8 relative on values of order 1 -> max-abs 2e-2, relative L2 6e-3 (measured ~3e-3)."""
import ctypes as C

import pytest
import torch

from chatterbox_b200 import _lib

pytestmark = pytest.mark.gpu


def _run(qkv, lens, tc):
    B, L, _ = qkv.shape
    lib = C.CDLL(_lib.LIB_PATH)
    lib.cbx_last_error.restype = C.c_char_p
    out = torch.full((B, L, 512), float("nan"), dtype=torch.bfloat16, device=qkv.device)
    lp = C.c_void_p(lens.data_ptr()) if lens is not None else None
    rc = lib.cbx_dbg_s3_attn(C.c_void_p(qkv.data_ptr()), C.c_void_p(out.data_ptr()), B, L, lp, tc, None)
    assert rc == 0, lib.cbx_last_error().decode()
    torch.cuda.synchronize()
    return out


def _ref(qkv, lens):
    B, L, _ = qkv.shape
    x = qkv.float().view(B, L, 3, 8, 64).permute(2, 0, 3, 1, 4)          # [3][B][8][L][64]
    s = x[0] @ x[1].transpose(-1, -2) * 0.125
    if lens is not None:
        key = torch.arange(L, device=qkv.device)[None, None, None, :] >= lens.clamp(1, L)[:, None, None, None]
        s = s.masked_fill(key, float("-inf"))
    return (torch.softmax(s, -1) @ x[2]).permute(0, 2, 1, 3).reshape(B, L, 512)


@pytest.mark.parametrize("tc", [1, 0])
@pytest.mark.parametrize("B,L,lens", [(2, 44, None), (2, 128, None), (2, 129, None), (2, 812, None), (2, 1800, None),
                                      (4, 300, [300, 300, 17, 17]), (6, 520, [520, 520, 257, 257, 128, 128]),
                                      (16, 300, [300, 300, 17, 17, 200, 200, 129, 129, 128, 128, 64, 64, 300, 300, 1, 1]),
                                      (40, 257, None),          # 640 work items on 148 persistent CTAs, odd tile count
                                      (24, 520, [520, 520, 1, 1, 130, 130, 257, 257, 400, 400, 128, 128] * 2)])
def test_attention_matches_fp32_torch(tc, B, L, lens):
    g = torch.Generator().manual_seed(L * 7 + B)
    qkv = (torch.randn(B, L, 1536, generator=g) * 1.5).to(torch.bfloat16).cuda()
    lens_t = torch.tensor(lens, dtype=torch.int32, device="cuda") if lens else None
    got = _run(qkv, lens_t, tc).float()
    want = _ref(qkv, lens_t)
    if lens:                      # rows past an utterance's length are padding: the reference zeroes them downstream
        for b, n in enumerate(lens):
            got[b, n:] = 0; want[b, n:] = 0
    assert torch.isfinite(got).all()
    rel = float((got - want).norm() / want.norm()); mx = float((got - want).abs().max())
    print(f"{'tcgen05' if tc else 'mma.sync'} attention B={B} L={L} lens={lens}: rel-L2 {rel:.2e} max-abs {mx:.2e}")
    assert rel < 6e-3 and mx < 2e-2


def test_late_key_blocks_far_above_the_running_maximum():
    """Online-softmax rescaling across key blocks: scores that jump by hundreds from one 128-key block to the next (and fall
    again) exercise the running-maximum correction of the block-late P V fold in both kernels."""
    g = torch.Generator().manual_seed(11)
    qkv = torch.randn(2, 700, 1536, generator=g)
    qkv[:, 300:, 512:1024] *= 40.0            # keys of blocks >= 2: scores ~ 40 x larger
    qkv[:, 600:, 512:1024] *= 0.01            # and small again in the last block
    qkv = qkv.to(torch.bfloat16).cuda()
    want = _ref(qkv, None)
    for tc in (1, 0):
        got = _run(qkv, None, tc).float()
        assert torch.isfinite(got).all()
        rel = float((got - want).norm() / want.norm())
        print(f"{'tcgen05' if tc else 'mma.sync'} attention, score jumps between key blocks: rel-L2 {rel:.2e}")
        assert rel < 6e-3


def test_tcgen05_and_mma_kernels_agree_and_repeat():
    g = torch.Generator().manual_seed(3)
    qkv = torch.randn(2, 812, 1536, generator=g).to(torch.bfloat16).cuda()
    a, b2, c = _run(qkv, None, 1), _run(qkv, None, 1), _run(qkv, None, 0)
    assert torch.equal(a, b2)
    assert (a.float() - c.float()).abs().max() < 2e-2
