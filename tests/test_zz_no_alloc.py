"""The allocation-free entry point `cbx_hift_source_static` (include/cbx.h section 3b; no counterpart in the reference)
against the existing `cbx_hift_source`, through the C-ABI, on two calls in
a row.  Its scratch is a static __device__ array that keeps the previous call's contents, so the second call uses a
different f0 and a SHORTER length on purpose.  Collected last (file name) so a failure here hides no other test."""
import ctypes as C

import pytest
import torch

from chatterbox_b200 import synthetic as S

TOL = 2e-2      # the relative-L2 tolerance of the HiFT tests (tests/test_hift_gpu.py); the two paths share their kernels,
                # so the expected difference is exactly 0


def test_symbol_is_exported_and_limit_is_stated():
    from chatterbox_b200 import _lib
    assert _lib.lib().cbx_hift_source_static_max_frames() == 4096


@pytest.mark.gpu
def test_static_scratch_source_matches_the_handle_path_twice_in_a_row():
    from chatterbox_b200 import _lib
    from chatterbox_b200.hift import HiFTGenerator
    lib = _lib.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    sd = S.make_state_dict("hift")
    h = HiFTGenerator(sd, dev, max_frames=256)
    lw = sd["m_source.l_linear.weight"].reshape(-1).to(dev, torch.float32).contiguous()
    lb = sd["m_source.l_linear.bias"].reshape(-1).to(dev, torch.float32).contiguous()
    assert lw.numel() == 9 and lb.numel() == 1
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    for call, (L, seed) in enumerate([(200, 3), (37, 4)]):
        g = torch.Generator().manual_seed(seed)
        f0 = (torch.rand(1, L, generator=g) * 300.0)            # voiced and (below 10 Hz) unvoiced frames
        f0[0, ::7] = 3.0
        phase = torch.zeros(1, 9, 1)
        phase[:, 1:] = (torch.rand(1, 8, 1, generator=g) * 2 - 1) * torch.pi
        noise = torch.randn(1, 9, 480 * L, generator=g)
        for ph, nz in ((phase, noise), (None, None)):           # injected randomness, then the Philox path
            want = h.source(f0, ph, nz, seed=seed)
            f0d = f0.to(dev).contiguous()
            phd = ph.to(dev).reshape(-1).contiguous() if ph is not None else None
            nzd = nz.to(dev).reshape(9, -1).contiguous() if nz is not None else None
            got = torch.full((1, 1, 480 * L), float("nan"), dtype=torch.float32, device=dev)
            _lib.check(lib.cbx_hift_source_static(dev.index, lw.data_ptr(), lb.data_ptr(), f0d.data_ptr(), L,
                                                  phd.data_ptr() if phd is not None else None,
                                                  nzd.data_ptr() if nzd is not None else None, seed, got.data_ptr(), stream))
            torch.cuda.synchronize()
            assert torch.isfinite(got).all()
            rel = float((got - want).norm() / want.norm())
            print(f"call {call} (L = {L}, injected = {ph is not None}): rel-L2 vs cbx_hift_source {rel:.3e}, "
                  f"max-abs {float((got - want).abs().max()):.3e}")
            assert rel <= TOL and float((got - want).abs().max()) <= 1e-6
    # sizes beyond the static scratch are rejected with a clear error, nothing is launched
    big = lib.cbx_hift_source_static_max_frames() + 1
    f0d = torch.zeros(big, device=dev)
    out = torch.zeros(480 * big, device=dev)
    rc = lib.cbx_hift_source_static(dev.index, lw.data_ptr(), lb.data_ptr(), f0d.data_ptr(), big, None, None, 0, out.data_ptr(), stream)
    assert rc != 0 and b"exceeds the static scratch" in lib.cbx_last_error()
