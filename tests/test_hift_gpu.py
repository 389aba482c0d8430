"""GPU parity of the HiFT vocoder (through the C-ABI) vs the reference's own outputs
(tests/golden/hift_golden.safetensors) and vs the CPU oracle.

Tolerance (north_star / SURVEY 8c): relative L2 <= 2e-2 on the waveform for the bf16 tensor-core path, with
the stochastic source signal injected (`HiFTGenerator.decode(x, s)`, hifigan.py:412)."""
import os

import pytest
import torch
from safetensors.torch import load_file

from chatterbox_b200 import synthetic as S

pytestmark = pytest.mark.gpu
GOLD = load_file(os.path.join(os.path.dirname(__file__), "golden", "hift_golden.safetensors"))
TOL = 2e-2


def _rel(a, b):
    return float((a - b).norm() / b.norm())


@pytest.fixture(scope="module")
def hift():
    from chatterbox_b200.hift import HiFTGenerator
    return HiFTGenerator(S.make_state_dict("hift"), "cuda", max_frames=256)


@pytest.mark.parametrize("L", [12, 40])
def test_f0_matches_reference_golden(hift, L):
    f0 = hift.f0_predictor(S.mel_input(L, seed=7 + L)).cpu()
    r = _rel(f0, GOLD[f"f0_L{L}"])
    print(f"f0 L={L} rel-L2 vs reference {r:.3e}")
    assert r < TOL


def test_source_matches_reference_with_injected_randomness(hift):
    s = hift.source(GOLD["f0_L12"], GOLD["phase_L12"], GOLD["noise_L12"]).cpu()
    r = _rel(s, GOLD["s_L12"])
    print(f"source L=12 rel-L2 vs reference {r:.3e}")
    assert r < 1e-2


def test_source_philox_statistics(hift):
    """Without injection: unvoiced frames carry N(0, (0.1/3)^2)-driven noise, voiced frames a bounded harmonic sum."""
    f0 = torch.zeros(1, 20)
    f0[:, 10:] = 120.0
    s = hift.source(f0, seed=5).cpu()[0, 0]
    assert s.abs().max() <= 1.0 and s[:4800].std() > 1e-3 and s[4800:].std() > s[:4800].std()
    s2 = hift.source(f0, seed=5).cpu()[0, 0]
    s3 = hift.source(f0, seed=6).cpu()[0, 0]
    assert torch.equal(s, s2) and not torch.equal(s, s3)


@pytest.mark.parametrize("L", [12, 40])
def test_decode_matches_reference_golden(hift, L):
    wav = hift.decode(S.mel_input(L, seed=7 + L), GOLD[f"s_L{L}"]).cpu()
    assert wav.shape == (1, 480 * L)
    r = _rel(wav, GOLD[f"wav_L{L}"])
    print(f"decode L={L} rel-L2 vs reference {r:.3e}")
    assert r < TOL


@pytest.mark.parametrize("L", [1, 2, 3, 17, 129])
def test_decode_ragged_lengths_vs_oracle(hift, L):
    from oracle import hift_oracle as HO
    w = HO.fold_weight_norm(S.make_state_dict("hift"))
    mel = S.mel_input(L, seed=50 + L)
    s = torch.tanh(torch.randn(1, 1, 480 * L, generator=torch.Generator().manual_seed(L)) * 0.3)
    fade = L >= 2      # the reference's trim_fade (s3gen.py:296) itself needs >= 960 samples
    with torch.inference_mode():
        ref, _ = HO.inference(w, mel, s=s, apply_trim_fade=fade)
    wav = hift.decode(mel, s, apply_fade=fade).cpu()
    r = _rel(wav, ref)
    print(f"decode+fade L={L} rel-L2 vs oracle {r:.3e}")
    assert r < TOL
    if fade:
        assert float(wav[0, :480].abs().max()) == 0.0      # trim_fade: first 20 ms are silent


def test_inference_end_to_end_uses_own_source(hift):
    mel = S.mel_input(16, seed=3)
    wav, src = hift.inference(mel, seed=11)
    assert wav.shape == (1, 7680) and src.shape == (1, 1, 7680)
    wav2 = hift.decode(mel, src)
    assert torch.equal(wav, wav2)
    assert float(wav.abs().max()) <= 0.99 + 1e-6
    from chatterbox_b200._lib import CbxError
    with pytest.raises(CbxError):
        hift.f0_predictor(torch.zeros(1, 80, 1000))


def test_decode_at_baseline_size_vs_oracle():
    """Config 2 size: 512 mel frames (10.24 s of audio) against the CPU oracle with the same source signal."""
    from chatterbox_b200.hift import HiFTGenerator
    from oracle import hift_oracle as HO
    L = 512
    sd = S.make_state_dict("hift")
    big = HiFTGenerator(sd, "cuda", max_frames=L)
    w = HO.fold_weight_norm(sd)
    mel = S.mel_input(L, seed=77)
    s = torch.tanh(torch.randn(1, 1, 480 * L, generator=torch.Generator().manual_seed(L)) * 0.3)
    with torch.inference_mode():
        ref, _ = HO.inference(w, mel, s=s, apply_trim_fade=True)
    wav = big.decode(mel, s, apply_fade=True).cpu()
    r = _rel(wav, ref)
    print(f"decode+fade L={L} rel-L2 vs oracle {r:.3e}")
    assert r < TOL
