"""CPU: the HiFT oracle restatement against golden vectors produced by the UNMODIFIED reference
(`oracle/make_golden.py hift`), plus known-answer checks (STFT/iSTFT round trip, Snake, weight-norm fold)."""
import math
import os

import torch
from safetensors.torch import load_file

from chatterbox_b200 import synthetic as S
from oracle import hift_oracle as HO

GOLD = load_file(os.path.join(os.path.dirname(__file__), "golden", "hift_golden.safetensors"))


def _w():
    return HO.fold_weight_norm(S.make_state_dict("hift"))


def test_weight_norm_fold_matches_torch_parametrisation():
    import torch.nn as nn
    from torch.nn.utils.parametrizations import weight_norm
    sd = S.make_state_dict("hift")
    c = weight_norm(nn.ConvTranspose1d(512, 256, 16, 8, padding=4))
    c.load_state_dict({k[len("ups.0."):]: v for k, v in sd.items() if k.startswith("ups.0.")})
    assert torch.allclose(c.weight, HO.fold_weight_norm(sd)["ups.0.weight"], atol=1e-7)


def test_f0_and_source_match_reference_golden():
    w = _w()
    for L in (12, 40):
        mel = S.mel_input(L, seed=7 + L)
        with torch.inference_mode():
            f0 = HO.f0_predictor(w, mel)
        assert (f0 - GOLD[f"f0_L{L}"]).abs().max() < 1e-2
    with torch.inference_mode():
        s = HO.sine_source(w, GOLD["f0_L12"], GOLD["phase_L12"], GOLD["noise_L12"])
    assert (s - GOLD["s_L12"]).abs().max() < 1e-5


def test_decode_matches_reference_golden():
    w = _w()
    for L in (12, 40):
        with torch.inference_mode():
            wav = HO.decode(w, S.mel_input(L, seed=7 + L), GOLD[f"s_L{L}"])
        assert wav.shape == (1, 480 * L)
        assert (wav - GOLD[f"wav_L{L}"]).norm() / GOLD[f"wav_L{L}"].norm() < 1e-5


def test_stft_istft_round_trip_and_snake():
    x = torch.randn(1, 960)
    spec = HO.stft(x)
    assert spec.shape == (1, 18, 241)
    re, im = spec[:, :9], spec[:, 9:]
    mag, ph = torch.sqrt(re * re + im * im), torch.atan2(im, re)
    y = HO.istft(mag, ph)
    assert (y - x).abs().max() < 1e-4
    a = torch.tensor([0.5, 2.0])
    v = torch.tensor([[[0.3], [-1.2]]])
    exp = torch.tensor([[[0.3 + math.sin(0.15) ** 2 / 0.5], [-1.2 + math.sin(-2.4) ** 2 / 2.0]]])
    assert torch.allclose(HO.snake(v, a), exp, atol=1e-6)


def test_trim_fade_shape():
    tf = HO.trim_fade()
    assert tf.shape == (960,) and float(tf[:480].abs().max()) == 0.0 and abs(float(tf[-1]) - 1.0) < 1e-6
