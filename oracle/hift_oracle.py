"""CPU restatement of the reference's HiFT-GAN vocoder path (TEST INFRASTRUCTURE — checker only).

Plain torch fp32 on CPU, no dependence on `/root/reference` or scipy, so it travels to the GPU box.
Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may
import it; the product path (`chatterbox_b200/`) never does.

Parity pin: `oracle/make_golden.py hift` (build container) checks every function here against the
UNMODIFIED reference `HiFTGenerator` / `ConvRNNF0Predictor` with identical synthetic weights and commits
the reference's outputs under `tests/golden/`; `tests/test_oracle_hift.py` re-checks this restatement
against those files.  The reference ships no golden vectors for this path.

What each function follows (paths relative to /root/reference/src/chatterbox/models/s3gen):
  fold_weight_norm   torch weight_norm parametrisation (hifigan.py:28,120-142,331-348,375; f0_predictor.py:29-47)
  f0_predictor       f0_predictor.py:52-55 (5 x {conv k3 + ELU}, Linear(512->1), abs)
  sine_source        hifigan.py:200-231 (SineGen) + :267-283 (SourceModuleHnNSF) with the random phase /
                     noise tensors passed in explicitly (the reference draws them from the global RNG)
  stft / istft       hifigan.py:396-410 (torch.stft / torch.istft, n_fft 16, hop 4, periodic hann)
  resblock, snake    hifigan.py:106-166, 34-84
  decode             hifigan.py:412-444
  inference          hifigan.py:462-474 (+ s3gen.py:233-236,296 trim_fade)
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
import torch.nn.functional as F

UP_RATES, UP_KERNELS = (8, 5, 3), (16, 11, 7)
N_FFT, HOP = 16, 4
SR = 24000
NB_HARM = 8
SINE_AMP, NOISE_STD, VOICED_THR = 0.1, 0.003, 10.0


def fold_weight_norm(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """`…parametrizations.weight.original0/1` (g, v) -> plain `…weight` = v * g / ||v|| (norm over all dims but 0)."""
    out = {}
    for k, v in sd.items():
        if k.endswith("parametrizations.weight.original1"):
            base = k[: -len("parametrizations.weight.original1")]
            g = sd[base + "parametrizations.weight.original0"]
            n = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))
            out[base + "weight"] = v * (g / n)
        elif k.endswith("parametrizations.weight.original0"):
            continue
        else:
            out[k] = v
    return out


def hann_periodic(n: int = N_FFT) -> torch.Tensor:
    return 0.5 - 0.5 * torch.cos(2 * math.pi * torch.arange(n, dtype=torch.float32) / n)


def f0_predictor(w, mel: torch.Tensor) -> torch.Tensor:
    """mel (1, 80, L) -> f0 (1, L)."""
    x = mel
    for j in (0, 2, 4, 6, 8):
        x = F.elu(F.conv1d(x, w[f"f0_predictor.condnet.{j}.weight"], w[f"f0_predictor.condnet.{j}.bias"], padding=1))
    x = F.linear(x.transpose(1, 2), w["f0_predictor.classifier.weight"], w["f0_predictor.classifier.bias"])
    return torch.abs(x.squeeze(-1))


def sine_source(w, f0: torch.Tensor, phase_vec: torch.Tensor, noise: torch.Tensor) -> torch.Tensor:
    """f0 (1, L) per mel frame; phase_vec (1, 9, 1) (row 0 must be 0); noise (1, 9, 480 L) ~ N(0,1)
    -> s (1, 1, 480 L): nearest x480 upsample, 9 harmonics, tanh(Linear(9 -> 1))."""
    up = 480
    f0s = f0[:, None, :].repeat_interleave(up, dim=2)                        # (1,1,T)  nn.Upsample(nearest)
    fmat = torch.cat([f0s * (i + 1) / SR for i in range(NB_HARM + 1)], dim=1)  # (1,9,T)
    theta = 2 * math.pi * (torch.cumsum(fmat, dim=-1) % 1)
    sine = SINE_AMP * torch.sin(theta + phase_vec)
    uv = (f0s > VOICED_THR).float()
    noise_amp = uv * NOISE_STD + (1 - uv) * SINE_AMP / 3
    sine = sine * uv + noise_amp * noise
    merged = torch.tanh(F.linear(sine.transpose(1, 2), w["m_source.l_linear.weight"], w["m_source.l_linear.bias"]))
    return merged.transpose(1, 2)                                            # (1,1,T)


def stft(s: torch.Tensor) -> torch.Tensor:
    """s (1, T) -> (1, 18, T/4 + 1): real parts of the 9 bins then imaginary parts (center=True, reflect pad)."""
    spec = torch.stft(s, N_FFT, HOP, N_FFT, window=hann_periodic().to(s.device), return_complex=True)
    spec = torch.view_as_real(spec)
    return torch.cat([spec[..., 0], spec[..., 1]], dim=1)


def istft(mag: torch.Tensor, phase: torch.Tensor) -> torch.Tensor:
    mag = torch.clip(mag, max=1e2)
    spec = torch.complex(mag * torch.cos(phase), mag * torch.sin(phase))
    return torch.istft(spec, N_FFT, HOP, N_FFT, window=hann_periodic().to(spec.device))


def snake(x: torch.Tensor, alpha: torch.Tensor) -> torch.Tensor:
    a = alpha[None, :, None]
    return x + (1.0 / (a + 1e-9)) * torch.sin(x * a).pow(2)


def resblock(w, p: str, x: torch.Tensor, k: int) -> torch.Tensor:
    for idx, d in enumerate((1, 3, 5)):
        xt = snake(x, w[p + f"activations1.{idx}.alpha"])
        xt = F.conv1d(xt, w[p + f"convs1.{idx}.weight"], w[p + f"convs1.{idx}.bias"], dilation=d, padding=(k * d - d) // 2)
        xt = snake(xt, w[p + f"activations2.{idx}.alpha"])
        xt = F.conv1d(xt, w[p + f"convs2.{idx}.weight"], w[p + f"convs2.{idx}.bias"], padding=(k - 1) // 2)
        x = xt + x
    return x


def decode(w, mel: torch.Tensor, s: torch.Tensor, stop_after: Optional[str] = None) -> torch.Tensor:
    """mel (1, 80, L), s (1, 1, 480 L) -> wav (1, 480 L)."""
    s_stft = stft(s.squeeze(1))
    x = F.conv1d(mel, w["conv_pre.weight"], w["conv_pre.bias"], padding=3)
    src_k = (7, 7, 11)
    for i in range(3):
        x = F.leaky_relu(x, 0.1)
        u, k = UP_RATES[i], UP_KERNELS[i]
        x = F.conv_transpose1d(x, w[f"ups.{i}.weight"], w[f"ups.{i}.bias"], stride=u, padding=(k - u) // 2)
        if i == 2:
            x = F.pad(x, (1, 0), mode="reflect")
        stride = (15, 3, 1)[i]
        pad = (7, 1, 0)[i]
        si = F.conv1d(s_stft, w[f"source_downs.{i}.weight"], w[f"source_downs.{i}.bias"], stride=stride, padding=pad)
        si = resblock(w, f"source_resblocks.{i}.", si, src_k[i])
        x = x + si
        xs = None
        for j, k in enumerate((3, 7, 11)):
            r = resblock(w, f"resblocks.{i * 3 + j}.", x, k)
            xs = r if xs is None else xs + r
        x = xs / 3
        if stop_after == f"stage{i}":
            return x
    x = F.leaky_relu(x)
    x = F.conv1d(x, w["conv_post.weight"], w["conv_post.bias"], padding=3)
    mag = torch.exp(x[:, : N_FFT // 2 + 1])
    phase = torch.sin(x[:, N_FFT // 2 + 1:])
    return torch.clamp(istft(mag, phase), -0.99, 0.99)


def trim_fade() -> torch.Tensor:
    """s3gen.py:233-236: 20 ms of zeros then a 20 ms raised-cosine fade-in."""
    n = SR // 50
    tf = torch.zeros(2 * n)
    tf[n:] = (torch.cos(torch.linspace(torch.pi, 0, n)) + 1) / 2
    return tf


def inference(w, mel: torch.Tensor, s: Optional[torch.Tensor] = None, phase_vec=None, noise=None,
              apply_trim_fade: bool = True):
    """HiFTGenerator.inference with the source either injected (`s`) or built from explicit random tensors;
    then S3Token2Wav.inference's trim_fade (s3gen.py:296).  Returns (wav (1, 480 L), s)."""
    if s is None:
        s = sine_source(w, f0_predictor(w, mel), phase_vec, noise)
    wav = decode(w, mel, s)
    if apply_trim_fade:
        tf = trim_fade().to(wav.device)
        wav = wav.clone()
        wav[:, : len(tf)] *= tf
    return wav, s
