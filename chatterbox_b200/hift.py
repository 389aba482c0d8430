"""Host-side mirror of the reference's HiFT-GAN vocoder (`HiFTGenerator`, hifigan.py:291-474) over the C-ABI.

Same state-dict keys as the reference (weight-norm parametrisation `…parametrizations.weight.original0/1`
or already-folded `…weight`), same `inference(speech_feat, cache_source)` / `decode(x, s)` signatures; the
convolutions, Snake ResBlocks, STFT / iSTFT and the sine source run in `libcbx.so` (csrc/s3_hift.cu).
Weight norm is folded once at load (the reference renormalises every conv weight on every call).
There is no CPU path: construction fails without CUDA + libcbx.so.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import torch

from . import _lib

UP_RATES, UP_KERNELS, UP_TAPS = (8, 5, 3), (16, 11, 7), (2, 3, 3)


def fold_weight_norm(sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """g * v / ||v|| with the norm over every dim but 0 (torch weight_norm, dim=0)."""
    out = {}
    for k, v in sd.items():
        if k.endswith("parametrizations.weight.original1"):
            base = k[: -len("parametrizations.weight.original1")]
            g = sd[base + "parametrizations.weight.original0"].float()
            v = v.float()
            n = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))
            out[base + "weight"] = v * (g / n)
        elif not k.endswith("parametrizations.weight.original0"):
            out[k] = v
    return out


def _conv_as_gemm(w: torch.Tensor, cin_pad: Optional[int] = None, n_pad: Optional[int] = None) -> torch.Tensor:
    """Conv1d weight (Cout, Cin, k) -> [Cout][k*Cin'] (column = tap*Cin' + ci), optionally zero-padded."""
    co, ci, k = w.shape
    if cin_pad and cin_pad > ci:
        w = torch.cat([w, torch.zeros(co, cin_pad - ci, k)], 1)
    m = w.permute(0, 2, 1).reshape(co, -1)
    if n_pad and n_pad > co:
        m = torch.cat([m, torch.zeros(n_pad - co, m.shape[1])], 0)
    return m


def _convT_as_gemm(w: torch.Tensor, stride: int, taps: int) -> torch.Tensor:
    """ConvTranspose1d weight (Cin, Cout, k) -> polyphase matrix [(r*Cout + co)][(m*Cin + ci)] = w[ci][co][r + stride*m]."""
    ci, co, k = w.shape
    out = torch.zeros(stride, co, taps, ci)
    for r in range(stride):
        for m in range(taps):
            j = r + stride * m
            if j < k:
                out[r, :, m, :] = w[:, :, j].t()
    return out.reshape(stride * co, taps * ci)


def _rb_prefixes() -> List[str]:
    return [f"source_resblocks.{i}." for i in range(3)] + [f"resblocks.{i}." for i in range(9)]


def pack_hift(sd: Dict[str, torch.Tensor], device):
    w = {k: v.detach().float().cpu() for k, v in fold_weight_norm(sd).items()}
    bf: List[torch.Tensor] = []
    f: List[torch.Tensor] = []
    bf.append(_conv_as_gemm(w["f0_predictor.condnet.0.weight"], cin_pad=128))
    bf += [_conv_as_gemm(w[f"f0_predictor.condnet.{j}.weight"]) for j in (2, 4, 6, 8)]
    bf.append(_conv_as_gemm(w["conv_pre.weight"], cin_pad=128))
    bf += [_convT_as_gemm(w[f"ups.{i}.weight"], UP_RATES[i], UP_TAPS[i]) for i in range(3)]
    for p in _rb_prefixes():
        bf += [_conv_as_gemm(w[p + f"convs1.{j}.weight"]) for j in range(3)]
        bf += [_conv_as_gemm(w[p + f"convs2.{j}.weight"]) for j in range(3)]
    bf.append(_conv_as_gemm(w["conv_post.weight"], n_pad=32))
    f += [w[f"f0_predictor.condnet.{j}.bias"] for j in (0, 2, 4, 6, 8)]
    f += [w["f0_predictor.classifier.weight"].reshape(-1), w["f0_predictor.classifier.bias"], w["conv_pre.bias"]]
    f += [w[f"ups.{i}.bias"] for i in range(3)]
    f += [w[f"source_downs.{i}.weight"].permute(1, 2, 0).reshape(-1, w[f"source_downs.{i}.weight"].shape[0]) for i in range(3)]
    f += [w[f"source_downs.{i}.bias"] for i in range(3)]
    for p in _rb_prefixes():
        f += [w[p + f"convs1.{j}.bias"] for j in range(3)] + [w[p + f"convs2.{j}.bias"] for j in range(3)]
        f += [w[p + f"activations1.{j}.alpha"] for j in range(3)] + [w[p + f"activations2.{j}.alpha"] for j in range(3)]
    f += [torch.cat([w["conv_post.bias"], torch.zeros(32 - 18)]), w["m_source.l_linear.weight"].reshape(-1),
          w["m_source.l_linear.bias"]]
    assert len(bf) == 82 and len(f) == 164, (len(bf), len(f))

    def blob(ts, dtype, align):
        offs, n = [], 0
        for t in ts:
            offs.append(n)
            n += (t.numel() + align - 1) // align * align
        out = torch.zeros(n, dtype=dtype, device=device)
        for o, t in zip(offs, ts):
            out[o:o + t.numel()] = t.reshape(-1).to(device=device, dtype=dtype)
        return out, offs

    bfb, bfo = blob(bf, torch.bfloat16, 64)
    fb, fo = blob(f, torch.float32, 4)
    return bfb, fb, bfo, fo


class HiFTGenerator:
    def __init__(self, state_dict: Dict[str, torch.Tensor], device="cuda", *, max_frames: int = 2048):
        if not torch.cuda.is_available():
            raise _lib.CbxError("chatterbox_b200.hift needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self._lib = _lib.lib()
        self.bf, self.f32, bfo, fo = pack_hift(state_dict, self.device)
        self._bfo = (C.c_int64 * len(bfo))(*bfo)
        self._fo = (C.c_int64 * len(fo))(*fo)
        w = _lib.CbxHiftWeights(self.bf.data_ptr(), self.f32.data_ptr(), self._bfo, self._fo)
        h = C.c_void_p()
        _lib.check(self._lib.cbx_hift_create(self.device.index, int(max_frames), C.byref(w), C.byref(h)))
        self._h = h
        self.max_frames = max_frames
        self.sampling_rate = 24000

    def sibling(self) -> "HiFTGenerator":
        """A second handle (own workspace, own launch counter) over the SAME packed weights: lets several utterances'
        vocoder passes run on different streams at once (S3Gen.inference_batch)."""
        o = object.__new__(HiFTGenerator)
        o.device, o._lib, o.bf, o.f32, o._bfo, o._fo = self.device, self._lib, self.bf, self.f32, self._bfo, self._fo
        w = _lib.CbxHiftWeights(o.bf.data_ptr(), o.f32.data_ptr(), o._bfo, o._fo)
        h = C.c_void_p()
        _lib.check(o._lib.cbx_hift_create(o.device.index, int(self.max_frames), C.byref(w), C.byref(h)))
        o._h, o.max_frames, o.sampling_rate = h, self.max_frames, self.sampling_rate
        return o

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.cbx_hift_destroy(h)
            self._h = None

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _f(self, a):
        return a.to(self.device, torch.float32).contiguous()

    def launches(self) -> int:
        n = C.c_int64()
        _lib.check(self._lib.cbx_hift_launches(self._h, C.byref(n)))
        return n.value

    @torch.inference_mode()
    def f0_predictor(self, mel: torch.Tensor) -> torch.Tensor:
        """(1, 80, L) -> (1, L)   (f0_predictor.py:52-55)"""
        mel = self._f(mel)
        L = mel.shape[2]
        out = torch.empty(1, L, dtype=torch.float32, device=self.device)
        _lib.check(self._lib.cbx_hift_f0(self._h, mel.data_ptr(), L, out.data_ptr(), self._stream()))
        return out

    @torch.inference_mode()
    def source(self, f0: torch.Tensor, phase_vec: Optional[torch.Tensor] = None, noise: Optional[torch.Tensor] = None,
               seed: int = 0) -> torch.Tensor:
        """f0 (1, L) -> s (1, 1, 480 L): f0_upsamp + SourceModuleHnNSF (hifigan.py:465-469).  `phase_vec` (1,9,1) and
        `noise` (1,9,480 L) replace the reference's global-RNG draws when given."""
        f0 = self._f(f0)
        L = f0.shape[1]
        ph = self._f(phase_vec).reshape(-1) if phase_vec is not None else None
        nz = self._f(noise).reshape(9, -1) if noise is not None else None
        out = torch.empty(1, 1, 480 * L, dtype=torch.float32, device=self.device)
        _lib.check(self._lib.cbx_hift_source(self._h, f0.data_ptr(), L, ph.data_ptr() if ph is not None else None,
                                             nz.data_ptr() if nz is not None else None, int(seed) & (2 ** 64 - 1),
                                             out.data_ptr(), self._stream()))
        return out

    @torch.inference_mode()
    def decode(self, x: torch.Tensor, s: torch.Tensor, apply_fade: bool = False) -> torch.Tensor:
        """x mel (1, 80, L), s (1, 1, 480 L) -> wav (1, 480 L)   (hifigan.py:412-444)"""
        x, s = self._f(x), self._f(s)
        L = x.shape[2]
        assert s.numel() == 480 * L
        out = torch.empty(1, 480 * L, dtype=torch.float32, device=self.device)
        _lib.check(self._lib.cbx_hift_decode(self._h, x.data_ptr(), L, s.data_ptr(), 1 if apply_fade else 0, out.data_ptr(),
                                             self._stream()))
        return out

    @torch.inference_mode()
    def inference(self, speech_feat: torch.Tensor, cache_source: Optional[torch.Tensor] = None, *, seed: Optional[int] = None,
                  apply_fade: bool = False):
        """hifigan.py:462-474: returns (wav (1, 480 L), source (1, 1, 480 L)).  A non-empty `cache_source` covering the
        whole utterance is used as the source; partial caches (streaming) are not supported."""
        mel = self._f(speech_feat)
        L = mel.shape[2]
        src = None
        if cache_source is not None and cache_source.shape[-1] != 0:
            assert cache_source.shape[-1] == 480 * L, "partial cache_source (streaming) is not supported"
            src = self._f(cache_source)
        wav = torch.empty(1, 480 * L, dtype=torch.float32, device=self.device)
        s_out = torch.empty(1, 1, 480 * L, dtype=torch.float32, device=self.device)
        if seed is None:
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        _lib.check(self._lib.cbx_hift_inference(self._h, mel.data_ptr(), L, src.data_ptr() if src is not None else None,
                                                int(seed) & (2 ** 64 - 1), 1 if apply_fade else 0, wav.data_ptr(),
                                                s_out.data_ptr(), self._stream()))
        return wav, s_out
