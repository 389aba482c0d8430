"""Host-side mirror of the reference `S3Token2Wav` / `S3Gen` (src/chatterbox/models/s3gen/s3gen.py:212-298):
speech tokens + reference-voice dict -> mel (flow-matching) -> waveform (HiFT), same method names and
return values.  The token front-end (`flow_frontend.py` -> csrc/s3_front.cu), the CFM solve and the vocoder all run in
`libcbx.so`.

Not provided (conditioning preparation, out of the hot path — SURVEY.md §2 rows 9, 12 and §8f row 4):
`embed_ref` (needs the absent third-party `s3tokenizer` network, CAM++ and librosa mel) — callers pass a
pre-computed `ref_dict` exactly as the reference's production path does (`tts.py:266-269`).
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch

from . import _lib, flow_frontend
from .cfm import CausalConditionalCFM, ConditionalDecoder
from .hift import HiFTGenerator

S3GEN_SR = 24000


class S3Gen:
    def __init__(self, flow_sd: Dict[str, torch.Tensor], hift_sd: Dict[str, torch.Tensor], device="cuda", *,
                 rand_noise: Optional[torch.Tensor] = None, max_mel_len: int = 4096, max_batch: int = 1,
                 vocoder_streams: int = 4):
        if not torch.cuda.is_available():
            raise _lib.CbxError("chatterbox_b200.S3Gen needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.front = flow_frontend.FlowFrontend(flow_sd, self.device, max_tokens=max(2, max_mel_len // 2))
        self.estimator = ConditionalDecoder(flow_sd, self.device, prefix="decoder.estimator.", max_len=max_mel_len,
                                            max_batch=max_batch)
        self.decoder = CausalConditionalCFM(self.estimator, rand_noise=rand_noise)
        self.mel2wav = HiFTGenerator(hift_sd, self.device, max_frames=max_mel_len)
        # inference_batch: the vocoder passes of a batch are independent and each one is a chain of small launches that
        # cannot fill the GPU, so they run vocoder_streams at a time, each on its own stream and handle (created on first use)
        self.vocoder_streams = max(1, int(vocoder_streams))
        self._voc_pool: list = []
        self._voc_streams: list = []
        self._front_pool: list = []        # the same for the token front-ends of a batch (flow_inference_batch)
        self.record_events = False
        self.events: Dict[str, torch.cuda.Event] = {}

    def _front(self, tok, prompt_token, prompt_feat, embedding):
        """flow.py:250-270 on the kernels of csrc/s3_front.cu (~100 launches on the current stream, no host sync)."""
        return self.front.flow_inputs(tok, prompt_token, prompt_feat, embedding)

    @classmethod
    def from_state_dict(cls, sd: Dict[str, torch.Tensor], device="cuda", **kw) -> "S3Gen":
        """`s3gen.safetensors` layout: `flow.*`, `mel2wav.*` (+ tokenizer / speaker_encoder entries, unused here)."""
        flow = {k[len("flow."):]: v for k, v in sd.items() if k.startswith("flow.")}
        hift = {k[len("mel2wav."):]: v for k, v in sd.items() if k.startswith("mel2wav.")}
        return cls(flow, hift, device, **kw)

    def embed_ref(self, *a, **k):
        raise NotImplementedError("embed_ref (reference-voice conditioning prep) is outside the accelerated path; "
                                  "pass a pre-computed ref_dict (Conditionals.gen)")

    def _ref(self, ref_dict):
        out = {}
        for k, v in ref_dict.items():
            if isinstance(v, np.ndarray):
                v = torch.from_numpy(v)
            out[k] = v.to(self.device) if torch.is_tensor(v) else v
        return out

    @torch.inference_mode()
    def flow_inference(self, speech_tokens, ref_wav=None, ref_sr=None, ref_dict: Optional[dict] = None, finalize: bool = True):
        assert ref_wav is None and ref_dict is not None, "pass a pre-computed ref_dict (embed_ref is not provided)"
        assert finalize, "streaming (finalize=False) is not supported"
        ref = self._ref(ref_dict)
        tok = speech_tokens.to(self.device)
        if tok.dim() == 1:
            tok = tok.unsqueeze(0)
        assert tok.shape[0] == 1
        ev = {k: torch.cuda.Event(enable_timing=True) for k in ("front0", "cfm0", "cfm1")} if self.record_events else None
        if ev:
            ev["front0"].record()
        mu, spks, cond, mel_len1 = self._front(tok.long(), ref["prompt_token"].long(), ref["prompt_feat"].float(),
                                               ref["embedding"].float())
        if ev:
            ev["cfm0"].record()
        L = mu.shape[2]
        feat, _ = self.decoder(mu=mu, mask=None, n_timesteps=10, spks=spks, cond=cond)     # None = all ones, no host sync
        if ev:
            ev["cfm1"].record()
            self.events.update(ev)
        return feat[:, :, mel_len1:].contiguous()

    @torch.inference_mode()
    def hift_inference(self, speech_feat, cache_source: Optional[torch.Tensor] = None, *, seed: Optional[int] = None,
                       apply_fade: bool = False):
        return self.mel2wav.inference(speech_feat, cache_source, seed=seed, apply_fade=apply_fade)

    @torch.inference_mode()
    def inference(self, speech_tokens, ref_wav=None, ref_sr=None, ref_dict: Optional[dict] = None,
                  cache_source: Optional[torch.Tensor] = None, finalize: bool = True, *, seed: Optional[int] = None):
        """-> (wav (1, n), source (1, 1, n)); trim_fade (s3gen.py:296) is applied inside the vocoder's last kernel."""
        mels = self.flow_inference(speech_tokens, ref_wav=ref_wav, ref_sr=ref_sr, ref_dict=ref_dict, finalize=finalize)
        ev = {k: torch.cuda.Event(enable_timing=True) for k in ("hift0", "hift1")} if self.record_events else None
        if ev:
            ev["hift0"].record()
        assert mels.shape[2] >= 2, "the reference's trim_fade needs at least 2 mel frames"
        wav, src = self.hift_inference(mels, cache_source, seed=seed, apply_fade=True)
        if ev:
            ev["hift1"].record()
            self.events.update(ev)
        return wav, src

    @torch.inference_mode()
    def flow_inference_batch(self, speech_tokens_list, ref_dicts):
        """Several utterances through ONE flow-matching solve (BASELINE config 4; the reference would call
        `flow.inference` once per request, server.py:223-232): token front-end per utterance, (mu, cond) padded to the
        longest mel length, `CausalConditionalCFM.forward_batch`, each utterance's own frames sliced out.
        -> list of mel (1, 80, 2 T_u)."""
        ev = {k: torch.cuda.Event(enable_timing=True) for k in ("front0", "cfm0", "cfm1")} if self.record_events else None
        if ev:
            ev["front0"].record()
        mus, spk, conds, l1s = [], [], [], []
        K = min(self.vocoder_streams, len(speech_tokens_list))
        while len(self._front_pool) < K:
            self._front_pool.append(self.front if not self._front_pool else self.front.sibling())
        while len(self._voc_streams) < K:
            self._voc_streams.append(torch.cuda.Stream(self.device))
        main = torch.cuda.current_stream(self.device)
        for i, (tok, rd) in enumerate(zip(speech_tokens_list, ref_dicts)):
            ref = self._ref(rd)
            tok = tok.to(self.device)
            if tok.dim() == 1:
                tok = tok.unsqueeze(0)
            if K <= 1:
                mu, spks, cond, l1 = self._front(tok.long(), ref["prompt_token"].long(), ref["prompt_feat"].float(),
                                                 ref["embedding"].float())
            else:              # ~100 small launches per utterance: K utterances at a time, each on its own stream / handle
                st = self._voc_streams[i % K]
                st.wait_stream(main)          # this utterance's tokens / reference tensors may just have been copied there
                with torch.cuda.stream(st):
                    mu, spks, cond, l1 = self._front_pool[i % K].flow_inputs(tok.long(), ref["prompt_token"].long(),
                                                                             ref["prompt_feat"].float(), ref["embedding"].float())
                for t in (mu, spks, cond):
                    t.record_stream(main)
            mus.append(mu); spk.append(spks); conds.append(cond); l1s.append(l1)
        if K > 1:
            for st in self._voc_streams[:K]:
                main.wait_stream(st)
        if ev:
            ev["cfm0"].record()
        lens = [int(m.shape[2]) for m in mus]
        L = max(lens)
        pad = lambda a: torch.nn.functional.pad(a, (0, L - a.shape[2]))
        feat = self.decoder.forward_batch(torch.cat([pad(m) for m in mus], 0), lens, 10, spks=torch.cat(spk, 0),
                                          cond=torch.cat([pad(c) for c in conds], 0))
        if ev:
            ev["cfm1"].record()
            self.events.update(ev)
        return [feat[u:u + 1, :, l1s[u]:lens[u]].contiguous() for u in range(len(mus))]

    @torch.inference_mode()
    def inference_batch(self, speech_tokens_list, ref_dicts, *, seed: Optional[int] = None):
        """-> list of (wav (1, n_u), source (1, 1, n_u)): batched flow matching, then the vocoder per utterance."""
        mels = self.flow_inference_batch(speech_tokens_list, ref_dicts)
        ev = {k: torch.cuda.Event(enable_timing=True) for k in ("hift0", "hift1")} if self.record_events else None
        if ev:
            ev["hift0"].record()
        K = min(self.vocoder_streams, len(mels))
        if K <= 1:
            outs = [self.hift_inference(m, None, seed=seed, apply_fade=True) for m in mels]
        else:
            if seed is None:       # one draw on the host, as the sequential path would make per utterance
                seed = int(torch.randint(0, 2 ** 62, (1,)).item())
            while len(self._voc_pool) < K:
                self._voc_pool.append(self.mel2wav if not self._voc_pool else self.mel2wav.sibling())
            while len(self._voc_streams) < K:
                self._voc_streams.append(torch.cuda.Stream(self.device))
            main = torch.cuda.current_stream(self.device)
            ready = torch.cuda.Event()
            ready.record(main)                                   # the mels come from the solve on the caller's stream
            outs = [None] * len(mels)
            for i, m in enumerate(mels):
                st = self._voc_streams[i % K]
                if i < K:
                    st.wait_event(ready)
                with torch.cuda.stream(st):
                    outs[i] = self._voc_pool[i % K].inference(m, None, seed=seed, apply_fade=True)
                for t in outs[i]:
                    t.record_stream(main)                        # allocated on the side stream, consumed on the caller's
            for st in self._voc_streams[:K]:
                main.wait_stream(st)
        if ev:
            ev["hift1"].record()
            self.events.update(ev)
        return outs

    def launches(self) -> int:
        return (self.front.launches() + self.estimator.launches() + self.mel2wav.launches()
                + sum(v.launches() for v in self._voc_pool[1:]) + sum(v.launches() for v in self._front_pool[1:]))
