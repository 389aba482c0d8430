"""Drop-in mirror of the reference public TTS API (`src/chatterbox/tts.py`): `ChatterboxTTS.from_pretrained /
from_local / generate(text, audio_prompt_path, exaggeration, cfg_weight, ...)`, `Conditionals`, `punc_norm` —
same names, arguments, defaults, side effects and return value (CPU float tensor (1, n) at 24 kHz).

T3 decode, the flow-matching solve and the vocoder run in `libcbx.so`.  Pieces outside the accelerated path
that depend on packages absent here are optional: the text tokenizer needs `tokenizer.json` (a checkpoint
asset); `prepare_conditionals` (voice encoder, S3 tokenizer, CAM++: once per voice) is not provided — pass
pre-computed `Conditionals` / `conds.pt`, the reference's own production path; the Perth watermark is
applied when `perth` is importable.  `from_state_dicts` builds the model from in-memory state dicts
(random-init weights offline); `generate(..., text_tokens=ids)` accepts pre-tokenised text.
"""
from __future__ import annotations

from dataclasses import dataclass
from pathlib import Path
from typing import Dict, Optional

import torch
import torch.nn.functional as F

from . import _lib
from .s3gen import S3GEN_SR, S3Gen
from .t3 import T3, T3Cond, T3Config

REPO_ID = "ResembleAI/chatterbox"
S3_SR = 16000
SPEECH_VOCAB_SIZE = 6561


def punc_norm(text: str) -> str:
    """reference tts.py:22-61"""
    if len(text) == 0:
        return "You need to add some text for me to talk."
    if text[0].islower():
        text = text[0].upper() + text[1:]
    text = " ".join(text.split())
    for old, new in [("...", ", "), ("…", ", "), (":", ","), (" - ", ", "), (";", ", "), ("—", "-"), ("–", "-"),
                     (" ,", ","), ("“", "\""), ("”", "\""), ("‘", "'"), ("’", "'")]:
        text = text.replace(old, new)
    text = text.rstrip(" ")
    if not any(text.endswith(p) for p in {".", "!", "?", "-", ","}):
        text += "."
    return text


def drop_invalid_tokens(x: torch.Tensor) -> torch.Tensor:
    """reference models/s3tokenizer/__init__.py:16-30: cut at SOS (6561) / EOS (6562)."""
    assert x.dim() == 1 or (x.dim() == 2 and x.shape[0] == 1), "only batch size of one allowed for now"
    x = x.reshape(-1)
    sos, eos = SPEECH_VOCAB_SIZE, SPEECH_VOCAB_SIZE + 1
    s = int((x == sos).nonzero()[0, 0]) + 1 if bool((x == sos).any()) else 0
    e = int((x == eos).nonzero()[0, 0]) if bool((x == eos).any()) else None
    return x[s:e]


@dataclass
class Conditionals:
    """reference tts.py:64-103"""
    t3: T3Cond
    gen: dict

    def to(self, device):
        self.t3 = self.t3.to(device=device)
        for k, v in self.gen.items():
            if torch.is_tensor(v):
                self.gen[k] = v.to(device=device)
        return self

    def save(self, fpath: Path):
        torch.save(dict(t3=self.t3.__dict__, gen=self.gen), fpath)

    @classmethod
    def load(cls, fpath, map_location="cpu"):
        if isinstance(map_location, str):
            map_location = torch.device(map_location)
        kwargs = torch.load(fpath, map_location=map_location, weights_only=True)
        return cls(T3Cond(**kwargs["t3"]), kwargs["gen"])


class EnTokenizer:
    """reference models/tokenizers/tokenizer.py:26-38 (HF `tokenizers` BPE file with [SPACE] substitution)."""

    def __init__(self, vocab_file_path):
        from tokenizers import Tokenizer
        self.tokenizer = Tokenizer.from_file(vocab_file_path)

    def text_to_tokens(self, text: str):
        return torch.IntTensor(self.encode(text)).unsqueeze(0)

    def encode(self, txt: str):
        return self.tokenizer.encode(txt.replace(" ", "[SPACE]")).ids


class ChatterboxTTS:
    ENC_COND_LEN = 6 * S3_SR
    DEC_COND_LEN = 10 * S3GEN_SR

    def __init__(self, t3: T3, s3gen: S3Gen, ve, tokenizer, device, conds: Optional[Conditionals] = None):
        self.sr = S3GEN_SR
        self.t3 = t3
        self.s3gen = s3gen
        self.ve = ve
        self.tokenizer = tokenizer
        self.device = device
        self.conds = conds
        try:
            import perth
            self.watermarker = perth.PerthImplicitWatermarker()
        except Exception:
            self.watermarker = None      # closed third-party model, applied after the accelerated path when present

    @classmethod
    def from_state_dicts(cls, t3_sd: Dict[str, torch.Tensor], flow_sd: Dict[str, torch.Tensor],
                         hift_sd: Dict[str, torch.Tensor], device="cuda", conds: Optional[Conditionals] = None,
                         tokenizer=None, *, rand_noise=None, max_speech_tokens: int = 1000,
                         max_text_tokens: int = T3Config.max_text_tokens,
                         max_prompt_tokens: int = 250) -> "ChatterboxTTS":
        max_pos = 34 + max_text_tokens + 4 + max_speech_tokens + 8
        t3 = T3(t3_sd, device, max_pos=max_pos)
        s3gen = S3Gen(flow_sd, hift_sd, device, rand_noise=rand_noise, max_mel_len=2 * (max_prompt_tokens + max_speech_tokens))
        if conds is not None:
            conds = conds.to(device)
        return cls(t3, s3gen, None, tokenizer, device, conds=conds)

    @classmethod
    def from_local(cls, ckpt_dir, device) -> "ChatterboxTTS":
        """reference tts.py:128-165: `t3_cfg.safetensors`, `s3gen.safetensors`, `tokenizer.json`, `conds.pt`."""
        from safetensors.torch import load_file
        ckpt_dir = Path(ckpt_dir)
        t3_state = load_file(ckpt_dir / "t3_cfg.safetensors")
        if "model" in t3_state.keys():
            t3_state = t3_state["model"][0]
        s3 = load_file(ckpt_dir / "s3gen.safetensors")
        tok = EnTokenizer(str(ckpt_dir / "tokenizer.json")) if (ckpt_dir / "tokenizer.json").exists() else None
        conds = None
        if (ckpt_dir / "conds.pt").exists():
            conds = Conditionals.load(ckpt_dir / "conds.pt", map_location="cpu")
        flow = {k[len("flow."):]: v for k, v in s3.items() if k.startswith("flow.")}
        hift = {k[len("mel2wav."):]: v for k, v in s3.items() if k.startswith("mel2wav.")}
        return cls.from_state_dicts(t3_state, flow, hift, device, conds=conds, tokenizer=tok)

    @classmethod
    def from_pretrained(cls, device) -> "ChatterboxTTS":
        """reference tts.py:167-180"""
        from huggingface_hub import hf_hub_download
        for fpath in ["ve.safetensors", "t3_cfg.safetensors", "s3gen.safetensors", "tokenizer.json", "conds.pt"]:
            local_path = hf_hub_download(repo_id=REPO_ID, filename=fpath)
        return cls.from_local(Path(local_path).parent, device)

    def _emotion_adv_host(self) -> float:
        """`conds.t3.emotion_adv` as a host float, read from the device once per T3Cond object (the reference
        compares against the tensor on every call, tts.py:225: a host sync per utterance when conds live on the GPU)."""
        c = self.conds.t3
        key = (id(c), id(c.emotion_adv))
        if getattr(self, "_emo_cache", (None, None))[0] != key:
            self._emo_cache = (key, float(c.emotion_adv.reshape(-1)[0]))
        return self._emo_cache[1]

    def prepare_conditionals(self, wav_fpath, exaggeration=0.5):
        raise NotImplementedError("prepare_conditionals (voice encoder, S3 tokenizer, CAM++ x-vector) is outside the "
                                  "accelerated path and its third-party networks are not bundled; load pre-computed "
                                  "Conditionals (conds.pt) instead")

    def generate(self, text=None, repetition_penalty=1.2, min_p=0.05, top_p=1.0, audio_prompt_path=None, exaggeration=0.5,
                 cfg_weight=0.5, temperature=0.8, *, text_tokens: Optional[torch.Tensor] = None, do_sample: bool = True,
                 max_new_tokens: int = 1000, stop_on_eos: bool = True, seed: Optional[int] = None,
                 alignment_analyzer: bool = False):
        if audio_prompt_path:
            self.prepare_conditionals(audio_prompt_path, exaggeration=exaggeration)
        else:
            assert self.conds is not None, "Please `prepare_conditionals` first or specify `audio_prompt_path`"
        if exaggeration != self._emotion_adv_host():
            _cond = self.conds.t3
            self.conds.t3 = T3Cond(speaker_emb=_cond.speaker_emb, cond_prompt_speech_tokens=_cond.cond_prompt_speech_tokens,
                                   emotion_adv=exaggeration * torch.ones(1, 1, 1)).to(device=self.device)
            self._emo_cache = ((id(self.conds.t3), id(self.conds.t3.emotion_adv)), float(exaggeration))
        if text_tokens is None:
            assert self.tokenizer is not None, "no tokenizer.json loaded: pass text_tokens=<ids> instead of text"
            text_tokens = self.tokenizer.text_to_tokens(punc_norm(text))
        text_tokens = torch.atleast_2d(text_tokens).to(self.device)
        if cfg_weight > 0.0:
            text_tokens = torch.cat([text_tokens, text_tokens], dim=0)
        text_tokens = F.pad(text_tokens, (1, 0), value=self.t3.hp.start_text_token)
        text_tokens = F.pad(text_tokens, (0, 1), value=self.t3.hp.stop_text_token)
        with torch.inference_mode():
            c = self.conds.t3
            cond = T3Cond(speaker_emb=c.speaker_emb, cond_prompt_speech_tokens=c.cond_prompt_speech_tokens,
                          emotion_adv=c.emotion_adv).to(device=self.device)
            speech_tokens = self.t3.inference(t3_cond=cond, text_tokens=text_tokens, max_new_tokens=max_new_tokens,
                                              temperature=temperature, cfg_weight=cfg_weight,
                                              repetition_penalty=repetition_penalty, min_p=min_p, top_p=top_p,
                                              do_sample=do_sample, stop_on_eos=stop_on_eos, seed=seed,
                                              alignment_analyzer=alignment_analyzer)
            speech_tokens = drop_invalid_tokens(speech_tokens[0])
            speech_tokens = speech_tokens[speech_tokens < SPEECH_VOCAB_SIZE]
            self.last_speech_tokens = speech_tokens
            wav, _ = self.s3gen.inference(speech_tokens=speech_tokens, ref_dict=self.conds.gen, seed=seed)
            wav = wav.squeeze(0).detach().cpu()
        if self.watermarker is not None:
            wav = torch.from_numpy(self.watermarker.apply_watermark(wav.numpy(), sample_rate=self.sr))
        return wav.unsqueeze(0)
