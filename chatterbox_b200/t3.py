"""Host-side mirror of the reference `T3` (src/chatterbox/models/t3/t3.py:31-371) over the C-ABI.

Same constructor state (reference state-dict keys), same `inference(...)` signature and return
value; the 30-layer backbone, logit processing, sampling and the decode loop run inside
`libcbx.so` (csrc/t3_decode.cu), and so does the once-per-utterance conditioning front-end
(`T3CondEnc` / `Perceiver` / embedding gathers / doubled BOS, t3.py:71-108,284-295: csrc/t3_front.cu, fp32).
There is no CPU path: construction fails without CUDA + libcbx.so.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, Optional

import torch

from . import _lib
from .pack import T3Packed, pack_t3


class T3Config:
    """reference t3/modules/t3_config.py:4-27"""
    start_text_token = 255
    stop_text_token = 0
    text_tokens_dict_size = 704
    max_text_tokens = 2048
    start_speech_token = 6561
    stop_speech_token = 6562
    speech_tokens_dict_size = 8194
    max_speech_tokens = 4096
    llama_config_name = "Llama_520M"
    input_pos_emb = "learned"
    speech_cond_prompt_len = 150
    encoder_type = "voice_encoder"
    speaker_embed_size = 256
    use_perceiver_resampler = True
    emotion_adv = True
    n_channels = 1024


@dataclass
class T3Cond:
    """reference t3/modules/cond_enc.py:11-38"""
    speaker_emb: torch.Tensor
    clap_emb: Optional[torch.Tensor] = None
    cond_prompt_speech_tokens: Optional[torch.Tensor] = None
    cond_prompt_speech_emb: Optional[torch.Tensor] = None
    emotion_adv: Optional[torch.Tensor] = 0.5

    def to(self, *, device=None, dtype=None):
        for k, v in self.__dict__.items():
            if torch.is_tensor(v):
                is_fp = type(v.view(-1)[0].item()) is not int
                setattr(self, k, v.to(device=device, dtype=dtype if is_fp else None))
        return self

    def save(self, fpath):
        torch.save(self.__dict__, fpath)

    @staticmethod
    def load(fpath, map_location="cpu"):
        return T3Cond(**torch.load(fpath, map_location=map_location, weights_only=True))


def _ensure_BOT_EOT(text_tokens: torch.Tensor, hp):
    B = text_tokens.size(0)
    assert (text_tokens == hp.start_text_token).int().sum() >= B, "missing start_text_token"
    assert (text_tokens == hp.stop_text_token).int().sum() >= B, "missing stop_text_token"


class T3:
    def __init__(self, state_dict: Dict[str, torch.Tensor], device="cuda", *, max_pos: int = 4096,
                 kv_fp32: bool = False, n_layers: int = 30, hp=T3Config(), gemm_prefill: bool = True):
        if not torch.cuda.is_available():
            raise _lib.CbxError("chatterbox_b200.T3 needs a CUDA device (no CPU fallback)")
        self.hp = hp
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.dim = 1024
        self._lib = _lib.lib()
        G = self._lib.cbx_t3_grid(self.device.index)
        if G <= 0:
            _lib.check(1)
        self.packed: T3Packed = pack_t3(state_dict, G, self.device, max_pos, n_layers, gemm_prefill=gemm_prefill)
        # front-end parameters (reference key names), fp32 on the device
        keep = ("cond_enc.", "text_emb.", "text_pos_emb.")
        self.fe = {k: v.to(self.device, torch.float32).contiguous() for k, v in state_dict.items() if k.startswith(keep)}
        self.fe["speech_emb.weight"] = self.packed.speech_emb
        self.fe["speech_pos_emb.emb.weight"] = self.packed.speech_pos_emb
        p = self.packed
        cfg = _lib.CbxT3Config(self.device.index, n_layers, max_pos, 1 if kv_fp32 else 0, int(p.speech_pos_emb.shape[0]))
        w = _lib.CbxT3Weights(p.wstream.data_ptr(), p.stride_units, p.ln1.data_ptr(), p.ln2.data_ptr(),
                              p.final_norm.data_ptr(), p.speech_emb.data_ptr(), p.speech_pos_emb.data_ptr(),
                              p.rope_cos.data_ptr(), p.rope_sin.data_ptr(),
                              p.wplain.data_ptr() if p.wplain is not None else None)
        h = C.c_void_p()
        _lib.check(self._lib.cbx_t3_create(C.byref(cfg), C.byref(w), C.byref(h)))
        self._h = h
        self.max_pos = max_pos
        self.last_logits: Optional[torch.Tensor] = None
        # optional CUDA-event stage timers (bench.py): set `record_events = True`, read `events`
        self.record_events = False
        self.events: Dict[str, torch.cuda.Event] = {}

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.cbx_t3_destroy(h)
            self._h = None
        hfe = getattr(self, "_hfe", None)
        if hfe:
            self._lib.cbx_t3_front_destroy(hfe)
            self._hfe = None

    # ---- conditioning front-end (t3.py:71-108, cond_enc.py:64-97, perceiver.py:113-212): csrc/t3_front.cu ----
    def _front_handle(self):
        if getattr(self, "_hfe", None) is None:
            fe = self.fe
            g = lambda k: fe[k].data_ptr() if k in fe else None          # fe tensors are contiguous fp32 and live as long as self
            pa = "cond_enc.perceiver.attn."
            w = _lib.CbxT3FrontWeights(g("cond_enc.spkr_enc.weight"), g("cond_enc.spkr_enc.bias"),
                                       g("cond_enc.perceiver.pre_attention_query"), g(pa + "norm.weight"), g(pa + "norm.bias"),
                                       g(pa + "to_q.weight"), g(pa + "to_q.bias"), g(pa + "to_k.weight"), g(pa + "to_k.bias"),
                                       g(pa + "to_v.weight"), g(pa + "to_v.bias"), g(pa + "proj_out.weight"), g(pa + "proj_out.bias"),
                                       g("cond_enc.emotion_adv_fc.weight"), g("text_emb.weight"), g("text_pos_emb.emb.weight"),
                                       g("speech_emb.weight"), g("speech_pos_emb.emb.weight"))
            h = C.c_void_p()
            _lib.check(self._lib.cbx_t3_front_create(self.device.index, 256, C.byref(w), C.byref(h)))
            self._hfe = h
        return self._hfe

    def _prefill_embeds(self, t3_cond: T3Cond, text_tokens: torch.Tensor, cfg_weight: float, n_bos: int):
        """-> (embeds (B, P, 1024) fp32 = [speaker | 32 perceiver rows | emotion | text | BOS x n_bos], len_cond).
        All inputs stay on the device (no host synchronisation)."""
        assert t3_cond.clap_emb is None, "clap_embed not implemented"
        assert t3_cond.emotion_adv is not None
        dev = self.device
        text_tokens = torch.atleast_2d(text_tokens).to(device=dev, dtype=torch.long).contiguous()
        B, N = text_tokens.shape
        if N > self.fe["text_pos_emb.emb.weight"].shape[0]:
            raise ValueError("text longer than the learned text position table")
        spk = t3_cond.speaker_emb.to(dev, torch.float32).reshape(-1).contiguous()
        emo = torch.as_tensor(t3_cond.emotion_adv, dtype=torch.float32, device=dev).reshape(-1).contiguous()
        toks = t3_cond.cond_prompt_speech_tokens
        pe = t3_cond.cond_prompt_speech_emb
        n_cond, toks_p, pe_p = 0, None, None
        if pe is not None:
            pe = pe.to(dev, torch.float32).reshape(-1, 1024).contiguous()
            n_cond, pe_p = pe.shape[0], pe.data_ptr()
        elif toks is not None:
            toks = toks.to(device=dev, dtype=torch.long).reshape(-1).contiguous()
            n_cond, toks_p = toks.numel(), toks.data_ptr()
        len_cond = 2 + (32 if n_cond else 0)
        out = torch.empty(B, len_cond + N + n_bos, 1024, dtype=torch.float32, device=dev)
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(self._lib.cbx_t3_front(self._front_handle(), spk.data_ptr(), toks_p, pe_p, n_cond, emo.data_ptr(),
                                          text_tokens.data_ptr(), B, N, 1 if cfg_weight > 0.0 else 0, n_bos, out.data_ptr(), stream))
        return out, len_cond

    def prepare_conditioning(self, t3_cond: T3Cond) -> torch.Tensor:
        """t3.py:71-78 -> (1, len_cond, 1024)."""
        sot_eot = torch.tensor([[self.hp.start_text_token, self.hp.stop_text_token]], device=self.device)
        out, len_cond = self._prefill_embeds(t3_cond, sot_eot, 0.0, 1)
        return out[:, :len_cond]

    def prepare_input_embeds(self, *, t3_cond: T3Cond, text_tokens, speech_tokens, cfg_weight: float = 0.0):
        """t3.py:80-108 for the initial speech tokens `T3.inference` passes (one BOS per row)."""
        st = torch.atleast_2d(speech_tokens)
        assert st.shape[1] == 1, "only the initial BOS speech token is supported here (T3.inference's own call)"
        out, len_cond = self._prefill_embeds(t3_cond, text_tokens, cfg_weight, 1)
        return out, len_cond

    # ---- the decode path ----
    @torch.inference_mode()
    def inference(self, *, t3_cond: T3Cond, text_tokens: torch.Tensor, initial_speech_tokens=None,
                  prepend_prompt_speech_tokens=None, num_return_sequences=1, max_new_tokens=None,
                  stop_on_eos=True, do_sample=True, temperature=0.8, min_p=0.05, top_p=1.00,
                  length_penalty=1.0, repetition_penalty=1.2, cfg_weight=0,
                  forced_tokens: Optional[torch.Tensor] = None, return_logits: bool = False,
                  seed: Optional[int] = None, alignment_analyzer: bool = False, alignment_layer_idx: int = 9,
                  alignment_inject: Optional[torch.Tensor] = None, alignment_debug: bool = False):
        """Reference signature (t3.py:203-225) plus three keyword extensions used by the parity tests:
        `forced_tokens` (teacher forcing), `return_logits` (keep raw per-step logits in
        `self.last_logits`), `seed` (Philox seed).  `do_sample=False` selects greedy argmax (the
        reference ignores the flag and always samples).
        `alignment_analyzer=True` runs `AlignmentStreamAnalyzer` (inference/alignment_stream_analyzer.py:29-154; the
        reference constructs the backend with `alignment_stream_analyzer=None`, t3.py:260, and keeps the `step` call
        commented out, t3_hf_backend.py:108-109) on the device: layer `alignment_layer_idx` head-mean attention over the
        text positions, EOS forced / suppressed before every token is sampled.  `alignment_debug=True` keeps the rows and
        per-step decisions in `self.last_alignment`; `alignment_inject` (rows, n_text) replaces the measured rows."""
        assert prepend_prompt_speech_tokens is None, "not implemented"
        _ensure_BOT_EOT(text_tokens, self.hp)
        text_tokens = torch.atleast_2d(text_tokens).to(dtype=torch.long, device=self.device)
        if initial_speech_tokens is None:
            initial_speech_tokens = self.hp.start_speech_token * torch.ones_like(text_tokens[:, :1])
        # conditioning prefix | text | BOS (| BOS: the reference appends BOS a second time on the CFG path, t3.py:284-295)
        embeds, len_cond = self._prefill_embeds(t3_cond, text_tokens, cfg_weight, 2 if cfg_weight > 0 else 1)
        B, P, _ = embeds.shape
        assert B in (1, 2), "T3.inference runs one utterance (optionally as a CFG pair)"
        if max_new_tokens is None:
            max_new_tokens = self.hp.max_speech_tokens
        max_new_tokens = int(max_new_tokens)
        if P + max_new_tokens > self.max_pos:
            raise ValueError(f"context {P}+{max_new_tokens} exceeds the KV slab (max_pos={self.max_pos})")
        if max_new_tokens + 1 > self.packed.speech_pos_emb.shape[0]:
            raise ValueError("max_new_tokens exceeds the learned speech position table")
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        tokens = torch.zeros(max_new_tokens, dtype=torch.int32, device=self.device)
        n_out = torch.zeros(1, dtype=torch.int32, device=self.device)
        logits = torch.empty(max_new_tokens, B, 8194, dtype=torch.float32, device=self.device) if return_logits else None
        forced = None
        if forced_tokens is not None:
            forced = forced_tokens.reshape(-1).to(device=self.device, dtype=torch.int32).contiguous()
            assert forced.numel() >= max_new_tokens
        sp = _lib.CbxSampleParams(float(cfg_weight), float(temperature), float(repetition_penalty), float(min_p),
                                  float(top_p), 0 if do_sample else 1, 1 if stop_on_eos else 0,
                                  int(seed if seed is not None else torch.initial_seed()) & (2 ** 64 - 1))
        al_rows = al_dec = al_inj = None
        if alignment_analyzer:
            n_text = int(text_tokens.shape[1])
            if alignment_inject is not None:
                al_inj = alignment_inject.to(self.device, torch.float32).contiguous()
                assert al_inj.shape[1] == n_text and al_inj.shape[0] >= max_new_tokens + 1
            if alignment_debug:
                al_rows = torch.zeros(max_new_tokens + 2, n_text, dtype=torch.float32, device=self.device)
                al_dec = torch.zeros(max_new_tokens, 4, dtype=torch.int32, device=self.device)
            _lib.check(self._lib.cbx_t3_set_alignment(self._h, int(alignment_layer_idx), int(len_cond), n_text,
                                                      al_inj.data_ptr() if al_inj is not None else None,
                                                      al_rows.data_ptr() if al_rows is not None else None,
                                                      al_dec.data_ptr() if al_dec is not None else None))
        else:
            _lib.check(self._lib.cbx_t3_set_alignment(self._h, -1, 0, 0, None, None, None))
        ev = None
        if self.record_events:
            ev = {k: torch.cuda.Event(enable_timing=True) for k in ("prefill0", "decode0", "decode1")}
            ev["prefill0"].record()
        _lib.check(self._lib.cbx_t3_prefill(self._h, embeds.data_ptr(), B, P, stream))
        if ev:
            ev["decode0"].record()
        _lib.check(self._lib.cbx_t3_decode(self._h, C.byref(sp), max_new_tokens,
                                           forced.data_ptr() if forced is not None else None,
                                           tokens.data_ptr(), n_out.data_ptr(),
                                           logits.data_ptr() if logits is not None else None, stream))
        if ev:
            ev["decode1"].record()
            self.events = ev
        self.last_prefill_len = P
        n = int(n_out.item())            # the only host sync of the utterance (a device error comes back as n < 0)
        if n < 0:
            raise _lib.CbxError(f"T3 megakernel reported a device error after {-n - 1} tokens (watchdog timeout)")
        self.last_logits = logits[:n] if logits is not None else None
        self.last_alignment = dict(rows=al_rows, decisions=al_dec[:n]) if al_dec is not None else None
        return tokens[:n].to(torch.long)[None]

    def state(self):
        pos, n_gen, err, launches = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int64()
        _lib.check(self._lib.cbx_t3_state(self._h, C.byref(pos), C.byref(n_gen), C.byref(err), C.byref(launches)))
        return dict(pos=pos.value, n_gen=n_gen.value, err=err.value, launches=launches.value)

    def check_device_state(self):
        s = self.state()
        if s["err"]:
            raise _lib.CbxError(f"T3 megakernel reported device error {s['err']} (watchdog timeout)")
        return s
