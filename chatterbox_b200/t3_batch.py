"""Batched T3 decode + request scheduler over the C-ABI (`cbx_t3_batch_*`, csrc/t3_batch.cu).

The reference serves one request at a time: `src/server/server.py:189-299` calls `tts_model.generate` synchronously
per request (:223,232) and `src/server/voice_manager.py:370-379` swaps the global `model.conds` to change voice.
Here up to 64 utterances share every decode step (BASELINE config 4): `T3Batch` owns U "slots", each with its own KV
slab, conditionals, sampling knobs, Philox stream and EOS state; `BatchScheduler` is the continuous-batching loop the
server would drive — admit requests into free slots (prompt prefill), step all running slots, retire finished ones.

Per-utterance arithmetic is the single-utterance path's (`T3.inference`, t3.py): hi/lo activation planes against bf16
weights, fp32 accumulation, same logit processing and sampler.  There is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

from . import _lib
from .t3 import T3, T3Cond, _ensure_BOT_EOT


@dataclass
class T3Request:
    """One utterance: what `ChatterboxTTS.generate` passes to `T3.inference` (tts.py:245-256)."""
    t3_cond: T3Cond
    text_tokens: torch.Tensor                 # (1 or 2, N) incl. SOT / EOT; one row is duplicated for the CFG pair
    max_new_tokens: int = 1000
    temperature: float = 0.8
    cfg_weight: float = 0.5
    repetition_penalty: float = 1.2
    min_p: float = 0.05
    top_p: float = 1.0
    do_sample: bool = True
    stop_on_eos: bool = True
    seed: int = 0
    forced_tokens: Optional[torch.Tensor] = None
    tag: object = None                        # caller's handle (request id)
    tokens: Optional[torch.Tensor] = field(default=None, repr=False)   # filled on retirement: (1, n) int64


class T3Batch:
    def __init__(self, t3: T3, n_slots: int, *, max_pos: int = 1024, max_new_cap: int = 1000, kv_fp32: bool = False,
                 keep_logits: bool = False):
        if t3.packed.wplain is None:
            raise _lib.CbxError("T3Batch needs the plain weight matrices: build T3 with gemm_prefill=True")
        self.t3 = t3
        self.device = t3.device
        self._lib = t3._lib
        self.n_slots, self.max_pos, self.max_new_cap, self.keep_logits = n_slots, max_pos, max_new_cap, keep_logits
        p = t3.packed
        if p.rope_cos.shape[0] < max_pos:
            raise ValueError(f"T3 was packed with RoPE tables for {p.rope_cos.shape[0]} positions < max_pos {max_pos}")
        cfg = _lib.CbxT3BatchConfig(self.device.index, p.n_layers, n_slots, max_pos, 1 if kv_fp32 else 0, max_new_cap,
                                    1 if keep_logits else 0, int(p.speech_pos_emb.shape[0]))
        w = _lib.CbxT3Weights(p.wstream.data_ptr(), p.stride_units, p.ln1.data_ptr(), p.ln2.data_ptr(),
                              p.final_norm.data_ptr(), p.speech_emb.data_ptr(), p.speech_pos_emb.data_ptr(),
                              p.rope_cos.data_ptr(), p.rope_sin.data_ptr(), p.wplain.data_ptr())
        h = C.c_void_p()
        _lib.check(self._lib.cbx_t3_batch_create(C.byref(cfg), C.byref(w), C.byref(h)))
        self._h = h
        self.prefill_len: Dict[int, int] = {}

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.cbx_t3_batch_destroy(h)
            self._h = None

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    @torch.inference_mode()
    def admit(self, slot: int, req: T3Request) -> int:
        """Prompt prefill of `req` into `slot` (t3.py:228-316).  Returns the prefill length P."""
        t3 = self.t3
        tt = torch.atleast_2d(req.text_tokens).to(dtype=torch.long, device=self.device)
        _ensure_BOT_EOT(tt, t3.hp)
        if tt.shape[0] == 1:
            tt = torch.cat([tt, tt], 0)
        # the engine always runs the CFG pair; with cfg_weight == 0 the second row is computed and ignored (its text
        # embeddings are NOT zeroed, t3.py:90-92, and there is a single BOS, t3.py:292-295)
        embeds, _ = t3._prefill_embeds(req.t3_cond, tt, req.cfg_weight, 2 if req.cfg_weight > 0 else 1)
        P = embeds.shape[1]
        n = int(req.max_new_tokens)
        forced = None
        if req.forced_tokens is not None:
            forced = req.forced_tokens.reshape(-1).to(device=self.device, dtype=torch.int32).contiguous()
            assert forced.numel() >= n
        sp = _lib.CbxSampleParams(float(req.cfg_weight), float(req.temperature), float(req.repetition_penalty),
                                  float(req.min_p), float(req.top_p), 0 if req.do_sample else 1,
                                  1 if req.stop_on_eos else 0, int(req.seed) & (2 ** 64 - 1))
        _lib.check(self._lib.cbx_t3_batch_prefill(self._h, slot, embeds.data_ptr(), P, C.byref(sp), n,
                                                  forced.data_ptr() if forced is not None else None, self._stream()))
        self.prefill_len[slot] = P
        return P

    @torch.inference_mode()
    def admit_many(self, slot0: int, reqs: List[T3Request]) -> List[int]:
        """Prompt prefill of `reqs` into the consecutive slots slot0, slot0 + 1, ...: runs of requests with the same prefill
        length, token budget and forcing mode go through the layers TOGETHER (cbx_t3_batch_prefill_many: 2 n P rows per GEMM);
        anything else falls back to `admit`.  Returns the prefill lengths."""
        t3 = self.t3
        prepped = []
        for req in reqs:
            tt = torch.atleast_2d(req.text_tokens).to(dtype=torch.long, device=self.device)
            _ensure_BOT_EOT(tt, t3.hp)
            if tt.shape[0] == 1:
                tt = torch.cat([tt, tt], 0)
            embeds, _ = t3._prefill_embeds(req.t3_cond, tt, req.cfg_weight, 2 if req.cfg_weight > 0 else 1)
            prepped.append(embeds)
        out, i = [], 0
        while i < len(reqs):
            P, n = prepped[i].shape[1], int(reqs[i].max_new_tokens)
            j = i + 1
            while (j < len(reqs) and prepped[j].shape[1] == P and int(reqs[j].max_new_tokens) == n
                   and (reqs[j].forced_tokens is None) == (reqs[i].forced_tokens is None)):
                j += 1
            run = reqs[i:j]
            emb = torch.cat([e.reshape(1, 2, P, -1) for e in prepped[i:j]], 0).contiguous().float()
            forced = None
            if run[0].forced_tokens is not None:
                forced = torch.stack([r.forced_tokens.reshape(-1)[:n].to(device=self.device, dtype=torch.int32) for r in run]).contiguous()
            sps = (_lib.CbxSampleParams * len(run))(*[
                _lib.CbxSampleParams(float(r.cfg_weight), float(r.temperature), float(r.repetition_penalty), float(r.min_p),
                                     float(r.top_p), 0 if r.do_sample else 1, 1 if r.stop_on_eos else 0, int(r.seed) & (2 ** 64 - 1))
                for r in run])
            _lib.check(self._lib.cbx_t3_batch_prefill_many(self._h, slot0 + i, len(run), emb.data_ptr(), P, sps, n,
                                                           forced.data_ptr() if forced is not None else None, self._stream()))
            for k in range(i, j):
                self.prefill_len[slot0 + k] = P
                out.append(P)
            i = j
        return out

    def step(self, n_steps: int = 1, slots: Optional[List[int]] = None) -> None:
        arr, n = None, 0
        if slots is not None:
            n = len(slots)
            arr = (C.c_int32 * n)(*slots)
        _lib.check(self._lib.cbx_t3_decode_batch(self._h, arr, n, int(n_steps), self._stream()))

    def poll(self):
        """Host sync.  -> (n_gen, done, pos) lists; done: 1 finished, 0 running, -1 free."""
        U = self.n_slots
        a, b, c = (C.c_int32 * U)(), (C.c_int32 * U)(), (C.c_int32 * U)()
        _lib.check(self._lib.cbx_t3_batch_poll(self._h, a, b, c, self._stream()))
        return list(a), list(b), list(c)

    def tokens(self, slot: int, n: int) -> torch.Tensor:
        out = torch.empty(n, dtype=torch.int32, device=self.device)
        _lib.check(self._lib.cbx_t3_batch_tokens(self._h, slot, out.data_ptr(), n, self._stream()))
        return out.to(torch.long)[None]

    def logits(self, slot: int, n: int) -> torch.Tensor:
        """keep_logits engines: raw per-step logits of `slot` as an (n, 2, 8194) fp32 tensor."""
        out = torch.empty(n, 2, 8194, dtype=torch.float32, device=self.device)
        _lib.check(self._lib.cbx_t3_batch_logits(self._h, slot, out.data_ptr(), n, self._stream()))
        return out

    def release(self, slot: int) -> None:
        _lib.check(self._lib.cbx_t3_batch_release(self._h, slot, self._stream()))

    def info(self):
        s, a, b = C.c_int32(), C.c_int64(), C.c_int64()
        _lib.check(self._lib.cbx_t3_batch_info(self._h, C.byref(s), C.byref(a), C.byref(b)))
        return dict(kv_splits=s.value, launches=a.value, launches_per_step=b.value)


class BatchScheduler:
    """Continuous batching over a `T3Batch`: the loop a server puts behind `/generate` instead of one synchronous
    `generate` per request (reference src/server/server.py:189-299).  `submit` queues requests; `run` admits queued
    requests into free slots (prompt prefill), advances every running slot `poll_every` steps per host round trip,
    retires finished slots (tokens copied out, slot released) and calls `on_finish(request)` — where the caller chains
    S3Gen — until the queue and the slots are empty.  Requests keep their own conditionals and knobs."""

    def __init__(self, engine: T3Batch, poll_every: int = 8):
        self.engine = engine
        self.poll_every = poll_every
        self.queue: List[T3Request] = []
        self.slots: List[Optional[T3Request]] = [None] * engine.n_slots
        self.steps = 0

    def submit(self, req: T3Request) -> None:
        self.queue.append(req)

    def _admit(self) -> None:
        """Queued requests into free slots, in queue order; a run of consecutive free slots is filled with ONE admit_many call
        (prompts of equal length then share their prefill pass)."""
        s, U = 0, self.engine.n_slots
        while s < U and self.queue:
            if self.slots[s] is not None:
                s += 1
                continue
            e = s
            while e < U and self.slots[e] is None and e - s < len(self.queue):
                e += 1
            reqs = [self.queue.pop(0) for _ in range(e - s)]
            if len(reqs) == 1:
                self.engine.admit(s, reqs[0])
            else:
                self.engine.admit_many(s, reqs)
            self.slots[s:e] = reqs
            s = e

    def run(self, on_finish=None) -> List[T3Request]:
        finished: List[T3Request] = []
        eng = self.engine
        while self.queue or any(r is not None for r in self.slots):
            self._admit()
            eng.step(self.poll_every)
            self.steps += self.poll_every
            n_gen, done, _ = eng.poll()
            for s, req in enumerate(self.slots):
                if req is not None and done[s] == 1:
                    req.tokens = eng.tokens(s, n_gen[s])
                    eng.release(s)
                    self.slots[s] = None
                    finished.append(req)
                    if on_finish is not None:
                        on_finish(req)
        return finished
