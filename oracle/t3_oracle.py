"""CPU restatement of the reference's T3 inference path (TEST INFRASTRUCTURE — checker only).

Plain torch fp32 on CPU, no dependence on `/root/reference` or `transformers`, so it travels to
the GPU box.  Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline /
`--impl reference` legs may import it; the product path (`chatterbox_b200/`) never does.

Parity pin: `oracle/make_golden.py` (run in the build container) checks every function here
against the UNMODIFIED reference modules imported from `/root/reference` (through
`oracle/ref_import.py`) with identical synthetic weights, and commits the reference's outputs
under `tests/golden/`; `tests/test_oracle_t3.py` re-checks this restatement against those files.
The reference itself ships no golden vectors for this path.

What each function follows (paths relative to /root/reference/src/chatterbox/models/t3):
  rope_inv_freq        transformers `_compute_llama3_parameters` with llama_configs.py:22-30
  cond_enc             modules/cond_enc.py:64-97, modules/perceiver.py:113-212, t3.py:71-78
  prepare_input_embeds t3.py:80-108 (+ the doubled BOS of t3.py:284-295)
  LlamaOracle          HF LlamaModel/LlamaDecoderLayer (transformers 4.46.3 pinned by
                       pyproject.toml:17): RMSNorm eps 1e-5, q/k/v/o no bias, rotate_half RoPE,
                       fp32 softmax, SwiGLU; speech_head of inference/t3_hf_backend.py:103
  process_logits       t3.py:320-337 with HF RepetitionPenalty / MinP / TopP processors
  inference            t3.py:203-371 (greedy variant: argmax in place of multinomial at :341;
                       teacher-forced variant: feed given tokens, record raw logits)
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

H, NH, HD, INTER = 1024, 16, 64, 4096
EPS = 1e-5
START_SPEECH, STOP_SPEECH = 6561, 6562


def rope_inv_freq() -> torch.Tensor:
    theta, factor, low_f, high_f, old_ctx = 500000.0, 8.0, 1.0, 4.0, 8192
    inv_freq = 1.0 / (theta ** (torch.arange(0, HD, 2, dtype=torch.int64).to(torch.float) / HD))
    low_wl, high_wl = old_ctx / low_f, old_ctx / high_f
    wavelen = 2 * math.pi / inv_freq
    inv_l = torch.where(wavelen > low_wl, inv_freq / factor, inv_freq)
    smooth = (old_ctx / wavelen - low_f) / (high_f - low_f)
    smoothed = (1 - smooth) * inv_l / factor + smooth * inv_l
    medium = ~(wavelen < high_wl) * ~(wavelen > low_wl)
    return torch.where(medium, smoothed, inv_l)


def rope_cos_sin(positions: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """(T,) int -> cos, sin of shape (T, 64) exactly as LlamaRotaryEmbedding.forward (fp32)."""
    freqs = positions.to(torch.float32)[:, None] * rope_inv_freq()[None, :]
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos(), emb.sin()


def _rotate_half(x):
    return torch.cat((-x[..., HD // 2:], x[..., : HD // 2]), dim=-1)


def _rmsnorm(x, w):
    var = x.pow(2).mean(-1, keepdim=True)
    return w * (x * torch.rsqrt(var + EPS))


# ---------------------------------------------------------------------------------------------
# conditioning front-end
# ---------------------------------------------------------------------------------------------
def _attn_block2(sd, pfx, x1, x2):
    """perceiver.AttentionBlock2.forward with flash (SDPA) attention, 4 heads x 256."""
    def lin(n, x):
        return F.linear(x, sd[pfx + n + ".weight"], sd[pfx + n + ".bias"])
    x1n = F.layer_norm(x1, (H,), sd[pfx + "norm.weight"], sd[pfx + "norm.bias"])
    x2n = F.layer_norm(x2, (H,), sd[pfx + "norm.weight"], sd[pfx + "norm.bias"])
    q, k, v = lin("to_q", x1n), lin("to_k", x2n), lin("to_v", x2n)

    def split(t):
        b, l, _ = t.shape
        return t.view(b, l, 4, H // 4).permute(0, 2, 1, 3)
    o = F.scaled_dot_product_attention(split(q), split(k), split(v))
    o = o.permute(0, 2, 1, 3).contiguous().view(x1.shape[0], x1.shape[1], H)
    return x1 + lin("proj_out", o)


def cond_enc(sd: Dict[str, torch.Tensor], cond: Dict[str, torch.Tensor]) -> torch.Tensor:
    """-> (1, 34, 1024): [speaker | 32 perceiver tokens | emotion]."""
    spk = F.linear(cond["speaker_emb"].view(-1, 256), sd["cond_enc.spkr_enc.weight"], sd["cond_enc.spkr_enc.bias"])[:, None]
    toks = cond["cond_prompt_speech_tokens"]
    pe = sd["speech_emb.weight"][toks] + sd["speech_pos_emb.emb.weight"][: toks.shape[1]][None]
    q = sd["cond_enc.perceiver.pre_attention_query"].expand(pe.shape[0], -1, -1)
    pre = _attn_block2(sd, "cond_enc.perceiver.attn.", q, pe)
    per = _attn_block2(sd, "cond_enc.perceiver.attn.", pre, pre)
    emo = F.linear(cond["emotion_adv"].view(-1, 1, 1), sd["cond_enc.emotion_adv_fc.weight"])
    return torch.cat((spk, per, emo), dim=1)


def prepare_input_embeds(sd, cond, text_tokens: torch.Tensor, cfg_weight: float) -> torch.Tensor:
    """text_tokens (B, N) incl. SOT/EOT, B = 2 when cfg_weight > 0 -> prefill embeds (B, P, 1024)
    = [cond(34) | text(N) | BOS | BOS] (single BOS when cfg_weight == 0, t3.py:292-295)."""
    ce = cond_enc(sd, cond)
    te = sd["text_emb.weight"][text_tokens].clone()
    if cfg_weight > 0.0:
        te[1].zero_()
    te = te + sd["text_pos_emb.emb.weight"][: text_tokens.shape[1]][None]
    bos = (sd["speech_emb.weight"][START_SPEECH] + sd["speech_pos_emb.emb.weight"][0])[None, None]
    B = text_tokens.shape[0]
    ce = ce.expand(B, -1, -1)
    bos = bos.expand(B, -1, -1)
    parts = [ce, te, bos] + ([bos] if cfg_weight > 0.0 else [])
    return torch.cat(parts, dim=1)


# ---------------------------------------------------------------------------------------------
# backbone
# ---------------------------------------------------------------------------------------------
def _bf16_round(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).to(torch.float32)


class LlamaOracle:
    """30-layer decoder with a growing KV cache.  `weight_round`/`kv_round` emulate the build's
    bf16 storage (weights rounded once, K/V rounded when cached) with fp32 arithmetic, so the
    CUDA bf16 mode can be compared tightly; both default to exact fp32 = the reference."""

    def __init__(self, sd: Dict[str, torch.Tensor], n_layers: int = 30, weight_round: bool = False,
                 kv_round: bool = False):
        r = _bf16_round if weight_round else (lambda t: t)
        self.n_layers = n_layers
        self.kv_round = kv_round
        self.L = []
        for i in range(n_layers):
            p = f"tfmr.layers.{i}."
            self.L.append(dict(
                ln1=sd[p + "input_layernorm.weight"], ln2=sd[p + "post_attention_layernorm.weight"],
                wqkv=r(torch.cat([sd[p + f"self_attn.{n}.weight"] for n in ("q_proj", "k_proj", "v_proj")], 0)),
                wo=r(sd[p + "self_attn.o_proj.weight"]),
                wgu=r(torch.cat([sd[p + "mlp.gate_proj.weight"], sd[p + "mlp.up_proj.weight"]], 0)),
                wd=r(sd[p + "mlp.down_proj.weight"])))
        self.norm = sd["tfmr.norm.weight"]
        self.head = r(sd["speech_head.weight"])
        self.k: List[Optional[torch.Tensor]] = [None] * n_layers
        self.v: List[Optional[torch.Tensor]] = [None] * n_layers
        self.pos = 0
        self.attn_probe_layer: Optional[int] = None
        self.attn_probe: Optional[torch.Tensor] = None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x (B, T, 1024) at absolute positions [pos, pos+T) -> logits (B, T, 8194)."""
        B, T, _ = x.shape
        dev = x.device
        cos, sin = rope_cos_sin(torch.arange(self.pos, self.pos + T))
        cos, sin = cos.to(dev), sin.to(dev)
        C = self.pos + T
        mask = torch.full((T, C), float("-inf"), device=dev).triu(diagonal=self.pos + 1)
        for li, w in enumerate(self.L):
            h = _rmsnorm(x, w["ln1"])
            qkv = F.linear(h, w["wqkv"]).view(B, T, 3, NH, HD)
            q, k, v = (qkv[:, :, j].transpose(1, 2) for j in range(3))
            q = q * cos + _rotate_half(q) * sin
            k = k * cos + _rotate_half(k) * sin
            if self.kv_round:
                k, v = _bf16_round(k), _bf16_round(v)
            self.k[li] = k if self.k[li] is None else torch.cat([self.k[li], k], 2)
            self.v[li] = v if self.v[li] is None else torch.cat([self.v[li], v], 2)
            s = torch.matmul(q, self.k[li].transpose(2, 3)) * (HD ** -0.5) + mask
            p = torch.softmax(s, dim=-1, dtype=torch.float32)
            if self.attn_probe_layer == li:
                self.attn_probe = p
            o = torch.matmul(p, self.v[li]).transpose(1, 2).reshape(B, T, H)
            x = x + F.linear(o, w["wo"])
            h = _rmsnorm(x, w["ln2"])
            gu = F.linear(h, w["wgu"])
            x = x + F.linear(F.silu(gu[..., :INTER]) * gu[..., INTER:], w["wd"])
        self.pos += T
        return F.linear(_rmsnorm(x, self.norm), self.head)


# ---------------------------------------------------------------------------------------------
# logits processing and the decode loop
# ---------------------------------------------------------------------------------------------
def process_logits(logits: torch.Tensor, generated_ids: torch.Tensor, *, cfg_weight: float,
                   temperature: float, repetition_penalty: float, min_p: float, top_p: float) -> torch.Tensor:
    """logits (B, V) of the CFG pair -> processed (1, V).  Order of t3.py:320-337."""
    if cfg_weight > 0.0:
        c, u = logits[0:1], logits[1:2]
        logits = c + cfg_weight * (c - u)
    if temperature != 1.0:
        logits = logits / temperature
    # RepetitionPenaltyLogitsProcessor
    score = torch.gather(logits, 1, generated_ids)
    score = torch.where(score < 0, score * repetition_penalty, score / repetition_penalty)
    logits = logits.scatter(1, generated_ids, score)
    # MinPLogitsWarper (min_tokens_to_keep = 1)
    probs = torch.softmax(logits, dim=-1)
    remove = probs < min_p * probs.amax(dim=-1, keepdim=True)
    remove.scatter_(-1, torch.topk(probs, 1, dim=-1).indices, False)
    logits = logits.masked_fill(remove, float("-inf"))
    # TopPLogitsWarper (min_tokens_to_keep = 1)
    sl, si = torch.sort(logits, descending=False)
    cum = sl.softmax(dim=-1).cumsum(dim=-1)
    rm = cum <= (1 - top_p)
    rm[..., -1:] = False
    remove = rm.scatter(1, si, rm)
    return logits.masked_fill(remove, float("-inf"))


def inference(sd, cond, text_tokens: torch.Tensor, *, max_new_tokens: int, cfg_weight: float = 0.5,
              temperature: float = 0.8, repetition_penalty: float = 1.2, min_p: float = 0.05,
              top_p: float = 1.0, forced_tokens: Optional[torch.Tensor] = None, stop_on_eos: bool = True,
              n_layers: int = 30, weight_round: bool = False, kv_round: bool = False,
              return_model: bool = False, attn_probe_layer: Optional[int] = None, analyzer: Optional["AlignmentOracle"] = None):
    """Greedy (argmax) or teacher-forced restatement of T3.inference.
    Returns (tokens (1, n) int64, raw_logits (n, B, V) — the `output.logits[:, -1]` of each step)."""
    emb = prepare_input_embeds(sd, cond, text_tokens, cfg_weight)
    B = emb.shape[0]
    m = LlamaOracle(sd, n_layers=n_layers, weight_round=weight_round, kv_round=kv_round)
    if analyzer is not None and attn_probe_layer is None:
        attn_probe_layer = 9
    m.attn_probe_layer = attn_probe_layer
    m.attn_rows = []          # head-mean attention of the conditional row over the text columns, per forward pass
    ti, tj = 34, 34 + text_tokens.shape[1]

    def probe_rows(first: bool):
        a = m.attn_probe[0].mean(0)                       # (T_q, T_k)
        return a[tj:, ti:tj] if first else a[:, ti:tj]
    logits = m.forward(emb)[:, -1]
    if attn_probe_layer is not None:
        m.attn_rows.append(probe_rows(True).clone())
    gen = torch.tensor([[START_SPEECH]], dtype=torch.long)
    toks, raw = [], []
    for i in range(max_new_tokens):
        if analyzer is not None:          # t3_hf_backend.py:108-109 (the call the reference keeps commented out)
            logits = analyzer.step(logits[:, None], m.attn_rows[-1])[:, 0]
        raw.append(logits.clone())
        proc = process_logits(logits, gen, cfg_weight=cfg_weight, temperature=temperature,
                              repetition_penalty=repetition_penalty, min_p=min_p, top_p=top_p)
        nxt = proc.argmax(dim=-1, keepdim=True) if forced_tokens is None else forced_tokens[:, i:i + 1]
        toks.append(nxt)
        gen = torch.cat([gen, nxt], dim=1)
        if stop_on_eos and int(nxt) == STOP_SPEECH:
            break
        e = (sd["speech_emb.weight"][nxt] + sd["speech_pos_emb.emb.weight"][i + 1]).expand(B, -1, -1)
        logits = m.forward(e)[:, -1]
        if attn_probe_layer is not None:
            m.attn_rows.append(probe_rows(False).clone())
    out = (torch.cat(toks, dim=1), torch.stack(raw))
    return out + (m,) if return_model else out


# ---------------------------------------------------------------------------------------------
# AlignmentStreamAnalyzer (inference/alignment_stream_analyzer.py:29-154)
# ---------------------------------------------------------------------------------------------
class AlignmentOracle:
    """Restatement of `AlignmentStreamAnalyzer.step` (alignment_stream_analyzer.py:89-154) with O(1) state instead of
    the growing alignment matrix A (T, S): every quantity the reference reads off A is a running statistic —
      A[-2:, -2:].max()                      -> last-two-columns maximum of the two most recent rows
      A[:, :4].max()                         -> running maximum over the first four columns
      A[completed_at:, -3:].sum(0).max()     -> three running column sums over rows appended after completion
      A[completed_at:, :-5].max(1).sum()     -> running sum of row maxima over columns [:S-5] after completion
    `step(logits, rows)` takes the head-mean layer-9 attention of the conditional row over the text columns for the
    queries of this forward pass ((2, S) or (1, S) for the prompt, (1, S) per decode step) and returns the logits with
    EOS forced / suppressed exactly as the reference (±2**15).  tests/test_oracle_t3.py drives this and the reference
    class side by side (golden: tests/golden/t3_align_golden.safetensors)."""

    def __init__(self, n_text: int, eos_idx: int = STOP_SPEECH):
        self.S, self.eos = n_text, eos_idx
        self.frame, self.text_position = 0, 0
        self.started, self.started_at = False, None
        self.complete, self.completed_at = False, None
        self.T = 0
        self.last2 = [float("-inf"), float("-inf")]      # last-two-columns max of the previous / current row
        self.first4 = float("-inf")
        self.tail = [0.0, 0.0, 0.0]
        self.rep = 0.0
        self.history = []          # per step: (cur_text_posn, text_position, started, complete, forced, suppressed)

    def step(self, logits: torch.Tensor, rows: torch.Tensor) -> torch.Tensor:
        S = self.S
        rows = rows.clone().float()
        rows[:, self.frame + 1:] = 0                                       # monotonic masking (:107)
        for r in rows:
            self.T += 1
            self.last2 = [self.last2[1], float(r[-2:].max())]
            self.first4 = max(self.first4, float(r[:4].max()))
            if self.complete:                                              # rows at index >= completed_at
                for k in range(3):
                    self.tail[k] += float(r[S - 3 + k])
                if S > 5:
                    self.rep += float(r[:S - 5].max())
        cur = int(rows[-1].argmax())
        if -4 < cur - self.text_position < 7:                              # not a discontinuity (:118-121)
            self.text_position = cur
        a_last2 = max(self.last2) if self.T >= 2 else self.last2[1]
        false_start = (not self.started) and (a_last2 > 0.1 or self.first4 < 0.5)
        self.started = not false_start
        if self.started and self.started_at is None:
            self.started_at = self.T
        self.complete = self.complete or self.text_position >= S - 3
        if self.complete and self.completed_at is None:
            self.completed_at = self.T
        long_tail = self.complete and max(self.tail) >= 10
        repetition = self.complete and self.rep > 5
        forced = bool(long_tail or repetition)
        if forced:
            logits = -(2 ** 15) * torch.ones_like(logits)
            logits[..., self.eos] = 2 ** 15
        suppressed = cur < S - 3
        if suppressed:
            logits = logits.clone()
            logits[..., self.eos] = -2 ** 15
        self.frame += 1
        self.history.append((cur, self.text_position, self.started, self.complete, forced, suppressed))
        return logits
