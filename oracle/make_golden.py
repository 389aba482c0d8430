"""Generate the pinned vectors under tests/golden/ from the UNMODIFIED reference (container only).

TEST INFRASTRUCTURE.  Run:  python -m oracle.make_golden [t3] [t3w] [align] [cfm] [hift] [flow]

The reference modules are imported from /root/reference through `oracle/ref_import.py`, loaded
with the synthetic state dicts of `chatterbox_b200/synthetic.py`, and driven through their own
public entry points:
  * T3: `T3.inference` itself (t3.py:203-371).  Its only stochastic call, `torch.multinomial`
    (t3.py:341), is swapped for argmax (greedy) or for the next teacher-forced id while the call
    runs; raw logits are captured with a forward hook on `speech_head`.
  * CFM / HiFT: see the later sections of this file.
Every vector is also compared with the CPU restatement in `oracle/*.py` before it is written, so
a committed fixture certifies reference == restatement at generation time.
"""
from __future__ import annotations

import contextlib
import os
import sys
import time

import torch
from safetensors.torch import save_file

from chatterbox_b200 import synthetic as S
from oracle import ref_import as R
from oracle import t3_oracle as T3O

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
T3_LOGIT_STEPS = [0, 1, 2, 3, 7, 15, 31, 47]
T3_GREEDY_N = 48
T3_FORCED_N = 24
T3_FORCED_STEPS = [0, 5, 23]


@contextlib.contextmanager
def _patched_multinomial(fn):
    orig = torch.multinomial
    torch.multinomial = fn
    try:
        yield
    finally:
        torch.multinomial = orig


def _ref_t3_run(t3, cond_fields, text2, n, forced=None, **kw):
    """Run the reference T3.inference greedy / teacher-forced; return tokens, raw logits per step."""
    from chatterbox.models.t3.modules.cond_enc import T3Cond
    raw, step = [], [0]
    hook = t3.speech_head.register_forward_hook(lambda m, i, o: raw.append(o[:, -1].detach().clone()))

    def pick(probs, num_samples=1):
        i = step[0]
        step[0] += 1
        if forced is not None:
            return forced[:, i:i + 1].clone()
        return probs.argmax(dim=-1, keepdim=True)

    cond = T3Cond(**{k: v.clone() for k, v in cond_fields.items()})
    with _patched_multinomial(pick):
        toks = t3.inference(t3_cond=cond, text_tokens=text2, max_new_tokens=n, **kw)
    hook.remove()
    return toks, torch.stack(raw[:toks.shape[1]])


def make_t3():
    torch.manual_seed(0)
    t0 = time.time()
    sd = S.make_state_dict("t3")
    t3 = R.build_t3()
    t3.load_state_dict(sd)
    print(f"[t3] reference built + synthetic weights loaded in {time.time() - t0:.1f}s")
    cond = S.t3_cond()
    text = S.text_tokens()
    text2 = torch.cat([text, text], 0)
    params = dict(cfg_weight=0.5, temperature=0.8, repetition_penalty=1.2, min_p=0.05, top_p=1.0)
    out = {}

    # conditioning prefix and prefill embeddings: reference vs restatement
    from chatterbox.models.t3.modules.cond_enc import T3Cond
    with torch.inference_mode():
        ref_emb, len_cond = t3.prepare_input_embeds(
            t3_cond=T3Cond(**{k: v.clone() for k, v in cond.items()}), text_tokens=text2,
            speech_tokens=T3O.START_SPEECH * torch.ones_like(text2[:, :1]), cfg_weight=0.5)
    my_emb = T3O.prepare_input_embeds(sd, cond, text2, 0.5)
    assert len_cond == 34
    d = (ref_emb - my_emb[:, :-1]).abs().max().item()
    print(f"[t3] prepare_input_embeds max-abs diff restatement vs reference: {d:.3e}")
    assert d < 1e-4
    out["cond_emb"] = ref_emb[0, :34].clone()
    out["embeds_row1_text0"] = ref_emb[1, 34:36].clone()

    # greedy
    t0 = time.time()
    toks, raw = _ref_t3_run(t3, cond, text2, T3_GREEDY_N, **params)
    print(f"[t3] reference greedy {toks.shape[1]} tokens in {time.time() - t0:.1f}s: {toks[0, :12].tolist()}")
    my_toks, my_raw = T3O.inference(sd, cond, text2, max_new_tokens=T3_GREEDY_N, **params)
    assert torch.equal(toks, my_toks), (toks, my_toks)
    d = (raw - my_raw).abs().max().item()
    print(f"[t3] greedy tokens identical; raw logits max-abs diff {d:.3e}")
    assert d < 2e-3
    # top-2 gap of the processed logits along the greedy path
    gaps = []
    gen = torch.tensor([[T3O.START_SPEECH]])
    for i in range(toks.shape[1]):
        p = T3O.process_logits(raw[i], gen, **params)
        top2 = p.topk(2, dim=-1).values[0]
        gaps.append(float(top2[0] - top2[1]))
        gen = torch.cat([gen, toks[:, i:i + 1]], 1)
    print(f"[t3] min top-2 gap along greedy path: {min(gaps):.4e}")
    out["greedy_tokens"] = toks[0].to(torch.int32)
    out["greedy_logit_steps"] = torch.tensor(T3_LOGIT_STEPS, dtype=torch.int32)
    out["greedy_logits"] = raw[T3_LOGIT_STEPS].clone()
    out["greedy_top2_gaps"] = torch.tensor(gaps)

    # teacher-forced
    g = torch.Generator().manual_seed(77)
    forced = torch.randint(0, 6561, (1, T3_FORCED_N), generator=g)
    toks_f, raw_f = _ref_t3_run(t3, cond, text2, T3_FORCED_N, forced=forced, **params)
    assert torch.equal(toks_f, forced)
    _, my_raw_f = T3O.inference(sd, cond, text2, max_new_tokens=T3_FORCED_N, forced_tokens=forced, **params)
    d = (raw_f - my_raw_f).abs().max().item()
    print(f"[t3] teacher-forced raw logits max-abs diff {d:.3e}")
    assert d < 2e-3
    out["forced_tokens"] = forced[0].to(torch.int32)
    out["forced_logit_steps"] = torch.tensor(T3_FORCED_STEPS, dtype=torch.int32)
    out["forced_logits"] = raw_f[T3_FORCED_STEPS].clone()

    # logits processors alone (HF objects as t3.py:302-304 builds them), non-default top_p
    from transformers.generation.logits_process import (MinPLogitsWarper, RepetitionPenaltyLogitsProcessor,
                                                        TopPLogitsWarper)
    lg = raw[3]
    ids = torch.cat([torch.tensor([[T3O.START_SPEECH]]), toks[:, :3]], 1)
    c = lg[0:1] + 0.5 * (lg[0:1] - lg[1:2])
    c = c / 0.8
    c = RepetitionPenaltyLogitsProcessor(penalty=1.3)(ids, c)
    c = MinPLogitsWarper(min_p=0.02)(None, c)
    c = TopPLogitsWarper(top_p=0.9)(None, c)
    mine = T3O.process_logits(lg, ids, cfg_weight=0.5, temperature=0.8, repetition_penalty=1.3, min_p=0.02, top_p=0.9)
    assert torch.equal(torch.isinf(c), torch.isinf(mine))
    fin = ~torch.isinf(c)
    assert (c[fin] - mine[fin]).abs().max() < 1e-5
    print(f"[t3] processors: {int(fin.sum())} tokens kept of {c.numel()}")
    out["proc_in_logits"] = lg.clone()
    out["proc_ids"] = ids[0].to(torch.int32)
    out["proc_out"] = c.clone()

    save_file({k: v.contiguous() for k, v in out.items()}, os.path.join(GOLD, "t3_golden.safetensors"))
    print("[t3] wrote tests/golden/t3_golden.safetensors")


T3W_N = 256                      # the bench workload: 256 speech tokens (BASELINE configs 1-2)
T3W_FULL_STEPS = [0, 1, 7, 31, 63, 127, 191, 255]
T3W_SAMPLE_COLS = 384            # vocabulary columns kept for EVERY step (full rows only at T3W_FULL_STEPS)


def make_t3w():
    """The bench's own configuration on UN-ROUNDED fp32 weights (`make_state_dict("t3", bf16_exact=False)`, what a
    real t3_cfg.safetensors holds): 30 layers, 256 greedy + 256 teacher-forced steps through the reference
    `T3.inference`.  Full logit rows are kept at T3W_FULL_STEPS, a fixed sample of vocabulary columns at every
    step (a full 256 x 2 x 8194 fp32 dump would be 16.8 MB)."""
    torch.manual_seed(0)
    t0 = time.time()
    sd = S.make_state_dict("t3", bf16_exact=False)
    t3 = R.build_t3()
    t3.load_state_dict(sd)
    print(f"[t3w] reference built + fp32 synthetic weights loaded in {time.time() - t0:.1f}s")
    cond = S.t3_cond()
    text = S.text_tokens()
    text2 = torch.cat([text, text], 0)
    params = dict(cfg_weight=0.5, temperature=0.8, repetition_penalty=1.2, min_p=0.05, top_p=1.0)
    g = torch.Generator().manual_seed(4242)
    cols = torch.randperm(8194, generator=g)[:T3W_SAMPLE_COLS].sort().values
    out = {"sample_cols": cols.to(torch.int32), "full_steps": torch.tensor(T3W_FULL_STEPS, dtype=torch.int32)}

    t0 = time.time()
    toks, raw = _ref_t3_run(t3, cond, text2, T3W_N, **params)
    print(f"[t3w] reference greedy {toks.shape[1]} tokens in {time.time() - t0:.1f}s: {toks[0, :12].tolist()}")
    assert toks.shape[1] == T3W_N, "synthetic weights must not emit EOS inside the bench workload"
    gaps = []
    gen = torch.tensor([[T3O.START_SPEECH]])
    for i in range(toks.shape[1]):
        p = T3O.process_logits(raw[i], gen, **params)
        top2 = p.topk(2, dim=-1).values[0]
        gaps.append(float(top2[0] - top2[1]))
        gen = torch.cat([gen, toks[:, i:i + 1]], 1)
    print(f"[t3w] min / median top-2 gap along the greedy path: {min(gaps):.4e} / {sorted(gaps)[len(gaps) // 2]:.4e}")
    out["greedy_tokens"] = toks[0].to(torch.int32)
    out["greedy_top2_gaps"] = torch.tensor(gaps)
    out["greedy_logits_full"] = raw[T3W_FULL_STEPS].clone()
    out["greedy_logits_cols"] = raw[:, :, cols].clone()
    # the restatement on the same weights, every 32nd step (it is what the GPU box can run)
    my_toks, my_raw = T3O.inference(sd, cond, text2, max_new_tokens=64, forced_tokens=toks, **params)
    d = (raw[:64] - my_raw).abs().max().item()
    print(f"[t3w] restatement vs reference over the first 64 forced steps: raw logits max-abs {d:.3e}")
    assert d < 2e-3

    g = torch.Generator().manual_seed(78)
    forced = torch.randint(0, 6561, (1, T3W_N), generator=g)
    t0 = time.time()
    toks_f, raw_f = _ref_t3_run(t3, cond, text2, T3W_N, forced=forced, **params)
    print(f"[t3w] reference teacher-forced {toks_f.shape[1]} steps in {time.time() - t0:.1f}s")
    assert torch.equal(toks_f, forced)
    out["forced_tokens"] = forced[0].to(torch.int32)
    out["forced_logits_full"] = raw_f[T3W_FULL_STEPS].clone()
    out["forced_logits_cols"] = raw_f[:, :, cols].clone()
    save_file({k: v.contiguous() for k, v in out.items()}, os.path.join(GOLD, "t3_fp32w_golden.safetensors"))
    print("[t3w] wrote tests/golden/t3_fp32w_golden.safetensors")


def _ref_analyzer(S_text, eos):
    """The UNMODIFIED reference class on a stand-in transformer (its constructor only installs a forward hook)."""
    import types
    R.install()
    from chatterbox.models.t3.inference.alignment_stream_analyzer import AlignmentStreamAnalyzer

    class _Attn(torch.nn.Module):
        def forward(self, *a, **k):
            return None
    tf = types.SimpleNamespace(layers=[types.SimpleNamespace(self_attn=_Attn()) for _ in range(10)])
    return AlignmentStreamAnalyzer(tf, None, (34, 34 + S_text), alignment_layer_idx=9, eos_idx=eos)


def align_script(S_text: int, n_steps: int, seed: int, mode: str) -> torch.Tensor:
    """Synthetic head-mean attention rows over the text columns, (n_steps + 1, S): two prompt rows' worth is rows 0-1 of
    the first call, then one row per decode step.  mode "walk": a peak that walks monotonically through the text and
    then dwells on the last tokens (-> complete, long tail); "repeat": walks to the end and jumps back (-> repetition);
    "noise": near-uniform rows (-> false start forever, EOS suppressed)."""
    g = torch.Generator().manual_seed(seed)
    rows = torch.rand(n_steps + 1, S_text, generator=g) * 0.02
    if mode == "noise":
        return rows
    for t in range(n_steps + 1):
        if mode == "walk":
            c = min(S_text - 1, t // 2)
        else:
            c = min(S_text - 1, t // 2) if t < 2 * S_text + 6 else 3 + (t % 5)
        rows[t, c] = 0.9
    return rows


def make_align():
    """AlignmentStreamAnalyzer.step (alignment_stream_analyzer.py:89-154): the reference class and the O(1)-state
    restatement side by side, on crafted attention rows that reach every branch and on the layer-9 attention the
    reference model itself produces; decisions per step committed as golden."""
    eos = T3O.STOP_SPEECH
    out = {}
    for name, (S_text, n, mode) in {"walk": (12, 80, "walk"), "repeat": (16, 90, "repeat"), "noise": (20, 30, "noise")}.items():
        rows = align_script(S_text, n, 5, mode)
        ref, mine = _ref_analyzer(S_text, eos), T3O.AlignmentOracle(S_text, eos)
        dec = []
        for k in range(n):
            chunk = rows[0:2] if k == 0 else rows[k + 1:k + 2]
            full = torch.zeros(chunk.shape[0] if k else 34 + S_text + 2, 34 + S_text + 2 + k)
            if k == 0:
                full[34 + S_text:, 34:34 + S_text] = chunk
            else:
                full[:, 34:34 + S_text] = chunk
            ref.last_aligned_attn = full
            lg = torch.randn(2, 1, 8194, generator=torch.Generator().manual_seed(k))
            a = ref.step(lg.clone())
            b = mine.step(lg.clone(), chunk)
            assert torch.equal(a, b), (name, k)
            # "forced": every other logit is -2**15 (the reference's suppression runs AFTER the forcing and overrides the
            # EOS logit itself when the current position is early: alignment_stream_analyzer.py:141-151)
            forced = bool(a[0, 0, 0] == -2 ** 15 and a[0, 0, 1] == -2 ** 15)
            sup = bool(a[0, 0, eos] == -2 ** 15)
            dec.append([int(ref.text_position), int(ref.started), int(ref.complete), int(forced), int(sup)])
            assert (int(mine.text_position), int(mine.started), int(mine.complete)) == tuple(dec[-1][:3])
        dec = torch.tensor(dec, dtype=torch.int32)
        print(f"[align] {name}: {n} steps, reference == restatement; started at {ref.started_at}, completed at {ref.completed_at}, "
              f"forced EOS at steps {torch.nonzero(dec[:, 3]).flatten()[:3].tolist()}, suppressed {int(dec[:, 4].sum())} steps")
        out[f"{name}_rows"] = rows
        out[f"{name}_decisions"] = dec
    assert int(out["walk_decisions"][:, 3].sum()) > 0 and int(out["repeat_decisions"][:, 3].sum()) > 0
    # layer-9 attention of the reference model itself (teacher-forced, 2-layer... no: the full 30-layer synthetic model)
    torch.manual_seed(0)
    sd = S.make_state_dict("t3")
    t3 = R.build_t3()
    t3.load_state_dict(sd)
    cond, text = S.t3_cond(), S.text_tokens()
    text2 = torch.cat([text, text], 0)
    S_text = text.shape[1]
    params = dict(cfg_weight=0.5, temperature=0.8, repetition_penalty=1.2, min_p=0.05, top_p=1.0)
    # transformers 5.5 keeps SDPA under output_attentions=True and returns no weights (4.46.3, the reference's pin, falls
    # back to the eager path): select the eager path explicitly for this capture
    t3.tfmr.config._attn_implementation = "eager"
    maps = []
    hk = t3.tfmr.layers[9].self_attn.register_forward_hook(lambda m, i, o: maps.append(o[1][0].mean(0).detach().clone()))
    forced = torch.randint(0, 6561, (1, 12), generator=torch.Generator().manual_seed(91))
    _ref_t3_run(t3, cond, text2, 12, forced=forced, **params)
    hk.remove()
    assert len(maps) >= 12 and maps[0].shape[0] == 34 + S_text + 2, (len(maps), maps[0].shape)
    rows = [maps[0][34 + S_text:, 34:34 + S_text]] + [m[:, 34:34 + S_text] for m in maps[1:12]]
    _, _, mo = T3O.inference(sd, cond, text2, max_new_tokens=12, forced_tokens=forced, return_model=True,
                             attn_probe_layer=9, **params)
    mine_rows = mo.attn_rows
    for k, (a, b) in enumerate(zip(rows, mine_rows)):
        assert (a - b).abs().max() < 1e-5, k
    out["model_rows0"] = rows[0].clone()
    out["model_rows"] = torch.cat(rows[1:], 0)
    out["model_forced"] = forced[0].to(torch.int32)
    print(f"[align] layer-9 head-mean attention rows of the reference model == restatement over 12 steps "
          f"(row sums over the text columns: {float(rows[0][-1].sum()):.3f} ... {float(rows[-1].sum()):.3f})")
    save_file({k: v.contiguous() for k, v in out.items()}, os.path.join(GOLD, "t3_align_golden.safetensors"))
    print("[align] wrote tests/golden/t3_align_golden.safetensors")


def make_cfm():
    """Reference ConditionalDecoder / CausalConditionalCFM vs oracle/cfm_oracle.py; golden outputs."""
    from oracle import cfm_oracle as CO
    torch.manual_seed(0)
    sd_flow = S.make_state_dict("flow")
    flow = R.build_flow()
    missing, unexpected = flow.load_state_dict(sd_flow, strict=False)
    assert not unexpected and all(k.endswith("rand_noise") for k in missing), (missing, unexpected)
    cfm = flow.decoder
    est = cfm.estimator
    pfx = "decoder.estimator."
    out = {}
    # (1) one estimator call, batch 2 (CFG packing of flow_matching.py:95-108), L = 96
    L = 96
    inp = S.cfm_inputs(L)
    t = torch.tensor([0.3, 0.3])
    x_in = torch.cat([inp["z"], inp["z"]], 0)
    mu_in = torch.cat([inp["mu"], torch.zeros_like(inp["mu"])], 0)
    sp_in = torch.cat([inp["spks"], torch.zeros_like(inp["spks"])], 0)
    c_in = torch.cat([inp["cond"], torch.zeros_like(inp["cond"])], 0)
    mask = torch.ones(2, 1, L)
    with torch.inference_mode():
        ref = est(x_in, mask, mu_in, t, sp_in, c_in)
        mine = CO.estimator(sd_flow, x_in, mask, mu_in, t, sp_in, c_in, pfx)
    d = (ref - mine).abs().max().item()
    print(f"[cfm] estimator L={L}: max-abs restatement vs reference {d:.3e} (ref rms {ref.pow(2).mean().sqrt():.3f})")
    assert d < 1e-3
    out["est_t0.3_L96"] = ref.clone()
    # (2) the 10-step Euler solve through CausalConditionalCFM.forward with the injected noise, L = 96 and 130
    for L in (96, 130):
        inp = S.cfm_inputs(L, seed=6 + L)
        cfm.rand_noise = S.cfm_noise()
        t0 = time.time()
        with torch.inference_mode():
            ref, _ = cfm(mu=inp["mu"], mask=torch.ones(1, 1, L), n_timesteps=10, spks=inp["spks"], cond=inp["cond"])
            mine = CO.solve_euler(sd_flow, inp["z"], inp["mu"], inp["spks"], inp["cond"], pfx=pfx)
        rel = ((ref - mine).norm() / ref.norm()).item()
        print(f"[cfm] solve_euler L={L}: rel-L2 restatement vs reference {rel:.3e} ({time.time() - t0:.1f}s)")
        assert rel < 1e-4
        out[f"mel_L{L}"] = ref.clone()
    out["t_span"] = CO.t_schedule(10)
    save_file({k: v.contiguous() for k, v in out.items()}, os.path.join(GOLD, "cfm_golden.safetensors"))
    print("[cfm] wrote tests/golden/cfm_golden.safetensors")


def make_hift():
    """Reference HiFTGenerator / ConvRNNF0Predictor vs oracle/hift_oracle.py; golden outputs."""
    import numpy as np
    from torch.distributions.uniform import Uniform
    from oracle import hift_oracle as HO
    sd = S.make_state_dict("hift")
    hift = R.build_hift()
    missing, unexpected = hift.load_state_dict(sd, strict=False)
    assert not unexpected and all("stft_window" in k for k in missing), (missing, unexpected)
    w = HO.fold_weight_norm(sd)
    out = {}
    for L in (12, 40):
        mel = S.mel_input(L, seed=7 + L)
        with torch.inference_mode():
            f0_ref = hift.f0_predictor(mel)
            f0 = HO.f0_predictor(w, mel)
        d = (f0_ref - f0).abs().max().item()
        print(f"[hift] L={L} f0 range {f0_ref.min():.1f}..{f0_ref.max():.1f} Hz; restatement max-abs {d:.3e}")
        assert d < 1e-2
        # source: replicate the reference's RNG draws (hifigan.py:212-214,226) under one seed
        torch.manual_seed(1000 + L)
        with torch.inference_mode():
            s_ref, _, _ = hift.m_source(hift.f0_upsamp(f0_ref[:, None]).transpose(1, 2))
        s_ref = s_ref.transpose(1, 2)
        torch.manual_seed(1000 + L)
        pv = Uniform(low=-np.pi, high=np.pi).sample(sample_shape=(1, 9, 1))
        pv[:, 0, :] = 0
        nz = torch.randn(1, 9, 480 * L)
        s_mine = HO.sine_source(w, f0_ref, pv, nz)
        d = (s_ref - s_mine).abs().max().item()
        print(f"[hift] L={L} source max-abs restatement vs reference {d:.3e} (rms {s_ref.pow(2).mean().sqrt():.3f})")
        assert d < 1e-4
        with torch.inference_mode():
            wav_ref = hift.decode(x=mel, s=s_ref)
            wav = HO.decode(w, mel, s_ref)
        rel = ((wav_ref - wav).norm() / wav_ref.norm()).item()
        print(f"[hift] L={L} decode rel-L2 restatement vs reference {rel:.3e}; wav rms {wav_ref.pow(2).mean().sqrt():.4f}, "
              f"clamped {float((wav_ref.abs() >= 0.99).float().mean()):.3f}")
        assert rel < 1e-4
        out[f"f0_L{L}"] = f0_ref.clone()
        out[f"s_L{L}"] = s_ref.clone()
        out[f"phase_L{L}"] = pv.clone()
        out[f"wav_L{L}"] = wav_ref.clone()
        if L == 12:
            out["noise_L12"] = nz.clone()
            out["stage0_L12"] = HO.decode(w, mel, s_ref, stop_after="stage0")
            out["stage2_L12"] = HO.decode(w, mel, s_ref, stop_after="stage2")
    save_file({k: v.contiguous() for k, v in out.items()}, os.path.join(GOLD, "hift_golden.safetensors"))
    print("[hift] wrote tests/golden/hift_golden.safetensors")


def make_flow():
    """Reference CausalMaskedDiffWithXvec.inference: captures the CFM inputs the front-end produces and the
    final mel; checks oracle/flow_oracle.py (plain torch) against them."""
    from oracle import flow_oracle as FF
    sd = S.make_state_dict("flow")
    flow = R.build_flow()
    flow.load_state_dict(sd, strict=False)
    flow.decoder.rand_noise = S.cfm_noise()
    ref = S.s3gen_ref(seed=4, prompt_len=20)
    tok = S.speech_tokens(30, seed=5)
    cap = {}
    hk = flow.decoder.register_forward_pre_hook(lambda m, a, kw: cap.update({k: v.clone() for k, v in kw.items() if torch.is_tensor(v)}),
                                                with_kwargs=True)
    with torch.inference_mode():
        feat, _ = flow.inference(token=tok, token_len=torch.tensor([30]), prompt_token=ref["prompt_token"],
                                 prompt_token_len=ref["prompt_token_len"], prompt_feat=ref["prompt_feat"],
                                 prompt_feat_len=None, embedding=ref["embedding"], finalize=True)
        mu, spks, cond, l1 = FF.flow_inputs(sd, tok, ref["prompt_token"], ref["prompt_feat"], ref["embedding"])
    hk.remove()
    for name, mine in (("mu", mu), ("spks", spks), ("cond", cond)):
        d = (cap[name] - mine).abs().max().item()
        print(f"[flow] front-end {name}: max-abs vs reference {d:.3e} (rms {cap[name].pow(2).mean().sqrt():.3f})")
        assert d < 1e-3
    assert l1 == 40 and feat.shape == (1, 80, 60)
    out = {"mu": cap["mu"], "spks": cap["spks"], "cond": cap["cond"], "feat": feat.clone()}
    save_file({k: v.contiguous() for k, v in out.items()}, os.path.join(GOLD, "flow_golden.safetensors"))
    print("[flow] wrote tests/golden/flow_golden.safetensors")


if __name__ == "__main__":
    which = sys.argv[1:] or ["t3", "cfm", "hift", "flow"]
    torch.set_num_threads(os.cpu_count() or 8)
    if "t3" in which:
        make_t3()
    if "t3w" in which:
        make_t3w()
    if "align" in which:
        make_align()
    if "cfm" in which and "make_cfm" in globals():
        globals()["make_cfm"]()
    if "hift" in which and "make_hift" in globals():
        globals()["make_hift"]()
    if "flow" in which:
        make_flow()
