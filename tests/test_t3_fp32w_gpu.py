"""T3 parity on the configuration the bench actually runs: 30 layers, bf16 KV slab, 256 speech tokens, and weights
that are NOT bf16-representable (`make_state_dict("t3", bf16_exact=False)` — what a real fp32 t3_cfg.safetensors
holds).  The reference (tests/golden/t3_fp32w_golden.safetensors, written by `python -m oracle.make_golden t3w` from the
unmodified `T3.inference`) computes in fp32 on the fp32 weights; the build stores them as bf16, so these tests measure
the build's full quantisation error (weights + KV) instead of hiding it.

Tolerances.  BASELINE.json north_star states the bf16-mode tolerance as "e.g. max-abs 2e-2".  That figure holds on
bf16-representable weights (tests/test_t3_gpu.py: 4.7e-3 measured).  On un-rounded fp32 weights the measured per-step
max-abs over 256 teacher-forced steps is 1.4e-2 mean, 2.0e-2 at the worst sampled step and 2.3e-2 at the worst full row
(|logit| max 3.1): the bf16 STORAGE of the weights alone puts the error at the 2e-2 mark.  That this is storage rounding
and not kernel arithmetic is checked directly: against the oracle run with the same roundings (`weight_round=True,
kv_round=True`, fp32 math) the kernel agrees to 1.8e-3 (measured; what remains are K / V elements that sit on a bf16
rounding boundary and land one bf16 ulp apart when kernel and oracle differ in the last fp32 bit), an order of magnitude
below the storage error.  So: tolerance vs the fp32 reference stated here as 3e-2 (measured 2.3e-2 — above north_star's
example figure, reported, not hidden); vs the equally-rounded oracle 5e-3 (measured 1.8e-3); greedy choices
must agree with the reference wherever the reference's own top-2 gap exceeds what a 2e-2 logit error can move.
"""
import os

import pytest
import torch
from safetensors.torch import load_file

from chatterbox_b200 import synthetic as S

GOLD = load_file(os.path.join(os.path.dirname(__file__), "golden", "t3_fp32w_golden.safetensors"))
PARAMS = dict(cfg_weight=0.5, temperature=0.8, repetition_penalty=1.2, min_p=0.05, top_p=1.0)
TOL_BF16 = 2e-2          # north_star's example figure: used for the "clear gap" criterion of the greedy test
TOL_FP32W = 3e-2         # stated tolerance vs the fp32 reference on un-rounded weights (measured 2.3e-2)
TOL_ROUNDED = 5e-3       # vs the oracle with the build's own storage roundings (measured 1.8e-3)
N = 256


def test_oracle_matches_reference_on_fp32_weights():
    """CPU: the restatement on the un-rounded weights against the reference's logits (first steps; the generator
    script checked the first 64)."""
    from oracle import t3_oracle as T3O
    sd = S.make_state_dict("t3", bf16_exact=False)
    text2 = torch.cat([S.text_tokens()] * 2, 0)
    forced = GOLD["forced_tokens"].long()[None]
    _, raw = T3O.inference(sd, S.t3_cond(), text2, max_new_tokens=2, forced_tokens=forced, **PARAMS)
    cols = GOLD["sample_cols"].long()
    assert (raw[0] - GOLD["forced_logits_full"][0]).abs().max() < 1e-4
    assert (raw[1][:, cols] - GOLD["forced_logits_cols"][1]).abs().max() < 1e-4
    w = sd["tfmr.layers.0.mlp.down_proj.weight"]
    assert not torch.equal(w, w.to(torch.bfloat16).float())       # really un-rounded


@pytest.fixture(scope="module")
def model():
    from chatterbox_b200.t3 import T3
    return T3(S.make_state_dict("t3", bf16_exact=False), "cuda", max_pos=512, kv_fp32=False)


def _cond():
    from chatterbox_b200.t3 import T3Cond
    return T3Cond(**{k: v.clone() for k, v in S.t3_cond().items()}).to(device="cuda")


@pytest.mark.gpu
def test_teacher_forced_256_steps_bf16_kv_fp32_weights(model):
    text2 = torch.cat([S.text_tokens()] * 2, 0)
    forced = GOLD["forced_tokens"].long()[None]
    toks = model.inference(t3_cond=_cond(), text_tokens=text2, max_new_tokens=N, forced_tokens=forced,
                           return_logits=True, do_sample=False, stop_on_eos=False, **PARAMS)
    assert torch.equal(toks.cpu(), forced)
    got = model.last_logits.cpu()
    cols = GOLD["sample_cols"].long()
    per_step = (got[:, :, cols] - GOLD["forced_logits_cols"]).abs().amax(dim=(1, 2))
    full = [(got[st] - GOLD["forced_logits_full"][j]).abs().max().item() for j, st in enumerate(GOLD["full_steps"].tolist())]
    scale = GOLD["forced_logits_full"].abs().max().item()
    print(f"teacher-forced, 30 layers, bf16 KV, fp32 (un-rounded) weights: per-step max-abs over {cols.numel()} sampled "
          f"columns: max {per_step.max():.3e} (step {int(per_step.argmax())}), mean {per_step.mean():.3e}; full rows at "
          f"steps {GOLD['full_steps'].tolist()}: {['%.2e' % v for v in full]}; |logit| max {scale:.2f}")
    assert per_step.max() < TOL_FP32W and max(full) < TOL_FP32W
    # the same steps against the oracle WITH the build's storage roundings (bf16 weights, bf16 K/V, fp32 math)
    from oracle import t3_oracle as T3O
    _, rr = T3O.inference(S.make_state_dict("t3", bf16_exact=False), S.t3_cond(), text2, max_new_tokens=3, forced_tokens=forced,
                          weight_round=True, kv_round=True, **PARAMS)
    e_r = (got[:3] - rr).abs().max().item()
    print(f"same build vs the oracle with bf16-rounded weights and K/V (fp32 math), steps 0-2: max-abs {e_r:.3e}")
    assert e_r < TOL_ROUNDED


@pytest.mark.gpu
def test_greedy_256_agrees_where_reference_gap_is_clear(model):
    from oracle import t3_oracle as T3O
    text2 = torch.cat([S.text_tokens()] * 2, 0)
    ref_t = GOLD["greedy_tokens"].long()[None]
    gaps = GOLD["greedy_top2_gaps"]
    # (a) teacher-forced along the reference's greedy path: the build's own arg-max at every step
    model.inference(t3_cond=_cond(), text_tokens=text2, max_new_tokens=N, forced_tokens=ref_t, return_logits=True,
                    do_sample=False, stop_on_eos=False, **PARAMS)
    got = model.last_logits.cpu()
    gen = torch.tensor([[T3O.START_SPEECH]])
    agree, clear, flips = 0, 0, []
    for i in range(N):
        pick = int(T3O.process_logits(got[i], gen, **PARAMS).argmax())
        ok = pick == int(ref_t[0, i])
        agree += ok
        # a raw-logit error e per row moves a processed logit by <= (1 + 2 w) e / T = 2.5 e, a gap of two by 5 e
        if float(gaps[i]) > 5 * TOL_BF16:
            clear += 1
            assert ok, (i, float(gaps[i]))
        elif not ok:
            flips.append((i, float(gaps[i])))
        gen = torch.cat([gen, ref_t[:, i:i + 1]], 1)
    # (b) free-running greedy: identical until the first near-tie
    toks = model.inference(t3_cond=_cond(), text_tokens=text2, max_new_tokens=N, do_sample=False, stop_on_eos=False, **PARAMS)
    same = (toks[0].cpu() == ref_t[0]).long()
    first_div = int(same.argmin()) if int(same.min()) == 0 else N
    print(f"greedy, fp32 weights vs bf16 build: arg-max agrees at {agree}/{N} steps along the reference path "
          f"({clear} steps with a clear gap, all agree); flips at near-ties: {flips[:6]}; free-running sequence identical "
          f"for the first {first_div} tokens")
    assert agree >= N - 8
