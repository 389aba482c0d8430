"""The T3 CPU restatement (oracle/t3_oracle.py) against vectors the UNMODIFIED reference produced
(tests/golden/t3_golden.safetensors, written by oracle/make_golden.py in the build container)."""
import os

import pytest
import torch
from safetensors.torch import load_file

from chatterbox_b200 import synthetic as S
from oracle import t3_oracle as T3O

GOLD = load_file(os.path.join(os.path.dirname(__file__), "golden", "t3_golden.safetensors"))
PARAMS = dict(cfg_weight=0.5, temperature=0.8, repetition_penalty=1.2, min_p=0.05, top_p=1.0)


@pytest.fixture(scope="module")
def sd():
    return S.make_state_dict("t3")


def test_rope_known_answers():
    inv = T3O.rope_inv_freq()
    assert inv.shape == (32,)
    assert abs(float(inv[0]) - 1.0) < 1e-7                       # high-frequency band untouched
    base_last = 1.0 / (500000.0 ** (62 / 64))
    assert abs(float(inv[-1]) / (base_last / 8.0) - 1.0) < 1e-5   # low-frequency band divided by factor 8
    cos, sin = T3O.rope_cos_sin(torch.tensor([0, 5]))
    assert torch.allclose(cos[0], torch.ones(64)) and torch.allclose(sin[0], torch.zeros(64))
    assert torch.equal(cos[1, :32], cos[1, 32:])                  # rotate_half (split-halves) layout


def test_cond_prefix_matches_reference(sd):
    text2 = torch.cat([S.text_tokens()] * 2, 0)
    emb = T3O.prepare_input_embeds(sd, S.t3_cond(), text2, 0.5)
    assert emb.shape == (2, 34 + 66 + 2, 1024)
    assert (emb[0, :34] - GOLD["cond_emb"]).abs().max() < 1e-5
    assert (emb[1, 34:36] - GOLD["embeds_row1_text0"]).abs().max() < 1e-6   # uncond row: pos-emb only
    assert torch.equal(emb[:, -1], emb[:, -2])                              # doubled BOS at learned pos 0


def test_processors_match_hf(sd):
    out = T3O.process_logits(GOLD["proc_in_logits"], GOLD["proc_ids"].long()[None], cfg_weight=0.5,
                             temperature=0.8, repetition_penalty=1.3, min_p=0.02, top_p=0.9)
    ref = GOLD["proc_out"]
    assert torch.equal(torch.isinf(out), torch.isinf(ref))
    fin = ~torch.isinf(ref)
    assert (out[fin] - ref[fin]).abs().max() < 1e-5


def test_greedy_and_logits_match_reference(sd):
    n = 8
    text2 = torch.cat([S.text_tokens()] * 2, 0)
    toks, raw = T3O.inference(sd, S.t3_cond(), text2, max_new_tokens=n, **PARAMS)
    assert torch.equal(toks[0].to(torch.int32), GOLD["greedy_tokens"][:n])
    for j, st in enumerate(GOLD["greedy_logit_steps"].tolist()):
        if st < n:
            assert (raw[st] - GOLD["greedy_logits"][j]).abs().max() < 1e-4


def test_teacher_forced_logits_match_reference(sd):
    n = 6
    text2 = torch.cat([S.text_tokens()] * 2, 0)
    forced = GOLD["forced_tokens"].long()[None]
    toks, raw = T3O.inference(sd, S.t3_cond(), text2, max_new_tokens=n, forced_tokens=forced, **PARAMS)
    for j, st in enumerate(GOLD["forced_logit_steps"].tolist()):
        if st < n:
            assert (raw[st] - GOLD["forced_logits"][j]).abs().max() < 1e-4


# ---- AlignmentStreamAnalyzer restatement vs the reference class's own decisions (oracle/make_golden.py align) ----
ALIGN = load_file(os.path.join(os.path.dirname(__file__), "golden", "t3_align_golden.safetensors"))


@pytest.mark.parametrize("name", ["walk", "repeat", "noise"])
def test_alignment_oracle_reproduces_reference_decisions(name):
    rows, dec = ALIGN[f"{name}_rows"], ALIGN[f"{name}_decisions"]
    S_text = rows.shape[1]
    an = T3O.AlignmentOracle(S_text)
    for k in range(dec.shape[0]):
        chunk = rows[0:2] if k == 0 else rows[k + 1:k + 2]
        out = an.step(torch.zeros(2, 1, 8194), chunk)
        forced = bool(out[0, 0, 0] == -2 ** 15)
        sup = bool(out[0, 0, T3O.STOP_SPEECH] == -2 ** 15)
        assert [an.text_position, int(an.started), int(an.complete), int(forced), int(sup)] == dec[k].tolist(), (name, k)
    if name != "noise":
        assert int(dec[:, 3].sum()) > 0          # the script reaches the forced-EOS branch


def test_alignment_rows_of_the_oracle_model_match_the_reference_model(sd):
    text2 = torch.cat([S.text_tokens()] * 2, 0)
    forced = ALIGN["model_forced"].long()[None]
    _, _, m = T3O.inference(sd, S.t3_cond(), text2, max_new_tokens=3, forced_tokens=forced, return_model=True,
                            attn_probe_layer=9, **PARAMS)
    assert (m.attn_rows[0] - ALIGN["model_rows0"]).abs().max() < 1e-5
    for k in range(1, 3):
        assert (m.attn_rows[k][0] - ALIGN["model_rows"][k - 1]).abs().max() < 1e-5
