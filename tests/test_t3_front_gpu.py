"""GPU parity of the T3 conditioning front-end kernels (csrc/t3_front.cu through cbx_t3_front) against the oracle's
`prepare_input_embeds` (oracle/t3_oracle.py, itself pinned against the reference's T3.prepare_input_embeds and the golden
`cond_emb` of the unmodified reference: tests/test_oracle_t3.py) — fp32 on both sides."""
import os

import pytest
import torch
from safetensors.torch import load_file

from chatterbox_b200 import synthetic as S

pytestmark = pytest.mark.gpu
GOLD = load_file(os.path.join(os.path.dirname(__file__), "golden", "t3_golden.safetensors"))


@pytest.fixture(scope="module")
def t3():
    from chatterbox_b200.t3 import T3
    return T3(S.make_state_dict("t3", n_layers=1), "cuda", max_pos=256, n_layers=1)


def _cond(seed=2, prompt_len=150, emo=0.5):
    from chatterbox_b200.t3 import T3Cond
    return T3Cond(**{k: v.clone() for k, v in S.t3_cond(seed=seed, prompt_len=prompt_len, exaggeration=emo).items()}).to(device="cuda")


@pytest.mark.parametrize("n_text,cfg,prompt_len,emo", [(64, 0.5, 150, 0.5), (3, 0.5, 7, 0.9), (40, 0.0, 150, 0.25), (5, 0.5, 256, 0.5)])
def test_prefill_embeds_match_the_oracle(t3, n_text, cfg, prompt_len, emo):
    from oracle import t3_oracle as T3O
    sd = S.make_state_dict("t3", n_layers=1)
    text = S.text_tokens(n_text, seed=n_text)
    tt = torch.cat([text, text], 0) if cfg > 0 else text
    want = T3O.prepare_input_embeds(sd, S.t3_cond(prompt_len=prompt_len, exaggeration=emo), tt, cfg)
    got, len_cond = t3._prefill_embeds(_cond(prompt_len=prompt_len, emo=emo), tt, cfg, 2 if cfg > 0 else 1)
    assert len_cond == 34 and got.shape == want.shape
    err = (got.cpu() - want).abs().max().item()
    print(f"N = {n_text + 2}, cfg {cfg}, {prompt_len} prompt tokens: prefill embeddings max-abs vs oracle {err:.3e} (|x| max {float(want.abs().max()):.2f})")
    assert err < 2e-5
    if cfg > 0:
        assert torch.equal(got[:, -1], got[:, -2])                      # doubled BOS
        assert not torch.equal(got[0, 34:34 + 2], got[1, 34:34 + 2])     # unconditional row: position embedding only


def test_conditioning_prefix_matches_the_reference_golden(t3):
    ce = t3.prepare_conditioning(_cond()).cpu()
    assert ce.shape == (1, 34, 1024)
    err = (ce[0] - GOLD["cond_emb"]).abs().max().item()
    print(f"conditioning prefix (speaker | perceiver | emotion) max-abs vs the reference's own output: {err:.3e}")
    assert err < 2e-5


def test_precomputed_prompt_embeddings_and_no_prompt(t3):
    """T3Cond.cond_prompt_speech_emb given directly (cond_enc.py:78-80), and a T3Cond without prompt tokens (prefix = speaker | emotion)."""
    from chatterbox_b200.t3 import T3Cond
    c = _cond(prompt_len=20)
    text = S.text_tokens(4).to("cuda")
    a, _ = t3._prefill_embeds(c, text, 0.0, 1)
    pe = t3.fe["speech_emb.weight"][c.cond_prompt_speech_tokens] + t3.fe["speech_pos_emb.emb.weight"][:20][None]
    c2 = T3Cond(speaker_emb=c.speaker_emb, cond_prompt_speech_tokens=c.cond_prompt_speech_tokens, cond_prompt_speech_emb=pe,
                emotion_adv=c.emotion_adv)
    b, _ = t3._prefill_embeds(c2, text, 0.0, 1)
    assert torch.equal(a, b)
    c3 = T3Cond(speaker_emb=c.speaker_emb, emotion_adv=c.emotion_adv)
    d, len_cond = t3._prefill_embeds(c3, text, 0.0, 1)
    assert len_cond == 2 and d.shape[1] == 2 + 6 + 1 and torch.equal(d[:, 0], a[:, 0]) and torch.equal(d[:, 1], a[:, 33])
