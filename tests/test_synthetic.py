"""Synthetic state dicts carry exactly the reference's state-dict keys and shapes."""
import json
import os

import pytest
import torch

from chatterbox_b200 import synthetic as S

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_manifest.json")


@pytest.mark.parametrize("family,spec", [("t3", S.t3_spec), ("flow", S.flow_spec), ("hift", S.hift_spec)])
def test_spec_matches_reference_manifest(family, spec):
    man = json.load(open(GOLD))[family]
    mine = {k: list(v) for k, v in spec().items()}
    assert set(mine) == set(man), (sorted(set(man) - set(mine))[:5], sorted(set(mine) - set(man))[:5])
    for k in man:
        assert mine[k] == man[k], (k, mine[k], man[k])


def test_init_is_a_function_of_key_and_seed_only():
    a = S.init_tensor("t3", "tfmr.layers.3.mlp.up_proj.weight", (8, 16), 7)
    b = S.init_tensor("t3", "tfmr.layers.3.mlp.up_proj.weight", (8, 16), 7)
    c = S.init_tensor("t3", "tfmr.layers.3.mlp.up_proj.weight", (8, 16), 8)
    assert torch.equal(a, b) and not torch.equal(a, c)
    # platform-stability pin: first values of a known tensor
    v = S.init_tensor("hift", "conv_post.bias", (18,), 1234)
    assert v.shape == (18,) and torch.isfinite(v).all()


def test_inputs():
    t = S.text_tokens()
    assert t.shape == (1, 66) and t[0, 0] == 255 and t[0, -1] == 0
    assert (t[0, 1:-1] != 255).all() and (t[0, 1:-1] >= 1).all() and (t[0, 1:-1] <= 703).all()
    c = S.t3_cond()
    assert c["cond_prompt_speech_tokens"].shape == (1, 150)
    assert abs(float(c["speaker_emb"].norm()) - 1.0) < 1e-5
    r = S.s3gen_ref()
    assert r["prompt_feat"].shape == (1, 300, 80)
