"""CPU: the CFM oracle restatement against the golden vectors produced by the UNMODIFIED reference
(`oracle/make_golden.py cfm`), plus known-answer checks of the pieces."""
import os

import torch
from safetensors.torch import load_file

from chatterbox_b200 import synthetic as S
from oracle import cfm_oracle as CO

GOLD = load_file(os.path.join(os.path.dirname(__file__), "golden", "cfm_golden.safetensors"))
PFX = "decoder.estimator."


def _sd():
    return S.make_state_dict("flow")


def test_t_schedule_is_cosine():
    t = CO.t_schedule(10)
    assert torch.allclose(t, GOLD["t_span"])
    assert t[0] == 0 and abs(float(t[-1]) - 1.0) < 1e-6
    assert abs(float(t[5]) - (1 - 2 ** -0.5)) < 1e-6


def test_estimator_matches_reference_golden():
    sd = _sd()
    L = 96
    inp = S.cfm_inputs(L)
    z2 = torch.cat([inp["z"], inp["z"]], 0)
    mu2 = torch.cat([inp["mu"], torch.zeros_like(inp["mu"])], 0)
    sp2 = torch.cat([inp["spks"], torch.zeros_like(inp["spks"])], 0)
    c2 = torch.cat([inp["cond"], torch.zeros_like(inp["cond"])], 0)
    with torch.inference_mode():
        out = CO.estimator(sd, z2, torch.ones(2, 1, L), mu2, torch.tensor([0.3, 0.3]), sp2, c2, PFX)
    assert (out - GOLD["est_t0.3_L96"]).abs().max() < 1e-4


def test_solve_euler_matches_reference_golden():
    sd = _sd()
    L = 130
    inp = S.cfm_inputs(L, seed=6 + L)
    with torch.inference_mode():
        mel = CO.solve_euler(sd, inp["z"], inp["mu"], inp["spks"], inp["cond"], pfx=PFX)
    rel = (mel - GOLD["mel_L130"]).norm() / GOLD["mel_L130"].norm()
    assert rel < 1e-5


def test_causality_of_convs_and_full_attention():
    """CausalConv1d only looks back (decoder.py:71-97); attention is bidirectional (decoder.py:269-270)."""
    sd = _sd()
    x = torch.randn(1, 256, 20)
    w, b = sd[PFX + "down_blocks.0.2.weight"], sd[PFX + "down_blocks.0.2.bias"]
    y0 = CO.causal_conv(x, w, b)
    x2 = x.clone()
    x2[:, :, 10:] += 1.0
    y1 = CO.causal_conv(x2, w, b)
    assert torch.equal(y0[:, :, :10], y1[:, :, :10]) and not torch.equal(y0[:, :, 10:], y1[:, :, 10:])
    h = torch.randn(1, 12, 256)
    h2 = h.clone()
    h2[:, 11] += torch.randn(256)
    p = PFX + "mid_blocks.0.1.0."
    a, c = CO.transformer_block(sd, p, h, None), CO.transformer_block(sd, p, h2, None)
    assert (a[:, 0] - c[:, 0]).abs().max() > 1e-6     # a late frame influences an early one
