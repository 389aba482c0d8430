"""CPU: the token -> (mu, spks, cond) front-end restatement (chatterbox_b200/flow_frontend.py, plain torch)
against the CFM inputs captured from the UNMODIFIED reference `CausalMaskedDiffWithXvec.inference`
(`oracle/make_golden.py flow`); GPU: front-end + CUDA CFM solve against the reference's final mel."""
import os

import pytest
import torch
from safetensors.torch import load_file

from chatterbox_b200 import flow_frontend as FF
from chatterbox_b200 import synthetic as S

GOLD = load_file(os.path.join(os.path.dirname(__file__), "golden", "flow_golden.safetensors"))


def _inputs():
    return S.s3gen_ref(seed=4, prompt_len=20), S.speech_tokens(30, seed=5)


def test_frontend_matches_reference_capture():
    sd = S.make_state_dict("flow")
    ref, tok = _inputs()
    with torch.inference_mode():
        mu, spks, cond, l1 = FF.flow_inputs(sd, tok, ref["prompt_token"], ref["prompt_feat"], ref["embedding"])
    assert l1 == 40 and mu.shape == (1, 80, 100)
    assert (mu - GOLD["mu"]).abs().max() < 1e-4
    assert (spks - GOLD["spks"]).abs().max() < 1e-6
    assert torch.equal(cond, GOLD["cond"])


def test_oracle_frontend_matches_reference_capture():
    """oracle/flow_oracle.py (the checker / CPU-baseline restatement) against the same reference capture, and the
    product's torch front-end against the oracle at another shape."""
    from oracle import flow_oracle as FO
    sd = S.make_state_dict("flow")
    ref, tok = _inputs()
    with torch.inference_mode():
        mu, spks, cond, l1 = FO.flow_inputs(sd, tok, ref["prompt_token"], ref["prompt_feat"], ref["embedding"])
        assert l1 == 40 and (mu - GOLD["mu"]).abs().max() < 1e-4 and (spks - GOLD["spks"]).abs().max() < 1e-6
        assert torch.equal(cond, GOLD["cond"])
        ref2, tok2 = S.s3gen_ref(seed=9, prompt_len=7), S.speech_tokens(13, seed=6)
        a = FO.flow_inputs(sd, tok2, ref2["prompt_token"], ref2["prompt_feat"], ref2["embedding"])
        b = FF.flow_inputs(sd, tok2, ref2["prompt_token"], ref2["prompt_feat"], ref2["embedding"])
    assert (a[0] - b[0]).abs().max() < 1e-5 and (a[1] - b[1]).abs().max() < 1e-6 and torch.equal(a[2], b[2])


def test_rel_shift_known_answer():
    x = torch.arange(2 * 5, dtype=torch.float32).view(1, 1, 2, 5)      # T = 2... uses (T, 2T-1) = (3,5) below
    x = torch.arange(3 * 5, dtype=torch.float32).view(1, 1, 3, 5)
    y = FF._rel_shift(x)
    # row i keeps columns (T-1-i) .. (2T-2-i): relative positions 0 .. T-1 to the left of the diagonal
    assert y.shape == (1, 1, 3, 3)
    assert y[0, 0].tolist() == [[2.0, 3.0, 4.0], [6.0, 7.0, 8.0], [10.0, 11.0, 12.0]]


@pytest.mark.gpu
def test_flow_inference_on_gpu_matches_reference_mel():
    from chatterbox_b200.s3gen import S3Gen
    ref, tok = _inputs()
    m = S3Gen(S.make_state_dict("flow"), S.make_state_dict("hift"), "cuda", rand_noise=S.cfm_noise(), max_mel_len=512)
    mel = m.flow_inference(tok, ref_dict=ref).cpu()
    assert mel.shape == GOLD["feat"].shape
    rel = float((mel - GOLD["feat"]).norm() / GOLD["feat"].norm())
    print(f"flow_inference rel-L2 vs reference {rel:.3e}")
    assert rel < 2e-2
