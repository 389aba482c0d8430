"""Flow front-end (token -> mu, spks, cond): CPU: the oracle restatement (oracle/flow_oracle.py) against the CFM inputs
captured from the UNMODIFIED reference `CausalMaskedDiffWithXvec.inference` (`oracle/make_golden.py flow`); GPU: the kernels
of csrc/s3_front.cu (through cbx_flowfe_run) against the same capture and against the oracle at the BASELINE sizes, and
front-end + CUDA CFM solve against the reference's final mel."""
import os

import pytest
import torch
from safetensors.torch import load_file

from chatterbox_b200 import synthetic as S
from oracle import flow_oracle as FO

GOLD = load_file(os.path.join(os.path.dirname(__file__), "golden", "flow_golden.safetensors"))
TOL_MU = 6e-3        # relative L2 of mu.  Measured on B200: 3.2e-3 .. 3.4e-3 at T = 1 .. 900 tokens, independent of the length:
                     # it is the bf16 rounding of the encoder weights (the oracle keeps them fp32); activations are hi / lo
                     # split bf16 pairs (~2^-17) with fp32 accumulation, attention and layer norms are fp32


def _inputs():
    return S.s3gen_ref(seed=4, prompt_len=20), S.speech_tokens(30, seed=5)


def _rel(a, b):
    return float((a - b).norm() / b.norm())


def test_oracle_frontend_matches_reference_capture():
    sd = S.make_state_dict("flow")
    ref, tok = _inputs()
    with torch.inference_mode():
        mu, spks, cond, l1 = FO.flow_inputs(sd, tok, ref["prompt_token"], ref["prompt_feat"], ref["embedding"])
    assert l1 == 40 and mu.shape == (1, 80, 100)
    assert (mu - GOLD["mu"]).abs().max() < 1e-4
    assert (spks - GOLD["spks"]).abs().max() < 1e-6
    assert torch.equal(cond, GOLD["cond"])


def test_rel_shift_known_answer():
    x = torch.arange(3 * 5, dtype=torch.float32).view(1, 1, 3, 5)
    y = FO._rel_shift(x)
    # row i keeps columns (T-1-i) .. (2T-2-i): what the attention kernel reads as p[T - 1 - i + j]
    assert y.shape == (1, 1, 3, 3)
    assert y[0, 0].tolist() == [[2.0, 3.0, 4.0], [6.0, 7.0, 8.0], [10.0, 11.0, 12.0]]


def test_pack_front_layout():
    from chatterbox_b200.flow_frontend import pack_front
    bf, f32, bfo, fo = pack_front(S.make_state_dict("flow"), "cpu")
    assert len(bfo) == 56 and len(fo) == 115 and all(o % 64 == 0 for o in bfo)
    assert bf.dtype == torch.bfloat16 and f32.dtype == torch.float32


@pytest.mark.gpu
def test_frontend_kernels_match_reference_capture():
    from chatterbox_b200.flow_frontend import FlowFrontend
    ref, tok = _inputs()
    fe = FlowFrontend(S.make_state_dict("flow"), "cuda", max_tokens=64)
    mu, spks, cond, l1 = fe.flow_inputs(tok, ref["prompt_token"], ref["prompt_feat"], ref["embedding"])
    r = _rel(mu.cpu(), GOLD["mu"])
    print(f"front-end kernels, T = 50 tokens: mu rel-L2 vs the reference capture {r:.3e}, spks max-abs "
          f"{float((spks.cpu() - GOLD['spks']).abs().max()):.2e}")
    assert l1 == 40 and r < TOL_MU
    assert (spks.cpu() - GOLD["spks"]).abs().max() < 1e-5 and torch.equal(cond.cpu(), GOLD["cond"])


@pytest.mark.gpu
@pytest.mark.parametrize("n_tok,n_prompt", [(256, 150), (750, 150), (1, 0), (33, 7)])
def test_frontend_kernels_match_oracle_at_baseline_sizes(n_tok, n_prompt):
    """T = 406 (config 2) and T = 900 (config 3) tokens, plus the smallest case and a ragged one (tile edges of the 32-query /
    32-key attention tiles and of the 128-row GEMM tile)."""
    from chatterbox_b200.flow_frontend import FlowFrontend
    sd = S.make_state_dict("flow")
    ref = S.s3gen_ref(seed=4, prompt_len=max(n_prompt, 1))
    ptk, pf = ref["prompt_token"][:, :n_prompt], ref["prompt_feat"][:, : 2 * n_prompt]
    tok = S.speech_tokens(n_tok, seed=5)
    with torch.inference_mode():
        want = FO.flow_inputs(sd, tok, ptk, pf, ref["embedding"])
    fe = FlowFrontend(sd, "cuda", max_tokens=1024)
    got = fe.flow_inputs(tok, ptk, pf, ref["embedding"])
    r = _rel(got[0].cpu(), want[0])
    print(f"front-end kernels, T = {n_tok + n_prompt} tokens (L = {2 * (n_tok + n_prompt)}): mu rel-L2 vs oracle {r:.3e}")
    assert r < TOL_MU and got[3] == want[3]
    assert (got[1].cpu() - want[1]).abs().max() < 1e-5 and torch.equal(got[2].cpu(), want[2])


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
