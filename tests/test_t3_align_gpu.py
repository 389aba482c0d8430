"""GPU parity of the on-device AlignmentStreamAnalyzer (csrc/t3_decode.cu align_step; reference
inference/alignment_stream_analyzer.py:29-154, BASELINE config 5) through the C-ABI (cbx_t3_set_alignment):
  * the head-mean layer-9 attention rows the kernels extract == the reference model's own rows (golden), and
  * the step() state machine, driven by injected attention scripts that reach every branch, takes the reference class's
    own decisions (golden) and stops the decode loop when EOS is forced."""
import os

import pytest
import torch
from safetensors.torch import load_file

from chatterbox_b200 import synthetic as S

pytestmark = pytest.mark.gpu
ALIGN = load_file(os.path.join(os.path.dirname(__file__), "golden", "t3_align_golden.safetensors"))
PARAMS = dict(cfg_weight=0.5, temperature=0.8, repetition_penalty=1.2, min_p=0.05, top_p=1.0)
EOS = 6562


def _cond(dev="cuda"):
    from chatterbox_b200.t3 import T3Cond
    return T3Cond(**{k: v.clone() for k, v in S.t3_cond().items()}).to(device=dev)


def test_measured_attention_rows_match_the_reference_model():
    """30 layers, bf16 KV, 12 teacher-forced steps of the golden utterance: prompt rows (2, S) and one row per step."""
    from chatterbox_b200.t3 import T3
    m = T3(S.make_state_dict("t3"), "cuda", max_pos=512)
    text2 = torch.cat([S.text_tokens()] * 2, 0)
    forced = ALIGN["model_forced"].long()[None]
    n = forced.shape[1]
    m.inference(t3_cond=_cond(), text_tokens=text2, max_new_tokens=n, forced_tokens=forced, do_sample=False, stop_on_eos=False,
                alignment_analyzer=True, alignment_debug=True, **PARAMS)
    rows = m.last_alignment["rows"].cpu()
    want = torch.cat([ALIGN["model_rows0"], ALIGN["model_rows"]], 0)            # 2 prompt rows + rows of the forwards of tokens 0..10
    got = rows[: want.shape[0]]
    err = (got - want).abs().max().item()
    print(f"layer-9 head-mean attention rows over {want.shape[1]} text columns, {want.shape[0]} rows: max-abs vs reference {err:.3e} "
          f"(row mass {float(want[-1].sum()):.3f})")
    assert err < 2e-3          # bf16 K in the slab against the reference's fp32 K; rows sum to ~0.6
    dec = m.last_alignment["decisions"].cpu()
    assert dec[:, 3].tolist() == list(range(2, 2 + n))       # rows consumed: 2 prompt rows, then one per step


@pytest.mark.parametrize("name", ["walk", "repeat", "noise"])
def test_state_machine_takes_the_reference_decisions_and_forces_eos(name):
    """Injected attention scripts (the golden's crafted rows): decisions per step == the reference class's; greedy decoding
    emits EOS exactly when the reference forces it (and the loop stops there)."""
    from chatterbox_b200.t3 import T3
    rows, dec = ALIGN[f"{name}_rows"], ALIGN[f"{name}_decisions"]
    S_text = rows.shape[1]
    sd = S.make_state_dict("t3", n_layers=10)
    m = T3(sd, "cuda", max_pos=256, n_layers=10)
    text2 = torch.cat([S.text_tokens(S_text - 2, seed=3)] * 2, 0)
    n = dec.shape[0]
    # the reference feeds rows[0:2] at the first call and rows[k + 1] at call k: that is rows in order
    toks = m.inference(t3_cond=_cond(), text_tokens=text2, max_new_tokens=n, do_sample=False, stop_on_eos=True,
                       alignment_analyzer=True, alignment_debug=True, alignment_inject=rows, **PARAMS)
    got = m.last_alignment["decisions"].cpu()
    k = got.shape[0]
    flags = dec[:k, 1] | (dec[:k, 2] << 1) | (dec[:k, 3] << 2) | (dec[:k, 4] << 3)
    assert got[:, 0].tolist() == dec[:k, 0].tolist(), name
    assert got[:, 1].tolist() == flags.tolist(), name
    # forced EOS that the suppression does not override ends the utterance at that step
    eos_steps = [i for i in range(n) if dec[i, 3] and not dec[i, 4]]
    if eos_steps:
        assert k == eos_steps[0] + 1 and int(toks[0, -1]) == EOS, (k, eos_steps[:3], toks[0, -3:])
    else:
        assert k == n
    # while EOS is suppressed it is never emitted
    assert all(int(toks[0, i]) != EOS for i in range(k) if dec[i, 4])
    print(f"{name}: {k} steps decoded, decisions identical to the reference class; forced-EOS stop at {eos_steps[:1]}")


def test_analyzer_off_is_the_default_and_changes_nothing():
    from chatterbox_b200.t3 import T3
    sd = S.make_state_dict("t3", n_layers=10)
    m = T3(sd, "cuda", max_pos=256, n_layers=10)
    text2 = torch.cat([S.text_tokens(20, seed=3)] * 2, 0)
    a = m.inference(t3_cond=_cond(), text_tokens=text2, max_new_tokens=16, do_sample=False, stop_on_eos=False, **PARAMS)
    b = m.inference(t3_cond=_cond(), text_tokens=text2, max_new_tokens=16, do_sample=False, stop_on_eos=False,
                    alignment_analyzer=True, **PARAMS)
    c = m.inference(t3_cond=_cond(), text_tokens=text2, max_new_tokens=16, do_sample=False, stop_on_eos=False, **PARAMS)
    assert torch.equal(a, c) and m.last_alignment is None
    assert b.shape == a.shape          # synthetic weights never pick EOS: suppression changes no greedy token here
    assert torch.equal(a, b)
