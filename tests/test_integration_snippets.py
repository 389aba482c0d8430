"""Executes the code blocks of INTEGRATION.md section 2 VERBATIM against libcbx.so (a wrong arity or a stale argument
list in the document fails here) and checks their results against the host mirrors of the same seams."""
import os
import re

import pytest
import torch

from chatterbox_b200 import synthetic as S

DOC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "INTEGRATION.md")


def _snippets():
    text = open(DOC).read()
    out = {}
    for m in re.finditer(r"<!-- snippet: (\w+) -->\n```python\n(.*?)```", text, re.S):
        out[m.group(1)] = m.group(2)
    return out


def test_document_has_the_four_snippets():
    sn = _snippets()
    assert set(sn) == {"stub", "t3", "cfm", "hift"}
    assert "t_vals" not in sn["cfm"] and "(C.c_float * n_steps)(*dts)" in sn["cfm"]      # the call carries both schedules


@pytest.mark.gpu
def test_snippets_run_verbatim_and_agree_with_the_host_mirrors():
    import ctypes as C
    from chatterbox_b200 import _lib
    from chatterbox_b200.cfm import CausalConditionalCFM, ConditionalDecoder
    from chatterbox_b200.hift import HiFTGenerator
    from chatterbox_b200.t3 import T3, T3Cond
    sn = _snippets()
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    ns = {"LIBCBX_PATH": _lib.LIB_PATH, "torch": torch, "stream": stream}
    exec(sn["stub"], ns)
    # ---- T3 ----
    params = dict(cfg_weight=0.5, temperature=0.8, repetition_penalty=1.2, min_p=0.05, top_p=1.0)
    t3 = T3(S.make_state_dict("t3", n_layers=2), dev, max_pos=128, n_layers=2)
    cond = T3Cond(**{k: v.clone() for k, v in S.t3_cond(prompt_len=20).items()}).to(device=dev)
    text2 = torch.cat([S.text_tokens(6)] * 2, 0).to(dev)
    want = t3.inference(t3_cond=cond, text_tokens=text2, max_new_tokens=8, do_sample=False, stop_on_eos=False, **params)
    emb, _ = t3.prepare_input_embeds(t3_cond=cond, text_tokens=text2, speech_tokens=6561 * torch.ones_like(text2[:, :1]), cfg_weight=0.5)
    bos = (t3.fe["speech_emb.weight"][6561] + t3.fe["speech_pos_emb.emb.weight"][0])[None, None]
    ns.update(embeds=torch.cat([emb, bos.expand(2, -1, -1)], 1), h_t3=t3._h, max_new_tokens=8, do_sample=False, stop_on_eos=False, **params)
    exec(sn["t3"], ns)
    assert torch.equal(ns["speech_tokens"], want)
    # ---- CFM ----
    sd = S.make_state_dict("flow")
    est = ConditionalDecoder(sd, dev, prefix="decoder.estimator.", max_len=64)
    cfm = CausalConditionalCFM(est, rand_noise=S.cfm_noise())
    L = 48
    inp = {k: v.to(dev).contiguous() for k, v in S.cfm_inputs(L).items()}
    want, _ = cfm(mu=inp["mu"], mask=None, n_timesteps=10, spks=inp["spks"], cond=inp["cond"])
    t_span = 1 - torch.cos(torch.linspace(0, 1, 11, dtype=torch.float32) * 0.5 * torch.pi)
    ns.update(h_cfm=est._h, z=inp["z"], mu=inp["mu"], spks=inp["spks"], cond=inp["cond"], L=L, t_span=t_span, inference_cfg_rate=0.7)
    exec(sn["cfm"], ns)
    assert float((ns["mel"] - want).norm() / want.norm()) < 1e-5
    # ---- HiFT ----
    hg = HiFTGenerator(S.make_state_dict("hift"), dev, max_frames=32)
    mel = S.mel_input(12, seed=2).to(dev).contiguous()
    want_wav, want_src = hg.inference(mel, None, seed=5, apply_fade=False)
    ns.update(h_hift=hg._h, mel=mel, L=12, seed=5)
    exec(sn["hift"], ns)
    assert float((ns["wav"] - want_wav).norm() / want_wav.norm()) < 1e-5 and torch.equal(ns["src"], want_src)
