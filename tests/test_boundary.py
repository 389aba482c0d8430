"""The drop-in boundary (chatterbox_b200.tts / vc / s3gen mirrors of the reference API): host logic on CPU,
the full T3 -> S3Gen -> HiFT pipeline on the GPU against the chained CPU oracles."""
import os

import pytest
import torch

from chatterbox_b200 import synthetic as S


def test_punc_norm_known_answers():
    from chatterbox_b200.tts import punc_norm
    assert punc_norm("") == "You need to add some text for me to talk."
    assert punc_norm("hello   world") == "Hello world."
    assert punc_norm("wait... what; no: yes") == "Wait,  what,  no, yes."     # replacements run after the space collapse
    assert punc_norm("“quoted” – dash — ok!") == "\"quoted\" - dash - ok!"
    assert punc_norm("trailing space ,here ") == "Trailing space,here."


@pytest.mark.refonly
def test_punc_norm_equals_reference_source():
    """Executes the reference's own punc_norm (extracted from tts.py with ast; the module itself needs librosa)."""
    import ast
    from chatterbox_b200.tts import punc_norm
    src = open("/root/reference/src/chatterbox/tts.py").read()
    fn = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "punc_norm")
    ns = {}
    exec(compile(ast.Module([fn], []), "ref_punc_norm", "exec"), ns)
    for t in ["", "a", "hello   world", "wait... what; no: yes", "x — y – z", "“q” ‘r’", "end , mid ;", "Ok?", "dash -",
              "multi\nline  text…", " lead", "colon: semi; dots..."]:
        assert punc_norm(t) == ns["punc_norm"](t), t


def test_drop_invalid_tokens_and_conditionals_roundtrip(tmp_path):
    from chatterbox_b200.tts import Conditionals, drop_invalid_tokens
    from chatterbox_b200.t3 import T3Cond
    assert drop_invalid_tokens(torch.tensor([6561, 5, 6, 6562, 7])).tolist() == [5, 6]
    assert drop_invalid_tokens(torch.tensor([[1, 2, 3]])).tolist() == [1, 2, 3]
    assert drop_invalid_tokens(torch.tensor([4, 6562])).tolist() == [4]
    c = Conditionals(T3Cond(**S.t3_cond()), {k: v for k, v in S.s3gen_ref().items()})
    p = tmp_path / "conds.pt"
    c.save(p)
    d = Conditionals.load(p)
    assert torch.equal(d.t3.cond_prompt_speech_tokens, c.t3.cond_prompt_speech_tokens)
    assert torch.equal(d.gen["prompt_feat"], c.gen["prompt_feat"]) and d.gen["prompt_feat_len"] is None


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_models_refuse_to_build_without_cuda():
    from chatterbox_b200._lib import CbxError
    from chatterbox_b200.tts import ChatterboxTTS
    with pytest.raises(CbxError):
        ChatterboxTTS.from_state_dicts({}, {}, {}, "cuda")


@pytest.fixture(scope="module")
def tts():
    from chatterbox_b200.t3 import T3Cond
    from chatterbox_b200.tts import ChatterboxTTS, Conditionals
    conds = Conditionals(T3Cond(**S.t3_cond(prompt_len=30)), S.s3gen_ref(seed=4, prompt_len=20))
    return ChatterboxTTS.from_state_dicts(S.make_state_dict("t3"), S.make_state_dict("flow"), S.make_state_dict("hift"),
                                          "cuda", conds=conds, rand_noise=S.cfm_noise(), max_speech_tokens=64,
                                          max_text_tokens=32, max_prompt_tokens=32)


@pytest.mark.gpu
def test_generate_end_to_end(tts):
    ids = S.text_tokens(10)[:, 1:-1]            # generate() adds SOT / EOT itself (tts.py:240-243)
    wav = tts.generate(text_tokens=ids, do_sample=False, max_new_tokens=24, stop_on_eos=False, seed=3)
    n_tok = int(tts.last_speech_tokens.numel())
    assert wav.device.type == "cpu" and wav.shape == (1, 960 * n_tok) and n_tok > 0
    assert torch.isfinite(wav).all() and float(wav.abs().max()) <= 0.99 + 1e-6
    assert float(wav[0, :480].abs().max()) == 0.0
    # sampling path + exaggeration update side effect (tts.py:225-231)
    wav2 = tts.generate(text_tokens=ids, exaggeration=0.7, max_new_tokens=12, stop_on_eos=False, seed=4)
    assert float(tts.conds.t3.emotion_adv[0, 0, 0]) == pytest.approx(0.7) and wav2.shape[0] == 1
    with pytest.raises(AssertionError):
        tts.generate(text="hello")            # no tokenizer.json offline


@pytest.mark.gpu
def test_s3gen_chain_matches_chained_oracles(tts):
    """tokens -> mel -> wav on the GPU vs flow front-end + CFM oracle + HiFT oracle on the CPU, with the GPU's
    own (stochastic) source signal injected into the oracle vocoder."""
    from oracle import flow_oracle as FF
    from oracle import cfm_oracle as CO, hift_oracle as HO
    tok = S.speech_tokens(24, seed=9)
    ref = S.s3gen_ref(seed=4, prompt_len=20)
    wav, src = tts.s3gen.inference(tok, ref_dict=ref, seed=5)
    mel_gpu = tts.s3gen.flow_inference(tok, ref_dict=ref).cpu()
    sd = S.make_state_dict("flow")
    with torch.inference_mode():
        mu, spks, cond, l1 = FF.flow_inputs(sd, tok, ref["prompt_token"], ref["prompt_feat"], ref["embedding"])
        mel = CO.solve_euler(sd, S.cfm_noise()[:, :, : mu.shape[2]], mu, spks, cond, pfx="decoder.estimator.")[:, :, l1:]
        w = HO.fold_weight_norm(S.make_state_dict("hift"))
        ref_wav, _ = HO.inference(w, mel, s=src.cpu())
        ref_wav_same_mel, _ = HO.inference(w, mel_gpu, s=src.cpu())
    r_mel = float((mel_gpu - mel).norm() / mel.norm())
    r_voc = float((wav.cpu() - ref_wav_same_mel).norm() / ref_wav_same_mel.norm())
    r_all = float((wav.cpu() - ref_wav).norm() / ref_wav.norm())
    print(f"mel rel-L2 {r_mel:.3e}; vocoder (same mel) rel-L2 {r_voc:.3e}; chained wav rel-L2 {r_all:.3e}")
    assert r_mel < 2e-2 and r_voc < 2e-2
    assert r_all < 1e-1      # chained: the vocoder amplifies the (in-tolerance) mel difference


@pytest.mark.gpu
def test_vc_generate(tts):
    from chatterbox_b200.vc import ChatterboxVC
    vc = ChatterboxVC(tts.s3gen, "cuda", ref_dict=S.s3gen_ref(seed=4, prompt_len=20))
    wav = vc.generate(speech_tokens=S.speech_tokens(20, seed=2), seed=1)
    assert wav.shape == (1, 960 * 20) and wav.device.type == "cpu"
    with pytest.raises(NotImplementedError):
        vc.generate("some.wav")


@pytest.mark.gpu
def test_front_end_repeated_calls_and_new_lengths(tts):
    """The front-end caches a plan (tensor maps + position projections) per token count: repeated calls of one shape with
    different tokens, interleaved with other lengths (plan cache churn), must reproduce the first answers."""
    ref = S.s3gen_ref(seed=4, prompt_len=20)
    s3 = tts.s3gen
    toks = [S.speech_tokens(31, seed=s) for s in (11, 12, 13)]
    want = [s3.flow_inference(t, ref_dict=ref).clone() for t in toks]
    for n in (5, 9, 17, 23, 29, 33, 37, 41, 43):          # more distinct lengths than cached plans
        s3.flow_inference(S.speech_tokens(n, seed=n), ref_dict=ref)
    got = [s3.flow_inference(t, ref_dict=ref) for t in toks]
    for w, g in zip(want, got):
        assert torch.equal(w, g)
    assert not torch.equal(want[0], want[1])


@pytest.mark.gpu
def test_s3gen_chain_at_the_bench_shape():
    """BASELINE configs[1] S3Gen stage at its own size: 256 speech tokens + a 150-token prompt -> L = 812 mel frames
    through the 10-step solve -> 512 frames through HiFT (10.24 s of audio), against the chained CPU oracles with the
    GPU's source signal injected.  Reports mel and waveform rel-L2."""
    from oracle import flow_oracle as FF
    from chatterbox_b200.s3gen import S3Gen
    from oracle import cfm_oracle as CO, hift_oracle as HO
    sd, sdh = S.make_state_dict("flow"), S.make_state_dict("hift")
    s3 = S3Gen(sd, sdh, "cuda", rand_noise=S.cfm_noise(), max_mel_len=1024)
    tok = S.speech_tokens(256, seed=5)
    ref = S.s3gen_ref(seed=4, prompt_len=150)
    wav, src = s3.inference(tok, ref_dict=ref, seed=5)
    mel_gpu = s3.flow_inference(tok, ref_dict=ref).cpu()
    assert mel_gpu.shape == (1, 80, 512) and wav.shape == (1, 480 * 512)
    with torch.inference_mode():
        mu, spks, cond, l1 = FF.flow_inputs(sd, tok, ref["prompt_token"], ref["prompt_feat"], ref["embedding"])
        assert mu.shape[2] == 812 and l1 == 300
        mel = CO.solve_euler(sd, S.cfm_noise()[:, :, :812], mu, spks, cond, pfx="decoder.estimator.")[:, :, l1:]
        w = HO.fold_weight_norm(sdh)
        ref_wav_same_mel, _ = HO.inference(w, mel_gpu, s=src.cpu())
        ref_wav, _ = HO.inference(w, mel, s=src.cpu())
    r_mel = float((mel_gpu - mel).norm() / mel.norm())
    r_voc = float((wav.cpu() - ref_wav_same_mel).norm() / ref_wav_same_mel.norm())
    r_all = float((wav.cpu() - ref_wav).norm() / ref_wav.norm())
    print(f"bench shape (256 tokens, L=812, 512 frames): mel rel-L2 {r_mel:.3e}; vocoder on the same mel {r_voc:.3e}; "
          f"chained waveform {r_all:.3e}")
    assert r_mel < 2e-2 and r_voc < 2e-2
    assert r_all < 1e-1      # the vocoder amplifies the (in-tolerance) mel difference; stated, not hidden


@pytest.mark.gpu
def test_s3gen_inference_batch_streams_and_ragged_lengths():
    """`S3Gen.inference_batch` (BASELINE config 4: one padded flow-matching solve for several utterances, then the vocoder per
    utterance on a pool of streams / handles).  (a) the vocoder pool must not change a bit against the one-stream path, nor
    against the vocoder called alone on the same mel — same kernels, same inputs, other streams / sibling handles; (b) the mel
    of every utterance of a ragged batch (40 / 64 / 17 / 64 / 33 tokens, two voices) must match its own single-utterance
    `flow_inference` to the tolerance the padded solve has in tests/test_cfm_gpu.py.  (Waveforms of (b) are not compared: the
    sine source integrates f0, so an in-tolerance mel difference drifts its phase — the chained test above injects the source
    for that reason.)"""
    from chatterbox_b200.s3gen import S3Gen
    sd, sdh = S.make_state_dict("flow"), S.make_state_dict("hift")
    ns = [40, 64, 17, 64, 33]
    toks = [S.speech_tokens(n, seed=20 + i)[0] for i, n in enumerate(ns)]
    refs = [S.s3gen_ref(seed=4 + (i & 1), prompt_len=30) for i in range(len(ns))]
    outs = {}
    for k in (1, 3):
        s3 = S3Gen(sd, sdh, "cuda", rand_noise=S.cfm_noise(), max_mel_len=512, max_batch=len(ns), vocoder_streams=k)
        for _ in range(2):       # second call: every graph / plan cache hot, pool reused
            outs[k] = [(w.cpu(), s.cpu()) for w, s in s3.inference_batch(toks, refs, seed=7)]
    assert len(s3._voc_pool) == 3 and len(s3._front_pool) == 3 and s3.launches() > 0
    for (w1, s1), (w3, s3_), n in zip(outs[1], outs[3], ns):
        assert w1.shape == (1, 480 * 2 * n) and torch.isfinite(w1).all()
        assert torch.equal(w1, w3) and torch.equal(s1, s3_)
    mels = s3.flow_inference_batch(toks, refs)
    for u, (m, t, r, n) in enumerate(zip(mels, toks, refs, ns)):
        assert m.shape == (1, 80, 2 * n)
        w_alone, _ = s3.hift_inference(m, None, seed=7, apply_fade=True)
        assert torch.equal(w_alone.cpu(), outs[3][u][0])
        alone = s3.flow_inference(t[None], ref_dict=r)
        rel = float((m - alone).norm() / alone.norm())
        print(f"flow_inference_batch vs flow_inference, {n} tokens: mel rel-L2 {rel:.3e}")
        assert rel < 5e-3
