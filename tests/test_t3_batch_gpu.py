"""GPU parity of the batched T3 decode engine (csrc/t3_batch.cu through cbx_t3_batch_* / cbx_t3_decode_batch):
every utterance of a batch against the reference's own outputs (tests/golden/t3_golden.safetensors), the CPU oracle,
and the same utterance run alone — with different texts, voices, lengths and knobs sharing the steps (BASELINE
config 4; reference caller src/server/server.py:189-299)."""
import os

import pytest
import torch
from safetensors.torch import load_file

from chatterbox_b200 import synthetic as S

pytestmark = pytest.mark.gpu

GOLD = load_file(os.path.join(os.path.dirname(__file__), "golden", "t3_golden.safetensors"))
PARAMS = dict(cfg_weight=0.5, temperature=0.8, repetition_penalty=1.2, min_p=0.05, top_p=1.0)
TOL_FP32KV = 1e-3
TOL_BF16 = 2e-2


def _cond(seed=2, prompt_len=150, dev="cuda"):
    from chatterbox_b200.t3 import T3Cond
    return T3Cond(**{k: v.clone() for k, v in S.t3_cond(seed=seed, prompt_len=prompt_len).items()}).to(device=dev)


def _req(text, cond, n, **kw):
    from chatterbox_b200.t3_batch import T3Request
    p = dict(PARAMS)
    p.update(kw)
    return T3Request(t3_cond=cond, text_tokens=text, max_new_tokens=n, **p)


def test_batch_of_different_utterances_matches_oracle_and_single_path():
    """2 layers, fp32 KV: four slots with different text lengths / voices / cfg weights, teacher-forced; each slot's raw
    logits against the CPU oracle and against the single-utterance megakernel."""
    from chatterbox_b200.t3 import T3
    from chatterbox_b200.t3_batch import T3Batch
    from oracle import t3_oracle as T3O
    sd = S.make_state_dict("t3", n_layers=2)
    t3 = T3(sd, "cuda", max_pos=256, kv_fp32=True, n_layers=2)
    eng = T3Batch(t3, 4, max_pos=256, max_new_cap=16, kv_fp32=True, keep_logits=True)
    specs = [(5, 2, 30, 0.5), (17, 7, 12, 0.5), (40, 9, 150, 0.3), (3, 11, 20, 0.5)]     # n_text, cond seed, prompt_len, cfg
    n = 9
    reqs, refs = [], []
    for k, (nt, cs, pl, cfg) in enumerate(specs):
        text2 = torch.cat([S.text_tokens(nt, seed=20 + k)] * 2, 0)
        forced = torch.randint(0, 6561, (1, n), generator=torch.Generator().manual_seed(50 + k))
        reqs.append(_req(text2, _cond(cs, pl), n, do_sample=False, forced_tokens=forced, cfg_weight=cfg))
        p = dict(PARAMS, cfg_weight=cfg)
        _, ref = T3O.inference(sd, S.t3_cond(seed=cs, prompt_len=pl), text2, max_new_tokens=n, forced_tokens=forced, n_layers=2, **p)
        refs.append(ref)
    for s, r in enumerate(reqs):
        eng.admit(s, r)
    eng.step(n)
    n_gen, done, pos = eng.poll()
    assert n_gen == [n] * 4 and done == [1] * 4
    for s, (r, ref) in enumerate(zip(reqs, refs)):
        got = eng.logits(s, n).cpu()
        err = (got - ref).abs().max().item()
        assert torch.equal(eng.tokens(s, n).cpu(), r.forced_tokens)
        t3.inference(t3_cond=r.t3_cond, text_tokens=r.text_tokens, max_new_tokens=n, forced_tokens=r.forced_tokens,
                     return_logits=True, do_sample=False, **dict(PARAMS, cfg_weight=r.cfg_weight))
        err1 = (got - t3.last_logits.cpu()).abs().max().item()
        print(f"slot {s} (P = {eng.prefill_len[s]}): logits max-abs vs oracle {err:.3e}, vs the single-utterance megakernel {err1:.3e}")
        assert err < TOL_FP32KV and err1 < TOL_FP32KV
        # prefill 34 + N + 2, one position per forward: the last token's forward is skipped (nothing would read it)
        assert pos[s] == eng.prefill_len[s] + n - 1


@pytest.fixture(scope="module")
def t3_30():
    from chatterbox_b200.t3 import T3
    return T3(S.make_state_dict("t3"), "cuda", max_pos=512, kv_fp32=False)


def test_golden_utterance_inside_a_batch_reproduces_reference_greedy_tokens(t3_30):
    """30 layers: the reference's 48 greedy tokens for the golden utterance, decoded in slot 2 of a 5-slot batch whose
    other slots run other utterances (fp32 KV: identical tokens; raw logits at the golden steps within 1e-3)."""
    from chatterbox_b200.t3_batch import T3Batch
    eng = T3Batch(t3_30, 5, max_pos=384, max_new_cap=64, kv_fp32=True, keep_logits=True)
    n = GOLD["greedy_tokens"].numel()
    gold_req = _req(torch.cat([S.text_tokens()] * 2, 0), _cond(), n, do_sample=False, stop_on_eos=False)
    others = [_req(torch.cat([S.text_tokens(10 + 7 * k, seed=3 + k)] * 2, 0), _cond(4 + k, 60 + 20 * k), n - 5 * k,
                   do_sample=False, stop_on_eos=False) for k in range(4)]
    order = [others[0], others[1], gold_req, others[2], others[3]]
    for s, r in enumerate(order):
        eng.admit(s, r)
    eng.step(n)
    n_gen, done, _ = eng.poll()
    assert done == [1] * 5 and n_gen == [r.max_new_tokens for r in order]
    toks = eng.tokens(2, n).cpu()[0].to(torch.int32)
    assert torch.equal(toks, GOLD["greedy_tokens"])
    got = eng.logits(2, n).cpu()
    for j, st in enumerate(GOLD["greedy_logit_steps"].tolist()):
        err = (got[st] - GOLD["greedy_logits"][j]).abs().max().item()
        print(f"golden utterance in slot 2 of 5, step {st}: raw logits max-abs vs reference {err:.3e}")
        assert err < TOL_FP32KV


def test_an_utterance_gives_the_same_tokens_alone_and_in_a_batch(t3_30):
    """bf16 KV (the serving mode), 30 layers, 64 greedy tokens: utterance A alone in a 1-slot engine, and in slot 3 of an
    8-slot engine next to seven other utterances.  Rows never mix in the GEMMs; what differs with the batch width is the
    KV split count of the attention (summation order), so logits agree to fp32 rounding and the greedy tokens are equal."""
    from chatterbox_b200.t3_batch import T3Batch
    n = 64
    mk = lambda: _req(torch.cat([S.text_tokens(33, seed=77)] * 2, 0), _cond(9, 100), n, do_sample=False, stop_on_eos=False)
    solo = T3Batch(t3_30, 1, max_pos=384, max_new_cap=n, keep_logits=True)
    solo.admit(0, mk())
    solo.step(n)
    t_solo, l_solo = solo.tokens(0, n).cpu(), solo.logits(0, n).cpu()
    s_solo = solo.info()["kv_splits"]
    del solo
    eng = T3Batch(t3_30, 8, max_pos=384, max_new_cap=n, keep_logits=True)
    for s in range(8):
        eng.admit(s, mk() if s == 3 else _req(torch.cat([S.text_tokens(8 + 5 * s, seed=s)] * 2, 0), _cond(20 + s, 40 + 10 * s), n,
                                                do_sample=(s % 2 == 0), seed=s, stop_on_eos=False))
    eng.step(n)
    assert eng.poll()[0] == [n] * 8
    t_b, l_b = eng.tokens(3, n).cpu(), eng.logits(3, n).cpu()
    d = (l_b - l_solo).abs().max().item()
    print(f"alone ({s_solo} KV splits) vs slot 3 of 8 ({eng.info()['kv_splits']} KV splits): logits max-abs {d:.3e}, "
          f"{int((t_b == t_solo).sum())}/{n} greedy tokens equal")
    assert torch.equal(t_b, t_solo) and d < 2e-4
    # and against the single-utterance megakernel (another summation order: tokens may part at a near-tie)
    one = t3_30.inference(t3_cond=_cond(9, 100), text_tokens=torch.cat([S.text_tokens(33, seed=77)] * 2, 0), max_new_tokens=n,
                          do_sample=False, stop_on_eos=False, return_logits=True, **PARAMS)
    same = int((one.cpu() == t_b).sum())
    d1 = (t3_30.last_logits.cpu()[0] - l_b[0]).abs().max().item()
    print(f"vs the single-utterance megakernel: {same}/{n} greedy tokens equal, step-0 logits max-abs {d1:.3e}")
    assert d1 < 2e-3 and same >= n // 2


def test_prompts_prefilled_together_match_prompts_prefilled_one_by_one(t3_30):
    """`admit_many` (cbx_t3_batch_prefill_many): six requests — four with the same prompt length (one pass of 2 x 4 x P rows
    through the layers, widest GEMM tiles, no split-K), two with other lengths (runs of one) — against the same requests
    admitted one by one into another engine.  The prompts' rows never mix; the GEMMs' tile / split-K choice differs with the
    row count, so step-0 logits agree to fp32 summation order and the greedy tokens are equal."""
    from chatterbox_b200.t3_batch import T3Batch
    n = 32
    def mk(k):
        nt, pl = (21, 60) if k in (1, 2, 3, 4) else (9 + 7 * k, 30 + 10 * k)
        return _req(torch.cat([S.text_tokens(nt, seed=50 + k)] * 2, 0), _cond(30 + k, pl), n, do_sample=False, stop_on_eos=False)
    one = T3Batch(t3_30, 6, max_pos=384, max_new_cap=n, keep_logits=True)
    for k in range(6):
        one.admit(k, mk(k))
    one.step(n)
    many = T3Batch(t3_30, 6, max_pos=384, max_new_cap=n, keep_logits=True)
    Ps = many.admit_many(0, [mk(k) for k in range(6)])
    assert Ps[1] == Ps[2] == Ps[3] == Ps[4] and Ps == [one.prefill_len[k] for k in range(6)]
    many.step(n)
    assert many.poll()[0] == [n] * 6
    for k in range(6):
        ta, tb = one.tokens(k, n).cpu(), many.tokens(k, n).cpu()
        d = (one.logits(k, n)[0] - many.logits(k, n)[0]).abs().max().item()
        print(f"slot {k} (P = {Ps[k]}): step-0 logits max-abs {d:.3e}, {int((ta == tb).sum())}/{n} greedy tokens equal")
        assert d < 2e-3 and int((ta == tb).sum()) >= n - 2


def test_eos_retires_a_slot_and_the_scheduler_refills_it():
    """Continuous batching: 7 requests of different lengths through a 3-slot engine; EOS (forced at different steps)
    retires slots early, freed slots are refilled from the queue, every request's tokens equal its forced prefix up to
    and including EOS (t3.py:343-348)."""
    from chatterbox_b200.t3 import T3
    from chatterbox_b200.t3_batch import BatchScheduler, T3Batch
    sd = S.make_state_dict("t3", n_layers=1)
    t3 = T3(sd, "cuda", max_pos=128, n_layers=1)
    eng = T3Batch(t3, 3, max_pos=128, max_new_cap=24)
    sch = BatchScheduler(eng, poll_every=4)
    want = {}
    for k in range(7):
        n = 6 + 2 * k
        forced = torch.randint(0, 6561, (1, n), generator=torch.Generator().manual_seed(k))
        eos_at = None
        if k % 2 == 1:
            eos_at = 2 + k
            forced[0, eos_at] = 6562
        sch.submit(_req(torch.cat([S.text_tokens(3 + k, seed=k)] * 2, 0), _cond(2 + k, 10 + k), n, forced_tokens=forced,
                        do_sample=False, tag=k))
        want[k] = forced[0, : (eos_at + 1 if eos_at is not None else n)]
    done = sch.run()
    assert sorted(r.tag for r in done) == list(range(7))
    for r in done:
        assert torch.equal(r.tokens.cpu()[0], want[r.tag]), (r.tag, r.tokens, want[r.tag])
    assert eng.poll()[1] == [-1, -1, -1]          # every slot free again


def test_batch_sampler_draws_lie_in_the_kept_set():
    """do_sample slots with per-slot seeds: every first-step draw lies in the set HF's min-p / top-p warpers keep for
    the slot's logits (t3.py:302-304,336-341), and the dominant tokens come up with their processed probabilities."""
    from chatterbox_b200.t3 import T3
    from chatterbox_b200.t3_batch import T3Batch
    from oracle import t3_oracle as T3O
    sd = S.make_state_dict("t3", n_layers=1)
    t3 = T3(sd, "cuda", max_pos=96, n_layers=1)
    U = 16
    eng = T3Batch(t3, U, max_pos=96, max_new_cap=4)
    text2 = torch.cat([S.text_tokens(3)] * 2, 0)
    p = dict(cfg_weight=0.5, temperature=0.25, repetition_penalty=1.2, min_p=0.05, top_p=0.9)
    _, raw = T3O.inference(sd, S.t3_cond(prompt_len=20), text2, max_new_tokens=1, n_layers=1, **p)
    probs = torch.softmax(T3O.process_logits(raw[0], torch.tensor([[T3O.START_SPEECH]]), **p), -1)[0]
    kept = set(torch.nonzero(probs > 0).flatten().tolist())
    draws = []
    for rnd in range(12):
        for s in range(U):
            eng.admit(s, _req(text2, _cond(2, 20), 1, do_sample=True, seed=1000 + rnd * U + s, **p))
        eng.step(1)
        n_gen, done, _ = eng.poll()
        assert n_gen == [1] * U and done == [1] * U
        draws += [int(eng.tokens(s, 1)[0, 0]) for s in range(U)]
    assert set(draws) <= kept, sorted(set(draws) - kept)[:5]
    top = torch.topk(probs, 2)
    N = len(draws)
    for pr, idx in zip(top.values.tolist(), top.indices.tolist()):
        f = draws.count(idx) / N
        assert abs(f - pr) < 5 * (pr * (1 - pr) / N) ** 0.5 + 0.02, (idx, f, pr)
