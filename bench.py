#!/usr/bin/env python
"""Benchmark of the Chatterbox inference hot path on B200 (driver contract: one JSON line on rank 0).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...   # the reference's algorithm on the host cores

A "step" is one utterance of BASELINE.json configs[1], the full generate() pipeline: a 64-token (+SOT/EOT)
synthetic sentence, a 6 s synthetic reference voice (150 conditioning tokens, 300 prompt mel frames),
cfg_weight 0.5, exaggeration 0.5, greedy, exactly 256 speech tokens (EOS suppressed), random-init weights
of the named architecture: T3 prefill + 256-token decode -> flow front-end -> 10-step CFM solve at
L = 812 mel frames -> HiFT vocoder for 512 frames (10.24 s of audio).  The headline metric is speech
tokens/s through that whole pipeline (whole job, all GPUs); T3-decode tokens/s, RTF and utterances/s are
reported beside it, and the roofline object describes the dominant kernel (the T3 decode megakernel).

Timing: W >= 3 untimed steps, then K steps between barrier + cuda synchronize, CUDA events on the
launching stream, MAX over ranks.  L2: every decode step streams 1.02 GB of weights (> 126 MB L2), so
inputs are larger than L2 by construction.  N > 1: independent utterances per rank ("weak"), no
data-path collective; NCCL only for the barrier and the max-over-ranks reduction.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

N_TEXT, PROMPT_LEN, N_SPEECH = 64, 150, 256
SAMPLE = dict(cfg_weight=0.5, temperature=0.8, repetition_penalty=1.2, min_p=0.05, top_p=1.0)
KV_BYTES_PER_POS = 30 * 2 * 1024 * 2 * 2          # layers x (K,V) x hidden x CFG rows x bf16
METRIC, UNIT = "speech_tokens_per_s", "tokens/s"
AUDIO_S = N_SPEECH / 25.0


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """Samples SM clock + throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index: int, period_s: float = 0.05):
        self.index, self.period = index, period_s
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {nv.nvmlClocksEventReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksEventReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksEventReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksEventReasonSwPowerCap: "sw_power_cap",
                 nv.nvmlClocksEventReasonHwPowerBrakeSlowdown: "hw_power_brake"}
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self):
        if self._thr is not None:
            self._stop.set()
            self._thr.join()
        return {"sm_mhz": (statistics.median(self.samples) if self.samples else None),
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def _ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of the decode launch (same 256-token workload) from the latest
    committed `ncu --set full` capture (profiles/*_t3_ncu_traffic.json); None if no capture is committed."""
    import glob
    files = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "*_t3_ncu_traffic.json")))
    if not files:
        return None
    try:
        with open(files[-1]) as f:
            return json.load(f)["traffic_bytes_per_launch"]
    except Exception:
        return None


def _decode_bytes(weight_bytes: int, P: int, n: int) -> int:
    """Algorithmic HBM bytes of one decode launch of n tokens starting at cache length P:
    per token: every weight once + KV read of the cached context + KV write of one position."""
    return n * weight_bytes + KV_BYTES_PER_POS * sum(P + i for i in range(n)) + KV_BYTES_PER_POS * n


# -------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline / cuda_eager_baseline: the oracle ports of the reference's algorithm, fp32 eager torch.
# One call = ONE FULL utterance of the bench workload (nothing is sampled down or extrapolated).
# -------------------------------------------------------------------------------------------------
_SD_CACHE = {}


def _sd(name):
    from chatterbox_b200 import synthetic as S
    if name not in _SD_CACHE:
        _SD_CACHE[name] = S.make_state_dict(name)
    return _SD_CACHE[name]


class OraclePipeline:
    """The reference's own math for the whole path — `T3.inference` (t3.py:203-371) greedy for 256 tokens ->
    `flow.inference` front-end (flow.py:236-282) -> 10-step Euler/CFG solve at L = 812 (flow_matching.py:72-123) ->
    `HiFTGenerator.inference` for 512 frames (hifigan.py:462-474) — through oracle/*.py on `device` ("cpu": the
    reference arm; "cuda:0": PyTorch eager on the same B200, the library-call bar of BASELINE.md section 1).
    Weights are moved / concatenated once at construction (model load), not per utterance."""

    def __init__(self, device: str, threads: int):
        from chatterbox_b200 import synthetic as S
        from oracle import hift_oracle as HO, t3_oracle as T3O
        self.dev = torch.device(device)
        if self.dev.type == "cpu":
            torch.set_num_threads(threads)
        mv = lambda d: {k: v.to(self.dev) for k, v in d.items()}
        self.sd_t3, self.sd_flow = mv(_sd("t3")), mv(_sd("flow"))
        self.w_hift = mv(HO.fold_weight_norm(_sd("hift")))
        self.cond = {k: v.to(self.dev) for k, v in S.t3_cond(prompt_len=PROMPT_LEN).items()}
        self.text2 = torch.cat([S.text_tokens(N_TEXT)] * 2, 0).to(self.dev)
        self.ref = {k: (v.to(self.dev) if torch.is_tensor(v) else v) for k, v in S.s3gen_ref(seed=4, prompt_len=PROMPT_LEN).items()}
        self.noise = S.cfm_noise().to(self.dev)
        self.llama = T3O.LlamaOracle(self.sd_t3)
        g = torch.Generator().manual_seed(11)
        self.phase = torch.zeros(1, 9, 1)
        self.phase[:, 1:] = (torch.rand(1, 8, 1, generator=g) * 2 - 1) * torch.pi
        self.phase = self.phase.to(self.dev)
        self.src_noise = torch.randn(1, 9, 480 * 2 * N_SPEECH, generator=g).to(self.dev)

    def _sync(self):
        if self.dev.type == "cuda":
            torch.cuda.synchronize(self.dev)

    def utterance(self):
        """-> (seconds, per-stage seconds, n_tokens, wav)"""
        from oracle import cfm_oracle as CO, flow_oracle as FO, hift_oracle as HO, t3_oracle as T3O
        sd, m = self.sd_t3, self.llama
        m.k = [None] * m.n_layers
        m.v = [None] * m.n_layers
        m.pos = 0
        parts = {}
        with torch.inference_mode():
            self._sync()
            t0 = time.perf_counter()
            emb = T3O.prepare_input_embeds(sd, self.cond, self.text2, SAMPLE["cfg_weight"])
            logits = m.forward(emb)[:, -1]
            self._sync()
            t1 = time.perf_counter()
            gen = torch.tensor([[T3O.START_SPEECH]], dtype=torch.long, device=self.dev)
            for i in range(N_SPEECH):
                proc = T3O.process_logits(logits, gen, **SAMPLE)
                nxt = proc.argmax(dim=-1, keepdim=True)
                gen = torch.cat([gen, nxt], dim=1)
                e = (sd["speech_emb.weight"][nxt] + sd["speech_pos_emb.emb.weight"][i + 1]).expand(2, -1, -1)
                logits = m.forward(e)[:, -1]
            toks = gen[:, 1:]
            toks = toks[toks < 6561][None]          # drop_invalid_tokens (tts.py:257-262)
            self._sync()
            t2 = time.perf_counter()
            mu, spks, cond, l1 = FO.flow_inputs(self.sd_flow, toks, self.ref["prompt_token"], self.ref["prompt_feat"],
                                                self.ref["embedding"])
            self._sync()
            t3 = time.perf_counter()
            L = mu.shape[2]
            mel = CO.solve_euler(self.sd_flow, self.noise[:, :, :L], mu, spks, cond, pfx="decoder.estimator.")[:, :, l1:]
            self._sync()
            t4 = time.perf_counter()
            n = 480 * mel.shape[2]
            wav, _ = HO.inference(self.w_hift, mel, phase_vec=self.phase, noise=self.src_noise[:, :, :n])
            self._sync()
            t5 = time.perf_counter()
        parts = dict(t3_prefill_s=t1 - t0, t3_decode_s=t2 - t1, front_s=t3 - t2, cfm_solve_s=t4 - t3, hift_s=t5 - t4)
        return t5 - t0, parts, int(toks.shape[1]), wav


def _oracle_sample_text(threads, dev):
    where = f"{threads} host threads" if dev == "cpu" else "cuda:0 (PyTorch eager, library kernels)"
    return (f"oracle ports (oracle/*.py = the reference's fp32 math) on {where}: ONE FULL utterance of the workload per step — "
            f"T3 prefill + {N_SPEECH} greedy tokens, flow front-end, 10 Euler/CFG steps at L=812, HiFT {2 * N_SPEECH} frames; "
            "nothing extrapolated")


def run_reference(args, rank: int):
    """`--impl reference`: the reference's CPU implementation of the path on the box's host cores, all threads, the
    whole workload every step.  Under torchrun rank 0 alone runs it (the other ranks exit): at N > 1 the line is still
    ONE host process, so only the N = 1 ratio compares like with like."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    pipe = OraclePipeline("cpu", threads)
    vals, parts_all = [], []
    for i in range(args.warmup + args.steps):
        utt_s, parts, n_tok, _ = pipe.utterance()
        assert n_tok == N_SPEECH
        if i >= args.warmup:
            vals.append(utt_s)
            parts_all.append(parts)
    utt_s = statistics.mean(vals)
    parts = {k: statistics.mean(p[k] for p in parts_all) for k in parts_all[0]}
    val = N_SPEECH / utt_s
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": utt_s * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": _config(),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": _oracle_sample_text(threads, "cpu")},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "rtf": utt_s / AUDIO_S, "utterances_per_s": 1.0 / utt_s, "stages_s": parts,
            "host_processes": 1}
    print(json.dumps(line), flush=True)


def _config():
    return {"workload": "configs[1]: full generate() pipeline, single utterance: 64+2 text tokens, 6 s voice (150 cond "
                        "tokens, 300 prompt mel frames), cfg 0.5, exaggeration 0.5, greedy, 256 speech tokens, CFM 10 "
                        "steps at L=812, HiFT 512 frames (10.24 s audio), random-init weights",
            "stages": ["t3_prefill", "t3_decode", "flow_frontend", "cfm_solve", "hift"],
            "utterances_per_step_per_gpu": 1, "speech_tokens": N_SPEECH,
            "l2": "inputs larger than L2 (1.02 GB of weights streamed per decode step)"}


# -------------------------------------------------------------------------------------------------
# this repo's arm
# -------------------------------------------------------------------------------------------------
def run_ours(args, rank: int, world: int, local_rank: int):
    import torch.distributed as dist
    from chatterbox_b200 import dispatch as D
    from chatterbox_b200 import synthetic as S
    from chatterbox_b200.t3 import T3Cond
    from chatterbox_b200.tts import ChatterboxTTS, Conditionals

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        D.barrier(world)
        torch.cuda.synchronize()

    # one replica per rank, an independent utterance per rank (seeds differ), same shapes
    conds = Conditionals(T3Cond(**S.t3_cond(seed=2 + 10 * rank, prompt_len=PROMPT_LEN)),
                         S.s3gen_ref(seed=4 + 10 * rank, prompt_len=PROMPT_LEN))
    tts = ChatterboxTTS.from_state_dicts(S.make_state_dict("t3"), S.make_state_dict("flow"), S.make_state_dict("hift"),
                                         dev, conds=conds, rand_noise=S.cfm_noise(), max_speech_tokens=N_SPEECH,
                                         max_text_tokens=N_TEXT + 2, max_prompt_tokens=PROMPT_LEN)
    t3, s3 = tts.t3, tts.s3gen
    t3.record_events = True
    s3.record_events = True
    ids_host = S.text_tokens(N_TEXT, seed=1 + rank)[:, 1:-1].to(torch.int32).pin_memory()    # generate() adds SOT/EOT
    text_dev = torch.cat([S.text_tokens(N_TEXT, seed=1 + rank)] * 2, 0).to(dev)
    gen_kw = dict(do_sample=False, stop_on_eos=False, max_new_tokens=N_SPEECH, seed=7, **SAMPLE)

    def step_resident():
        """the inner seams generate() goes through, inputs already in HBM, output left in HBM"""
        c = tts.conds.t3
        cond = T3Cond(speaker_emb=c.speaker_emb, cond_prompt_speech_tokens=c.cond_prompt_speech_tokens, emotion_adv=c.emotion_adv)
        toks = t3.inference(t3_cond=cond, text_tokens=text_dev, max_new_tokens=N_SPEECH, do_sample=False,
                            stop_on_eos=False, seed=7, **SAMPLE)
        wav, _ = s3.inference(speech_tokens=toks[0], ref_dict=tts.conds.gen, seed=7)
        return toks, wav

    def step_e2e():
        """the public call: host token ids in, host waveform out (tts.py:208-272)"""
        return tts.generate(text_tokens=ids_host, **gen_kw)

    def counters():
        return t3.state()["launches"] + s3.launches()

    def stage_times(ev):
        return {"t3_prefill": ev["prefill0"].elapsed_time(ev["decode0"]), "t3_decode": ev["decode0"].elapsed_time(ev["decode1"]),
                "flow_frontend": ev["front0"].elapsed_time(ev["cfm0"]), "cfm_solve": ev["cfm0"].elapsed_time(ev["cfm1"]),
                "hift": ev["hift0"].elapsed_time(ev["hift1"])}

    warm_stages = []            # iteration 0 = every stage eager (first sight of the shapes), 1 = graph capture, 2+ = replay
    for _ in range(max(args.warmup, 3)):
        toks, wav = step_resident()
        assert toks.shape[1] == N_SPEECH and int((toks >= 6561).sum()) == 0 and wav.shape[1] == 480 * 2 * N_SPEECH
        torch.cuda.synchronize()
        warm_stages.append(stage_times({**t3.events, **s3.events}))
    l0 = counters()

    # ---- timed region 1: inputs resident in HBM ----
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    torch.cuda.profiler.start()     # ncu --profile-from-start off: the launch list covers the timed region only
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    evs = []
    e0.record()
    for _ in range(args.steps):
        step_resident()
        evs.append({**t3.events, **s3.events})
    e1.record()
    barrier()
    torch.cuda.profiler.stop()
    clocks = sampler.stop()
    total_ms = e0.elapsed_time(e1)
    launches = counters() - l0

    def stage(a, b):
        return statistics.mean(ev[a].elapsed_time(ev[b]) for ev in evs)

    stages = {"t3_prefill": stage("prefill0", "decode0"), "t3_decode": stage("decode0", "decode1"),
              "flow_frontend": stage("front0", "cfm0"), "cfm_solve": stage("cfm0", "cfm1"), "hift": stage("hift0", "hift1")}

    # ---- timed region 2: end to end through the public API with HOST inputs / outputs ----
    for _ in range(2):
        step_e2e()
    barrier()
    e0.record()
    for _ in range(args.steps):
        out = step_e2e()
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    assert out.shape == (1, 480 * 2 * N_SPEECH) and out.device.type == "cpu"

    # ---- side arms (rank 0 reports them; NOT part of the headline): the default sampler, and lengths never seen before ----
    side = {}
    if rank == 0:
        c = tts.conds.t3
        cond = T3Cond(speaker_emb=c.speaker_emb, cond_prompt_speech_tokens=c.cond_prompt_speech_tokens, emotion_adv=c.emotion_adv)

        def utterance(n_tok, do_sample):
            toks = t3.inference(t3_cond=cond, text_tokens=text_dev, max_new_tokens=n_tok, do_sample=do_sample, stop_on_eos=False,
                                seed=11, **SAMPLE)
            s3.inference(speech_tokens=toks[0].clamp(max=6560), ref_dict=tts.conds.gen, seed=7)
            torch.cuda.synchronize()
            st = stage_times({**t3.events, **s3.events})
            return sum(st.values()), st
        # (a) do_sample=True (the product default: temperature / min-p / top-p / multinomial on CTA 0, fixed Philox seed)
        utterance(N_SPEECH, True)
        runs = [utterance(N_SPEECH, True) for _ in range(3)]
        ms = statistics.mean(r[0] for r in runs)
        side["sampling_arm"] = {"what": "same utterance with do_sample=True (temperature 0.8, min_p 0.05, top_p 1.0, repetition penalty 1.2, "
                                        "Philox multinomial), EOS suppressed, graphs hot", "ms_per_utterance": ms,
                                "tokens_per_s": N_SPEECH / (ms * 1e-3), "t3_decode_ms": statistics.mean(r[1]["t3_decode"] for r in runs)}
        # (b) speech-token counts never seen before: the mel length is new, so the CFM solve / front-end plan / HiFT run eagerly
        lens = [N_SPEECH - 5, N_SPEECH - 13, N_SPEECH - 19]
        vruns = [(n, *utterance(n, False)) for n in lens]
        side["variable_length_arm"] = {"what": "greedy utterances of 251 / 243 / 237 speech tokens, each length seen for the first time "
                                               "(no CUDA graph of that mel length yet: eager launches)",
                                       "tokens_per_s": sum(v[0] for v in vruns) / (sum(v[1] for v in vruns) * 1e-3),
                                       "ms_per_utterance": [v[1] for v in vruns],
                                       "stages_ms_mean": {k: statistics.mean(v[2][k] for v in vruns) for k in vruns[0][2]}}
        side["first_call_stages_ms"] = {"eager_first_call": warm_stages[0], "graph_capture_call": warm_stages[1]}
    barrier()

    total_ms, e2e_ms = D.max_over_ranks([total_ms, e2e_ms], dev, world)

    if rank == 0:
        peak, peak_src = _peaks()
        P = t3.last_prefill_len
        wbytes = t3.packed.weight_bytes_per_step()
        dbytes = _decode_bytes(wbytes, P, N_SPEECH)
        d_ms = stages["t3_decode"]
        achieved = dbytes / (d_ms * 1e-3) / 1e9
        tokens = N_SPEECH * args.steps * world
        step_s = total_ms / args.steps * 1e-3
        L = 2 * (PROMPT_LEN + N_SPEECH)
        cfm_flop = 10 * 2 * L * (132.18e6 + 114688.0 * L)              # per estimator call x 10 steps
        hift_flop = 612.3e6 * 2 * N_SPEECH
        line = {"metric": METRIC, "value": D.whole_job_rate(N_SPEECH * args.steps, world, total_ms), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic", "config": _config(),
                "e2e": {"value": tokens / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(ids_host.numel() * 4),
                        "d2h_bytes_per_step": int(out.numel() * 4) + 4},
                "gpu_launches": int(launches),
                "clocks": clocks,
                "roofline": {"kernel": "t3_megakernel<bf16,1> (one decode launch = 256 tokens)", "bound": "hbm",
                             "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": _ncu_traffic(), "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": dbytes, "launch_ms": d_ms,
                             "share_of_step": d_ms / (total_ms / args.steps)},
                "stages_ms": stages,
                "headline_arm": "greedy (do_sample=False), EOS suppressed, one fixed shape: every CUDA-graph cache (prefill, CFM solve) is hot "
                                "after warm-up; sampling_arm / variable_length_arm / first_call_stages_ms give the other paths",
                **side,
                "t3_decode_tokens_per_s": N_SPEECH / (d_ms * 1e-3),
                "cfm_tflops": cfm_flop / (stages["cfm_solve"] * 1e-3) / 1e12,
                "hift_tflops": hift_flop / (stages["hift"] * 1e-3) / 1e12,
                "utterances_per_s": args.steps * world / (total_ms * 1e-3),
                "rtf": step_s / AUDIO_S, "rtf_e2e": (e2e_ms / args.steps * 1e-3) / AUDIO_S}
        if not args.no_cpu_baseline and world == 1:
            # the same restatements as PyTorch eager on this B200 (library kernels, fp32): the bar BASELINE.md section 1 names
            eager = OraclePipeline(f"cuda:{local_rank}", 1)
            eager.utterance()
            runs = [eager.utterance() for _ in range(2)]
            es = statistics.mean(r[0] for r in runs)
            line["cuda_eager_baseline"] = {"value": N_SPEECH / es, "unit": UNIT, "kind": "port",
                                           "sample": _oracle_sample_text(1, "cuda"), "ms_per_utterance": es * 1e3,
                                           "stages_s": {k: statistics.mean(r[1][k] for r in runs) for k in runs[0][1]},
                                           "rtf": es / AUDIO_S}
            del eager
            torch.cuda.empty_cache()
            threads = os.cpu_count() or 1
            utt_s, parts, n_tok, _ = OraclePipeline("cpu", threads).utterance()
            line["cpu_baseline"] = {"value": N_SPEECH / utt_s, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": _oracle_sample_text(threads, "cpu"), "rtf": utt_s / AUDIO_S, "stages_s": parts}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# -------------------------------------------------------------------------------------------------
# --workload config4: BASELINE configs[3], 64 concurrent utterances per GPU (batched T3 decode), data-parallel ranks
# -------------------------------------------------------------------------------------------------
def run_config4(args, rank: int, world: int, local_rank: int):
    """A step = one batch of U utterances (independent seeds, same shapes as config 2) per GPU: the U prompts prefilled
    together into the slots of the batched engine, 256 decode steps shared by all slots (one CUDA graph per step), then S3Gen
    (front-end + 10-step CFM + HiFT) per utterance.  Reported: whole-job speech tokens/s through the pipeline (value),
    the batched-decode tokens/s per GPU, utterances/s, and the HBM roofline of the decode stage
    (1.02 GB of weights + sum over slots of 245 760 B x (context + 1) per step)."""
    import torch.distributed as dist
    from chatterbox_b200 import dispatch as D
    from chatterbox_b200 import synthetic as S
    from chatterbox_b200.s3gen import S3Gen
    from chatterbox_b200.t3 import T3, T3Cond
    from chatterbox_b200.t3_batch import T3Batch, T3Request

    U = args.slots
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        D.barrier(world)
        torch.cuda.synchronize()

    max_pos = 34 + N_TEXT + 2 + 2 + N_SPEECH + 8
    t3 = T3(S.make_state_dict("t3"), dev, max_pos=max_pos)
    eng = T3Batch(t3, U, max_pos=max_pos, max_new_cap=N_SPEECH)
    s3 = S3Gen(S.make_state_dict("flow"), S.make_state_dict("hift"), dev, rand_noise=S.cfm_noise(),
               max_mel_len=2 * (PROMPT_LEN + N_SPEECH), max_batch=args.s3_batch)
    reqs, refs = [], []
    for k in range(U):
        sd = 100 + k + 1000 * rank                     # independent seeds 100 + k
        cond = T3Cond(**S.t3_cond(seed=sd, prompt_len=PROMPT_LEN)).to(device=dev)
        text2 = torch.cat([S.text_tokens(N_TEXT, seed=sd)] * 2, 0).to(dev)
        reqs.append(T3Request(t3_cond=cond, text_tokens=text2, max_new_tokens=N_SPEECH, do_sample=False, stop_on_eos=False,
                              seed=sd, **SAMPLE))
        refs.append({k2: (v.to(dev) if torch.is_tensor(v) else v) for k2, v in S.s3gen_ref(seed=sd, prompt_len=PROMPT_LEN).items()})
    ev = {k: torch.cuda.Event(enable_timing=True) for k in ("p0", "d0", "d1", "s1")}

    def step():
        ev["p0"].record()
        eng.admit_many(0, reqs)                         # equal prompt lengths: the U prompts go through the layers together
        ev["d0"].record()
        eng.step(N_SPEECH)
        ev["d1"].record()
        toks = [eng.tokens(k, N_SPEECH) for k in range(U)]
        wavs = []
        for k0 in range(0, U, args.s3_batch):          # the flow-matching stage of s3_batch utterances as one solve
            ks = range(k0, min(U, k0 + args.s3_batch))
            outs = s3.inference_batch([toks[k][0][toks[k][0] < 6561] for k in ks], [refs[k] for k in ks], seed=7)
            wavs += [w for w, _ in outs]
        ev["s1"].record()
        for k in range(U):
            eng.release(k)
        return toks, wavs

    for _ in range(3):
        toks, wavs = step()
    assert all(int(t.shape[1]) == N_SPEECH for t in toks) and all(w.shape[1] == 480 * 2 * N_SPEECH for w in wavs)
    l0 = t3.state()["launches"] + eng.info()["launches"] + s3.launches()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    torch.cuda.profiler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st = {"prefill": [], "decode": [], "s3gen": []}
    e0.record()
    for _ in range(args.steps):
        step()
        torch.cuda.synchronize()
        st["prefill"].append(ev["p0"].elapsed_time(ev["d0"]))
        st["decode"].append(ev["d0"].elapsed_time(ev["d1"]))
        st["s3gen"].append(ev["d1"].elapsed_time(ev["s1"]))
    e1.record()
    barrier()
    torch.cuda.profiler.stop()
    clocks = sampler.stop()
    total_ms = e0.elapsed_time(e1)
    launches = t3.state()["launches"] + eng.info()["launches"] + s3.launches() - l0
    (total_ms,) = D.max_over_ranks([total_ms], dev, world)
    if rank == 0:
        peak, peak_src = _peaks()
        P = eng.prefill_len[0]
        wbytes = t3.packed.weight_bytes_per_step()
        # per decode step i (0-based): weights once + every slot's KV read (P + i positions) and write (1 position);
        # the last step's forward is skipped by the engine (its logits would never be read)
        dbytes = sum(wbytes + U * KV_BYTES_PER_POS * (P + i + 1) for i in range(N_SPEECH - 1))
        d_ms = statistics.mean(st["decode"])
        achieved = dbytes / (d_ms * 1e-3) / 1e9
        step_ms = total_ms / args.steps
        line = {"metric": METRIC, "value": D.whole_job_rate(U * N_SPEECH * args.steps, world, total_ms), "unit": UNIT,
                "n_gpus": world, "steps": args.steps, "warmup": 3, "ms_per_step": step_ms, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": f"configs[3]: {U} concurrent utterances per GPU, data-parallel ranks; each utterance as configs[1] "
                                       "(64+2 text tokens, 150 cond tokens, cfg 0.5, greedy, 256 speech tokens, CFM 10 steps at L=812, "
                                       "HiFT 512 frames), independent seeds; batched T3 decode (one CUDA graph per step), S3Gen per utterance",
                           "utterances_per_step_per_gpu": U, "speech_tokens": N_SPEECH,
                           "l2": "inputs larger than L2 (1.02 GB of weights + the slots' KV slabs streamed per decode step)"},
                "gpu_launches": int(launches), "clocks": clocks,
                "roofline": {"kernel": "batched decode stage (per step: graph of tcgen05 GEMMs + k_b_attn + k_b_sample)",
                             "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": dbytes,
                             "launch_ms": d_ms, "share_of_step": d_ms / step_ms},
                "stages_ms": {k: statistics.mean(v) for k, v in st.items()},
                "t3_decode_tokens_per_s_per_gpu": U * N_SPEECH / (d_ms * 1e-3),
                "t3_tokens_per_s_per_gpu_incl_prefill": U * N_SPEECH / ((d_ms + statistics.mean(st["prefill"])) * 1e-3),
                "ms_per_decode_step": d_ms / N_SPEECH,
                "utterances_per_s": U * args.steps * world / (total_ms * 1e-3),
                "rtf": (step_ms * 1e-3) / (U * AUDIO_S), "kv_splits": eng.info()["kv_splits"], "s3_batch": args.s3_batch}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="config2", choices=["config2", "config4"],
                    help="config2 (default, the headline): one utterance per GPU; config4: --slots concurrent utterances per GPU")
    ap.add_argument("--slots", type=int, default=64, help="config4: concurrent utterances per GPU (1..64)")
    ap.add_argument("--s3-batch", type=int, default=64, help="config4: utterances per batched flow-matching solve (64: 727 ms of S3Gen per 64 utterances; 32: 757; 16: 771)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this build has no CPU path (use --impl reference for the CPU arm)")
    if args.workload == "config4":
        run_config4(args, rank, world, local_rank)
        return
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
