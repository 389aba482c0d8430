#!/usr/bin/env python
"""Benchmark of the Chatterbox inference hot path on B200 (driver contract: one JSON line on rank 0).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...   # the reference's algorithm on the host cores

A "step" is one utterance of BASELINE.json configs[1]: a 64-token (+SOT/EOT) synthetic sentence, a
6 s synthetic reference voice (150 conditioning tokens), cfg_weight 0.5, exaggeration 0.5, greedy,
exactly 256 speech tokens (EOS suppressed), random-init weights of the named architecture.  The
stages that exist in this build run inside the step (see `config.stages`); the headline metric is
T3 speech tokens/s (whole job, all GPUs), with RTF and utterances/s reported beside it.

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
METRIC, UNIT = "t3_speech_tokens_per_s", "tokens/s"


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


def _decode_bytes(weight_bytes: int, P: int, n: int) -> int:
    """Algorithmic HBM bytes of one decode launch of n tokens starting at cache length P:
    per token: every weight once + KV read of the cached context + KV write of one position."""
    return n * weight_bytes + KV_BYTES_PER_POS * sum(P + i for i in range(n)) + KV_BYTES_PER_POS * n


# -------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the oracle port of the reference's algorithm on the host cores
# -------------------------------------------------------------------------------------------------
_SD_CACHE = {}


def cpu_t3_sample(n_decode: int, threads: int):
    """Prefill + n_decode greedy tokens of the full 30-layer model through the oracle restatement of
    T3.inference (oracle/t3_oracle.py; reference t3.py:203-371) in fp32 on `threads` host threads.
    Returns (prefill_s, per_token_s)."""
    from chatterbox_b200 import synthetic as S
    from oracle import t3_oracle as T3O
    torch.set_num_threads(threads)
    if "t3" not in _SD_CACHE:
        _SD_CACHE["t3"] = S.make_state_dict("t3")
    sd = _SD_CACHE["t3"]
    cond = S.t3_cond(prompt_len=PROMPT_LEN)
    text2 = torch.cat([S.text_tokens(N_TEXT)] * 2, 0)
    with torch.inference_mode():
        t0 = time.perf_counter()
        emb = T3O.prepare_input_embeds(sd, cond, text2, SAMPLE["cfg_weight"])
        m = T3O.LlamaOracle(sd)
        t0 = time.perf_counter()
        logits = m.forward(emb)[:, -1]
        t_prefill = time.perf_counter() - t0
        gen = torch.tensor([[T3O.START_SPEECH]], dtype=torch.long)
        t0 = time.perf_counter()
        for i in range(n_decode):
            proc = T3O.process_logits(logits, gen, **SAMPLE)
            nxt = proc.argmax(dim=-1, keepdim=True)
            gen = torch.cat([gen, nxt], dim=1)
            e = (sd["speech_emb.weight"][nxt] + sd["speech_pos_emb.emb.weight"][i + 1]).expand(2, -1, -1)
            logits = m.forward(e)[:, -1]
        t_tok = (time.perf_counter() - t0) / n_decode
    return t_prefill, t_tok


def run_reference(args, rank: int):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n_dec = args.ref_tokens
    times = []
    for i in range(args.warmup + args.steps):
        tp, tt = cpu_t3_sample(n_dec, threads)
        if i >= args.warmup:
            times.append((tp, tt))
    tp = statistics.mean(t[0] for t in times)
    tt = statistics.mean(t[1] for t in times)
    utt_s = tp + N_SPEECH * tt
    val = N_SPEECH / utt_s
    sample = (f"oracle port of T3.inference (fp32, torch CPU, {threads} threads): full prefill (P=102, CFG pair) + "
              f"{n_dec} greedy decode tokens per step, extrapolated to {N_SPEECH} tokens as "
              f"prefill_s + {N_SPEECH} x per_token_s")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": utt_s * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": _config(["t3"]),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def _config(stages):
    return {"workload": "configs[1]: single utterance, 64+2 text tokens, 6 s voice (150 cond tokens), cfg 0.5, "
                        "exaggeration 0.5, greedy, 256 speech tokens, random-init weights",
            "stages": stages, "utterances_per_step_per_gpu": 1, "speech_tokens": N_SPEECH,
            "l2": "inputs larger than L2 (1.02 GB of weights streamed per decode step)"}


# -------------------------------------------------------------------------------------------------
# this repo's arm
# -------------------------------------------------------------------------------------------------
def run_ours(args, rank: int, world: int, local_rank: int):
    import torch.distributed as dist
    from chatterbox_b200 import synthetic as S
    from chatterbox_b200.t3 import T3, T3Cond

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # independent utterance per rank (seeds differ), same shapes
    sd = S.make_state_dict("t3")
    t3 = T3(sd, dev, max_pos=512, kv_fp32=False)
    del sd
    t3.record_events = True
    cond_host = {k: v.pin_memory() for k, v in S.t3_cond(seed=2 + 10 * rank, prompt_len=PROMPT_LEN).items()}
    text_host = torch.cat([S.text_tokens(N_TEXT, seed=1 + rank)] * 2, 0).pin_memory()
    cond_dev = T3Cond(**{k: v.to(dev) for k, v in cond_host.items()})
    text_dev = text_host.to(dev)
    kw = dict(max_new_tokens=N_SPEECH, do_sample=False, stop_on_eos=False, **SAMPLE)

    def step_resident():
        c = T3Cond(speaker_emb=cond_dev.speaker_emb, cond_prompt_speech_tokens=cond_dev.cond_prompt_speech_tokens,
                   emotion_adv=cond_dev.emotion_adv)
        return t3.inference(t3_cond=c, text_tokens=text_dev, **kw)

    def step_e2e():
        c = T3Cond(**{k: v.to(dev, non_blocking=True) for k, v in cond_host.items()})
        toks = t3.inference(t3_cond=c, text_tokens=text_host.to(dev, non_blocking=True), **kw)
        return toks.cpu()

    h2d = sum(v.numel() * v.element_size() for v in cond_host.values()) + text_host.numel() * text_host.element_size()
    d2h = N_SPEECH * 8 + 4

    for _ in range(max(args.warmup, 3)):
        toks = step_resident()
        assert toks.shape[1] == N_SPEECH
    l0 = t3.state()["launches"]

    # ---- timed region 1: inputs resident in HBM ----
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    torch.cuda.profiler.start()     # ncu --profile-from-start off: the launch list covers the timed region only
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dec_ms, pre_ms = [], []
    e0.record()
    for _ in range(args.steps):
        step_resident()
        ev = t3.events
        dec_ms.append((ev["decode0"], ev["decode1"]))
        pre_ms.append((ev["prefill0"], ev["decode0"]))
    e1.record()
    barrier()
    torch.cuda.profiler.stop()
    clocks = sampler.stop()
    total_ms = e0.elapsed_time(e1)
    launches = t3.state()["launches"] - l0
    dec_ms = [a.elapsed_time(b) for a, b in dec_ms]
    pre_ms = [a.elapsed_time(b) for a, b in pre_ms]

    # ---- timed region 2: end to end through the public call with HOST inputs/outputs ----
    for _ in range(2):
        step_e2e()
    barrier()
    e0.record()
    for _ in range(args.steps):
        out = step_e2e()
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    assert out.shape[1] == N_SPEECH

    t = torch.tensor([total_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = t.tolist()

    if rank == 0:
        peak, peak_src = _peaks()
        P = t3.last_prefill_len
        wbytes = t3.packed.weight_bytes_per_step()
        dbytes = _decode_bytes(wbytes, P, N_SPEECH)
        d_ms = statistics.mean(dec_ms)
        achieved = dbytes / (d_ms * 1e-3) / 1e9
        tokens = N_SPEECH * args.steps * world
        line = {"metric": METRIC, "value": tokens / (total_ms * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic", "config": _config(["t3_prefill", "t3_decode"]),
                "e2e": {"value": tokens / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h},
                "gpu_launches": int(launches),
                "clocks": clocks,
                "roofline": {"kernel": "t3_megakernel<bf16> (decode launch: 256 tokens)", "bound": "hbm",
                             "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": None, "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": dbytes, "launch_ms": d_ms},
                "stages_ms": {"t3_prefill": statistics.mean(pre_ms), "t3_decode": d_ms},
                "decode_tokens_per_s": N_SPEECH / (d_ms * 1e-3),
                "utterances_per_s": args.steps * world / (total_ms * 1e-3),
                "rtf_t3_only": (total_ms / args.steps * 1e-3) / (N_SPEECH / 25.0)}
        if not args.no_cpu_baseline and world == 1:
            threads = os.cpu_count() or 1
            tp, tt = cpu_t3_sample(args.ref_tokens, threads)
            v = N_SPEECH / (tp + N_SPEECH * tt)
            line["cpu_baseline"] = {
                "value": v, "unit": UNIT, "cores": threads, "kind": "port",
                "sample": f"oracle port of T3.inference fp32 on {threads} threads: full prefill + {args.ref_tokens} "
                          f"greedy tokens, extrapolated to {N_SPEECH} tokens (prefill {tp:.2f} s, {tt * 1e3:.1f} ms/token)"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ref-tokens", type=int, default=16, help="decode tokens per CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this build has no CPU path (use --impl reference for the CPU arm)")
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
