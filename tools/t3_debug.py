"""Bring-up diagnostics for the T3 megakernel (run on the GPU box): staged comparison with the CPU
oracle at 1, 2 and 30 layers, fp32-KV and bf16-KV, prefill chunk edge cases.  Writes JSON lines to
gpurun_out/t3_debug.log.  Test infrastructure (imports oracle/)."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from chatterbox_b200 import synthetic as S
from chatterbox_b200.t3 import T3, T3Cond
from oracle import t3_oracle as T3O

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
os.makedirs(OUT, exist_ok=True)
LOG = open(os.path.join(OUT, "t3_debug.log"), "a")


def log(**kw):
    s = json.dumps(kw)
    print(s, flush=True)
    LOG.write(s + "\n")
    LOG.flush()


def run_case(n_layers, kv_fp32, n_text, n_tok, cfg_weight, sd):
    text = S.text_tokens(n_text)
    tt = torch.cat([text, text], 0) if cfg_weight > 0 else text
    cond = S.t3_cond()
    g = torch.Generator().manual_seed(5)
    forced = torch.randint(0, 6561, (1, n_tok), generator=g)
    params = dict(cfg_weight=cfg_weight, temperature=0.8, repetition_penalty=1.2, min_p=0.05, top_p=1.0)
    t0 = time.time()
    _, ref = T3O.inference(sd, cond, tt, max_new_tokens=n_tok, forced_tokens=forced, n_layers=n_layers,
                           kv_round=not kv_fp32, **params)
    t_or = time.time() - t0
    m = T3(sd, "cuda", max_pos=512, kv_fp32=kv_fp32, n_layers=n_layers)
    c = T3Cond(**{k: v.clone() for k, v in cond.items()}).to(device="cuda")
    t0 = time.time()
    toks = m.inference(t3_cond=c, text_tokens=tt, max_new_tokens=n_tok, forced_tokens=forced, return_logits=True,
                       do_sample=False, **params)
    torch.cuda.synchronize()
    t_gpu = time.time() - t0
    got = m.last_logits.float().cpu()
    errs = [(got[i] - ref[i]).abs().max().item() for i in range(got.shape[0])]
    log(case="forced", n_layers=n_layers, kv_fp32=kv_fp32, n_text=n_text, P=34 + n_text + 2 + (2 if cfg_weight > 0 else 1),
        cfg=cfg_weight, n=int(toks.shape[1]), err_first=errs[0], err_max=max(errs), err_last=errs[-1],
        ref_absmax=ref.abs().max().item(), state=m.state(), t_oracle=round(t_or, 2), t_gpu=round(t_gpu, 3))
    del m
    return max(errs)


def main():
    torch.manual_seed(0)
    log(case="env", gpu=torch.cuda.get_device_name(0), sms=torch.cuda.get_device_properties(0).multi_processor_count)
    stages = [(1, True, 2, 3, 0.5), (1, True, 3, 3, 0.0), (2, True, 5, 4, 0.5), (2, False, 5, 4, 0.5)]
    for (nl, kvf, nt, ntok, cfg) in stages:
        sd = S.make_state_dict("t3", n_layers=nl)
        try:
            e = run_case(nl, kvf, nt, ntok, cfg, sd)
        except Exception as ex:  # noqa
            log(case="exception", n_layers=nl, kv_fp32=kvf, msg=repr(ex)[:500])
            return 1
        if not (e < 5e-2):
            log(case="stop", reason="error too large, later stages skipped", err=e)
            return 1
    if "--full" in sys.argv:
        sd = S.make_state_dict("t3")
        run_case(30, True, 64, 8, 0.5, sd)
        run_case(30, False, 64, 8, 0.5, sd)
    return 0


if __name__ == "__main__":
    sys.exit(main())
