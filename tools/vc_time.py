"""Config 3 (BASELINE.json): S3Gen-only voice conversion of a 30 s source — 750 speech tokens + a 6 s / 10 s target-voice
prompt -> flow front-end, CFM 10 steps at L = 2 (Tp + 750), HiFT 1500 frames.  Stage times by CUDA events (third call)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chatterbox_b200 import synthetic as S
from chatterbox_b200.s3gen import S3Gen

T = int(sys.argv[1]) if len(sys.argv) > 1 else 750
Tp = int(sys.argv[2]) if len(sys.argv) > 2 else 150
s3 = S3Gen(S.make_state_dict("flow"), S.make_state_dict("hift"), "cuda", rand_noise=S.cfm_noise(), max_mel_len=2 * (T + Tp) + 64)
ref = S.s3gen_ref(seed=4, prompt_len=Tp)
tok = S.speech_tokens(T, seed=5)
s3.record_events = True
for _ in range(3):
    wav, _ = s3.inference(speech_tokens=tok, ref_dict=ref, seed=7)
torch.cuda.synchronize()
ev = s3.events
L = 2 * (T + Tp)
front = ev["front0"].elapsed_time(ev["cfm0"]); cfm = ev["cfm0"].elapsed_time(ev["cfm1"]); hift = ev["hift0"].elapsed_time(ev["hift1"])
flop = 10 * 2 * L * (132.18e6 + 114688.0 * L)
audio = T / 25.0
print(f"VC T={T} Tp={Tp} L={L}: front-end {front:.2f} ms, CFM solve {cfm:.2f} ms ({flop / cfm / 1e9:.1f} TFLOP/s), HiFT {hift:.2f} ms; "
      f"total {front + cfm + hift:.1f} ms for {audio:.1f} s of audio (RTF {(front + cfm + hift) / 1e3 / audio:.4f}), wav {tuple(wav.shape)}")
