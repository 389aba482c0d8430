"""Drop-in mirror of the reference voice-conversion API (`src/chatterbox/vc.py`): `ChatterboxVC.from_pretrained /
from_local / set_target_voice / generate(audio, target_voice_path)`.  S3Gen (flow matching + vocoder) runs in
`libcbx.so`.  The S3 speech tokenizer (third-party `s3tokenizer` network, absent here) is the VC front-end and
outside the accelerated path: `generate(..., speech_tokens=ids)` accepts pre-computed 25 Hz tokens."""
from __future__ import annotations

from pathlib import Path
from typing import Optional

import torch

from .s3gen import S3GEN_SR, S3Gen

REPO_ID = "ResembleAI/chatterbox"
S3_SR = 16000


class ChatterboxVC:
    ENC_COND_LEN = 6 * S3_SR
    DEC_COND_LEN = 10 * S3GEN_SR

    def __init__(self, s3gen: S3Gen, device, ref_dict: Optional[dict] = None):
        self.sr = S3GEN_SR
        self.s3gen = s3gen
        self.device = device
        try:
            import perth
            self.watermarker = perth.PerthImplicitWatermarker()
        except Exception:
            self.watermarker = None
        self.ref_dict = None if ref_dict is None else {k: v.to(device) if torch.is_tensor(v) else v for k, v in ref_dict.items()}

    @classmethod
    def from_local(cls, ckpt_dir, device) -> "ChatterboxVC":
        from safetensors.torch import load_file
        ckpt_dir = Path(ckpt_dir)
        ref_dict = None
        if (ckpt_dir / "conds.pt").exists():
            ref_dict = torch.load(ckpt_dir / "conds.pt", map_location="cpu", weights_only=True)["gen"]
        s3gen = S3Gen.from_state_dict(load_file(ckpt_dir / "s3gen.safetensors"), device)
        return cls(s3gen, device, ref_dict=ref_dict)

    @classmethod
    def from_pretrained(cls, device) -> "ChatterboxVC":
        from huggingface_hub import hf_hub_download
        for fpath in ["s3gen.safetensors", "conds.pt"]:
            local_path = hf_hub_download(repo_id=REPO_ID, filename=fpath)
        return cls.from_local(Path(local_path).parent, device)

    def set_target_voice(self, wav_fpath):
        raise NotImplementedError("set_target_voice needs S3Gen.embed_ref (conditioning prep, outside the accelerated "
                                  "path); construct ChatterboxVC with a pre-computed ref_dict")

    def generate(self, audio=None, target_voice_path=None, *, speech_tokens: Optional[torch.Tensor] = None,
                 seed: Optional[int] = None):
        if target_voice_path:
            self.set_target_voice(target_voice_path)
        else:
            assert self.ref_dict is not None, "Please `prepare_conditionals` first or specify `target_voice_path`"
        if speech_tokens is None:
            raise NotImplementedError("tokenising source audio needs the third-party S3TokenizerV2 network; pass "
                                      "speech_tokens=<25 Hz S3 token ids>")
        with torch.inference_mode():
            wav, _ = self.s3gen.inference(speech_tokens=speech_tokens, ref_dict=self.ref_dict, seed=seed)
            wav = wav.squeeze(0).detach().cpu()
        if self.watermarker is not None:
            wav = torch.from_numpy(self.watermarker.apply_watermark(wav.numpy(), sample_rate=self.sr))
        return wav.unsqueeze(0)
