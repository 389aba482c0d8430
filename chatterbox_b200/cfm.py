"""Host-side mirror of the reference's S3Gen flow-matching decoder over the C-ABI.

`CausalConditionalCFM.forward(mu, mask, n_timesteps, temperature, spks, cond)` and
`ConditionalDecoder.forward(x, mask, mu, t, spks, cond)` keep the reference signatures
(src/chatterbox/models/s3gen/flow_matching.py:194-218, decoder.py:233-317); the estimator's 14 resnet
blocks, 56 transformer blocks and the Euler/CFG loop run in `libcbx.so` (csrc/s3_cfm.cu: tcgen05 GEMMs
with fused epilogues + flash attention).  State-dict keys are the reference's (`decoder.estimator.*`
inside the flow module), so a real `s3gen.safetensors` loads through the same packer.
There is no CPU path: construction fails without CUDA + libcbx.so.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import torch

from . import _lib

NRES, NTFM = 14, 56
CFG_RATE = 0.7     # reference configs.py:8 (inference_cfg_rate)


def _res_prefixes() -> List[str]:
    return ["down_blocks.0.0."] + [f"mid_blocks.{m}.0." for m in range(12)] + ["up_blocks.0.0."]


def _tfm_prefixes() -> List[str]:
    out = [f"down_blocks.0.1.{j}." for j in range(4)]
    for m in range(12):
        out += [f"mid_blocks.{m}.1.{j}." for j in range(4)]
    return out + [f"up_blocks.0.1.{j}." for j in range(4)]


def _conv_as_gemm(w: torch.Tensor) -> torch.Tensor:
    """Conv1d weight (Cout, Cin, k) -> GEMM matrix [Cout][k*Cin] with column index tap*Cin + ci."""
    return w.permute(0, 2, 1).reshape(w.shape[0], -1)


def pack_estimator(sd: Dict[str, torch.Tensor], prefix: str, device):
    """-> (bf16 blob, f32 blob, bf16 offsets, f32 offsets) in the order csrc/s3_cfm.cu indexes."""
    g = lambda k: sd[prefix + k].detach().to(torch.float32)
    bf: List[torch.Tensor] = []
    f: List[torch.Tensor] = []
    for p in _res_prefixes():
        bf += [_conv_as_gemm(g(p + "block1.block.0.weight")), _conv_as_gemm(g(p + "block2.block.0.weight")),
               _conv_as_gemm(g(p + "res_conv.weight"))]
        f += [g(p + "block1.block.0.bias"), g(p + "block1.block.2.weight"), g(p + "block1.block.2.bias"),
              g(p + "block2.block.0.bias"), g(p + "block2.block.2.weight"), g(p + "block2.block.2.bias"),
              g(p + "res_conv.bias")]
    for p in _tfm_prefixes():
        bf += [torch.cat([g(p + "attn1.to_q.weight"), g(p + "attn1.to_k.weight"), g(p + "attn1.to_v.weight")], 0),
               g(p + "attn1.to_out.0.weight"), g(p + "ff.net.0.proj.weight"), g(p + "ff.net.2.weight")]
        f += [g(p + "norm1.weight"), g(p + "norm1.bias"), g(p + "attn1.to_out.0.bias"), g(p + "norm3.weight"),
              g(p + "norm3.bias"), g(p + "ff.net.0.proj.bias"), g(p + "ff.net.2.bias")]
    bf += [_conv_as_gemm(g("down_blocks.0.2.weight")), _conv_as_gemm(g("up_blocks.0.2.weight")),
           _conv_as_gemm(g("final_block.block.0.weight")), _conv_as_gemm(g("final_proj.weight"))]
    f += [g("down_blocks.0.2.bias"), g("up_blocks.0.2.bias"), g("final_block.block.0.bias"),
          g("final_block.block.2.weight"), g("final_block.block.2.bias"), g("final_proj.bias")]
    f += [g("time_mlp.linear_1.weight"), g("time_mlp.linear_1.bias"), g("time_mlp.linear_2.weight"),
          g("time_mlp.linear_2.bias"),
          torch.stack([g(p + "mlp.1.weight") for p in _res_prefixes()]),
          torch.stack([g(p + "mlp.1.bias") for p in _res_prefixes()])]
    assert len(bf) == NRES * 3 + NTFM * 4 + 4 and len(f) == NRES * 7 + NTFM * 7 + 6 + 6

    def blob(ts, dtype, align):
        offs, n = [], 0
        for t in ts:
            offs.append(n)
            n += (t.numel() + align - 1) // align * align
        out = torch.zeros(n, dtype=dtype, device=device)
        for o, t in zip(offs, ts):
            out[o:o + t.numel()] = t.reshape(-1).to(device=device, dtype=dtype)
        return out, offs

    bfb, bfo = blob(bf, torch.bfloat16, 64)       # 128-byte aligned matrices (TMA global address alignment)
    fb, fo = blob(f, torch.float32, 4)
    return bfb, fb, bfo, fo


def euler_schedule(n_timesteps: int):
    """t fed to every step and its dt, with the reference's own fp32 arithmetic
    (flow_matching.py:215-217 cosine schedule; :90,118-121 running t / dt)."""
    t_span = torch.linspace(0, 1, n_timesteps + 1, dtype=torch.float32)
    t_span = 1 - torch.cos(t_span * 0.5 * torch.pi)
    t, dt = t_span[0], t_span[1] - t_span[0]
    ts, dts = [], []
    for step in range(1, n_timesteps + 1):
        ts.append(float(t))
        dts.append(float(dt))
        t = t + dt
        if step < n_timesteps:
            dt = t_span[step + 1] - t
    return ts, dts


def _check_mask(mask) -> None:
    """One un-padded utterance = an all-ones mask (decoder.py:25-33 then adds a zero bias).  `None` means exactly
    that and costs nothing; a host tensor is checked on the host; a DEVICE tensor is checked too, which is a host
    sync — the in-tree callers (s3gen.py) pass None."""
    if mask is not None:
        assert bool((mask == 1).all()), "only un-padded utterances (all-ones mask) are supported"


class ConditionalDecoder:
    """The estimator (decoder.py:100-317).  forward(x, mask, mu, t, spks, cond) with batch 2."""

    def __init__(self, state_dict: Dict[str, torch.Tensor], device="cuda", *, prefix: str = "", max_len: int = 4096,
                 max_batch: int = 1):
        if not torch.cuda.is_available():
            raise _lib.CbxError("chatterbox_b200.cfm needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self._lib = _lib.lib()
        self.bf, self.f32, bfo, fo = pack_estimator(state_dict, prefix, self.device)
        self._bfo = (C.c_int64 * len(bfo))(*bfo)
        self._fo = (C.c_int64 * len(fo))(*fo)
        w = _lib.CbxCfmWeights(self.bf.data_ptr(), self.f32.data_ptr(), self._bfo, self._fo)
        h = C.c_void_p()
        _lib.check(self._lib.cbx_cfm_create_batched(self.device.index, int(max_len), int(max_batch), C.byref(w), C.byref(h)))
        self._h = h
        self.max_len = max_len
        self.max_batch = max_batch

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.cbx_cfm_destroy(h)
            self._h = None

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def launches(self) -> int:
        n = C.c_int64()
        _lib.check(self._lib.cbx_cfm_launches(self._h, C.byref(n)))
        return n.value

    def cache_sizes(self):
        g, p = C.c_int32(), C.c_int32()
        _lib.check(self._lib.cbx_cfm_cache_sizes(self._h, C.byref(g), C.byref(p)))
        return g.value, p.value

    @torch.inference_mode()
    def forward(self, x, mask, mu, t, spks=None, cond=None):
        assert x.shape[0] == 2 and x.shape[1] == 80, "estimator runs the CFG pair (batch 2)"
        _check_mask(mask)
        assert float(t[0]) == float(t[1])
        L = x.shape[2]
        f = lambda a: a.to(self.device, torch.float32).contiguous()
        x, mu, spks, cond = f(x), f(mu), f(spks), f(cond)
        out = torch.empty(2, 80, L, dtype=torch.float32, device=self.device)
        _lib.check(self._lib.cbx_cfm_estimator(self._h, x.data_ptr(), mu.data_ptr(), spks.data_ptr(), cond.data_ptr(),
                                               float(t[0]), L, out.data_ptr(), self._stream()))
        return out

    __call__ = forward


class CausalConditionalCFM:
    """flow_matching.py:188-218.  `rand_noise` is the fixed (1, 80, 15000) noise tensor of the reference
    (drawn once at construction there; injectable here so parity runs share it)."""

    def __init__(self, estimator: ConditionalDecoder, rand_noise: Optional[torch.Tensor] = None,
                 inference_cfg_rate: float = CFG_RATE):
        self.estimator = estimator
        self.device = estimator.device
        self.inference_cfg_rate = inference_cfg_rate
        if rand_noise is None:
            rand_noise = torch.randn([1, 80, 50 * 300])
        self.rand_noise = rand_noise.to(self.device, torch.float32)
        self._sched = {}

    @torch.inference_mode()
    def forward(self, mu, mask, n_timesteps, temperature=1.0, spks=None, cond=None):
        assert mu.shape[0] == 1 and mu.shape[1] == 80
        _check_mask(mask)
        L = mu.shape[2]
        e = self.estimator
        f = lambda a: a.to(self.device, torch.float32).contiguous()
        z = (self.rand_noise[:, :, :L] * temperature).contiguous()
        mu, spks, cond = f(mu), f(spks), f(cond)
        if n_timesteps not in self._sched:
            ts, dts = euler_schedule(n_timesteps)
            self._sched[n_timesteps] = ((C.c_float * n_timesteps)(*ts), (C.c_float * n_timesteps)(*dts))
        ts, dts = self._sched[n_timesteps]
        out = torch.empty(1, 80, L, dtype=torch.float32, device=self.device)
        _lib.check(e._lib.cbx_cfm_solve(e._h, z.data_ptr(), mu.data_ptr(), spks.data_ptr(), cond.data_ptr(), L,
                                        int(n_timesteps), ts, dts, float(self.inference_cfg_rate), out.data_ptr(),
                                        e._stream()))
        return out, None

    __call__ = forward

    @torch.inference_mode()
    def forward_batch(self, mu, lens, n_timesteps, temperature=1.0, spks=None, cond=None):
        """The same solve for nb utterances at once (BASELINE config 4): mu, cond (nb, 80, L) padded to the longest
        utterance, spks (nb, 80), lens = true mel lengths (list of ints, or None when all are L).  This is the reference's
        padded-batch call `forward(mu, mask=~make_pad_mask(lens), ...)` (flow_matching.py:194-218, utils/mask.py:167-193):
        every utterance starts from the same `rand_noise[:, :, :L]`.  Frames past lens[u] of the result are unspecified
        (the reference zeroes them): slice them off."""
        nb, _, L = mu.shape
        e = self.estimator
        assert mu.shape[1] == 80 and nb <= e.max_batch, (mu.shape, e.max_batch)
        f = lambda a: a.to(self.device, torch.float32).contiguous()
        z = (self.rand_noise[:, :, :L] * temperature).expand(nb, -1, -1).contiguous()
        mu, spks, cond = f(mu), f(spks), f(cond)
        if n_timesteps not in self._sched:
            ts, dts = euler_schedule(n_timesteps)
            self._sched[n_timesteps] = ((C.c_float * n_timesteps)(*ts), (C.c_float * n_timesteps)(*dts))
        ts, dts = self._sched[n_timesteps]
        larr = None
        if lens is not None and any(int(x) != L for x in lens):
            larr = (C.c_int32 * nb)(*[int(x) for x in lens])
        out = torch.empty(nb, 80, L, dtype=torch.float32, device=self.device)
        _lib.check(e._lib.cbx_cfm_solve_batch(e._h, nb, z.data_ptr(), mu.data_ptr(), spks.data_ptr(), cond.data_ptr(), larr, L,
                                              int(n_timesteps), ts, dts, float(self.inference_cfg_rate), out.data_ptr(),
                                              e._stream()))
        return out
