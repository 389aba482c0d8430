"""Token -> (mu, spks, cond) front-end of the S3Gen flow module over the C-ABI (`cbx_flowfe_*`, csrc/s3_front.cu): the part
of `CausalMaskedDiffWithXvec.inference` (reference src/chatterbox/models/s3gen/flow.py:236-282) that runs ONCE per utterance
before the flow-matching loop — x-vector projection, token embedding, the CosyVoice2 `UpsampleConformerEncoder`
(transformer/upsample_encoder.py:237-304) and `encoder_proj`.

State-dict keys are the reference's (`input_embedding.*`, `spk_embed_affine_layer.*`, `encoder.*`, `encoder_proj.*`), so a
real `s3gen.safetensors` packs through the same function.  There is no CPU path (the torch restatement of this stage lives in
`oracle/flow_oracle.py`, test infrastructure only).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List

import torch

from . import _lib

N_LAYERS0, N_LAYERS1 = 6, 4


def _conv_as_gemm(w: torch.Tensor) -> torch.Tensor:
    """Conv1d weight (Cout, Cin, k) -> GEMM matrix [Cout][k * Cin] with column index tap * Cin + ci."""
    return w.permute(0, 2, 1).reshape(w.shape[0], -1)


def pack_front(sd: Dict[str, torch.Tensor], device):
    """-> (bf16 blob, f32 blob, bf16 offsets, f32 offsets) in the order csrc/s3_front.cu indexes (BF_* / F_* enums)."""
    g = lambda k: sd[k].detach().to(torch.float32)
    layers = [f"encoder.encoders.{i}." for i in range(N_LAYERS0)] + [f"encoder.up_encoders.{i}." for i in range(N_LAYERS1)]
    bf: List[torch.Tensor] = [g("encoder.embed.out.0.weight"), _conv_as_gemm(g("encoder.pre_lookahead_layer.conv1.weight")),
                              _conv_as_gemm(g("encoder.pre_lookahead_layer.conv2.weight"))]
    f: List[torch.Tensor] = [g("input_embedding.weight"), g("encoder.embed.out.0.bias"), g("encoder.embed.out.1.weight"),
                             g("encoder.embed.out.1.bias"), g("encoder.pre_lookahead_layer.conv1.bias"),
                             g("encoder.pre_lookahead_layer.conv2.bias")]
    for p in layers:
        a = p + "self_attn."
        bf += [torch.cat([g(a + "linear_q.weight"), g(a + "linear_k.weight"), g(a + "linear_v.weight")], 0), g(a + "linear_pos.weight"),
               g(a + "linear_out.weight"), g(p + "feed_forward.w_1.weight"), g(p + "feed_forward.w_2.weight")]
        f += [g(p + "norm_mha.weight"), g(p + "norm_mha.bias"),
              torch.cat([g(a + "linear_q.bias"), g(a + "linear_k.bias"), g(a + "linear_v.bias")], 0), g(a + "linear_out.bias"),
              g(a + "pos_bias_u").reshape(-1), g(a + "pos_bias_v").reshape(-1), g(p + "norm_ff.weight"), g(p + "norm_ff.bias"),
              g(p + "feed_forward.w_1.bias"), g(p + "feed_forward.w_2.bias")]
    bf += [_conv_as_gemm(g("encoder.up_layer.conv.weight")), g("encoder.up_embed.out.0.weight"), g("encoder_proj.weight")]
    f += [g("encoder.up_layer.conv.bias"), g("encoder.up_embed.out.0.bias"), g("encoder.up_embed.out.1.weight"),
          g("encoder.up_embed.out.1.bias"), g("encoder.after_norm.weight"), g("encoder.after_norm.bias"), g("encoder_proj.bias"),
          g("spk_embed_affine_layer.weight"), g("spk_embed_affine_layer.bias")]
    assert len(bf) == 56 and len(f) == 115

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


class FlowFrontend:
    """`flow_inputs(token, prompt_token, prompt_feat, embedding) -> (mu (1,80,L), spks (1,80), cond (1,80,L), mel_len1)`:
    flow.py:250-270."""

    def __init__(self, flow_sd: Dict[str, torch.Tensor], device="cuda", *, max_tokens: int = 2048):
        if not torch.cuda.is_available():
            raise _lib.CbxError("chatterbox_b200.flow_frontend needs a CUDA device (no CPU fallback)")
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self._lib = _lib.lib()
        self.bf, self.f32, bfo, fo = pack_front(flow_sd, self.device)
        self._bfo = (C.c_int64 * len(bfo))(*bfo)
        self._fo = (C.c_int64 * len(fo))(*fo)
        w = _lib.CbxFlowFrontWeights(self.bf.data_ptr(), self.f32.data_ptr(), self._bfo, self._fo)
        h = C.c_void_p()
        _lib.check(self._lib.cbx_flowfe_create(self.device.index, int(max_tokens), C.byref(w), C.byref(h)))
        self._h = h
        self.max_tokens = max_tokens

    def sibling(self) -> "FlowFrontend":
        """A second handle (own workspace) over the SAME packed weights, for a second stream (S3Gen.flow_inference_batch)."""
        o = object.__new__(FlowFrontend)
        o.device, o._lib, o.bf, o.f32, o._bfo, o._fo, o.max_tokens = self.device, self._lib, self.bf, self.f32, self._bfo, self._fo, self.max_tokens
        w = _lib.CbxFlowFrontWeights(o.bf.data_ptr(), o.f32.data_ptr(), o._bfo, o._fo)
        h = C.c_void_p()
        _lib.check(o._lib.cbx_flowfe_create(o.device.index, int(o.max_tokens), C.byref(w), C.byref(h)))
        o._h = h
        return o

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.cbx_flowfe_destroy(h)
            self._h = None

    def launches(self) -> int:
        n = C.c_int64()
        _lib.check(self._lib.cbx_flowfe_launches(self._h, C.byref(n)))
        return n.value

    @torch.inference_mode()
    def flow_inputs(self, token: torch.Tensor, prompt_token: torch.Tensor, prompt_feat: torch.Tensor, embedding: torch.Tensor):
        dev = self.device
        tok = token.to(dev, torch.long).reshape(-1).contiguous()
        ptk = prompt_token.to(dev, torch.long).reshape(-1).contiguous()
        pf = prompt_feat.to(dev, torch.float32).reshape(-1, 80).contiguous()
        emb = embedding.to(dev, torch.float32).reshape(-1).contiguous()
        T, Tp, Lp = tok.numel(), ptk.numel(), pf.shape[0]
        L = 2 * (T + Tp)
        mu = torch.empty(1, 80, L, dtype=torch.float32, device=dev)
        cond = torch.empty(1, 80, L, dtype=torch.float32, device=dev)
        spks = torch.empty(1, 80, dtype=torch.float32, device=dev)
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(self._lib.cbx_flowfe_run(self._h, ptk.data_ptr() if Tp else None, Tp, tok.data_ptr(), T,
                                            pf.data_ptr() if Lp else None, Lp, emb.data_ptr(), mu.data_ptr(), spks.data_ptr(),
                                            cond.data_ptr(), stream))
        return mu, spks, cond, Lp
