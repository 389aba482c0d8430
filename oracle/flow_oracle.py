"""CPU restatement of the S3Gen token front-end (TEST INFRASTRUCTURE — checker / CPU-baseline leg only).

Plain torch fp32, no dependence on `/root/reference`, so it travels to the GPU box.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` / cuda_eager legs may import it; the
product path (`chatterbox_b200/`) never does.

Follows `CausalMaskedDiffWithXvec.inference` (reference src/chatterbox/models/s3gen/flow.py:236-282) up to the
CFM call: x-vector projection, token embedding, `UpsampleConformerEncoder.forward`
(transformer/upsample_encoder.py:237-304; rel-pos attention transformer/attention.py:225-330 incl. `rel_shift`),
`encoder_proj`.  Parity pin: `python -m oracle.make_golden flow` runs the UNMODIFIED reference module on the same
synthetic weights and commits its `mu / spks / cond / feat` as tests/golden/flow_golden.safetensors;
tests/test_flow_frontend.py re-checks this file against them.
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import torch
import torch.nn.functional as F

HEADS, DK, D = 8, 64, 512


def _rel_pos_emb(T: int, device, dtype) -> torch.Tensor:
    """EspnetRelPositionalEncoding.position_encoding (embedding.py:224-294): (1, 2T-1, 512), relative
    positions T-1 ... -(T-1)."""
    pos = torch.arange(T - 1, -T, -1, dtype=torch.float32, device=device).unsqueeze(1)
    div = torch.exp(torch.arange(0, D, 2, dtype=torch.float32, device=device) * -(math.log(10000.0) / D))
    pe = torch.zeros(2 * T - 1, D, device=device)
    pe[:, 0::2] = torch.sin(pos * div)
    pe[:, 1::2] = torch.cos(pos * div)
    return pe.unsqueeze(0).to(dtype)


def _rel_shift(x: torch.Tensor) -> torch.Tensor:
    """attention.py:225-247."""
    b, h, t, n = x.shape
    xp = torch.cat([torch.zeros(b, h, t, 1, device=x.device, dtype=x.dtype), x], dim=-1)
    xp = xp.view(b, h, n + 1, t)
    return xp[:, :, 1:].reshape(b, h, t, n)[:, :, :, : n // 2 + 1]


def _layer(sd, p: str, x: torch.Tensor, pos_emb: torch.Tensor) -> torch.Tensor:
    """ConformerEncoderLayer (encoder_layer.py:155-236) without macaron / conv module, pre-norm, eps 1e-12."""
    B, T, _ = x.shape
    lin = lambda n, v, bias=True: F.linear(v, sd[p + n + ".weight"], sd[p + n + ".bias"] if bias else None)
    h = F.layer_norm(x, (D,), sd[p + "norm_mha.weight"], sd[p + "norm_mha.bias"], 1e-12)
    q = lin("self_attn.linear_q", h).view(B, T, HEADS, DK)
    k = lin("self_attn.linear_k", h).view(B, T, HEADS, DK).transpose(1, 2)
    v = lin("self_attn.linear_v", h).view(B, T, HEADS, DK).transpose(1, 2)
    pe = lin("self_attn.linear_pos", pos_emb, bias=False).view(1, -1, HEADS, DK).transpose(1, 2)
    qu = (q + sd[p + "self_attn.pos_bias_u"]).transpose(1, 2)
    qv = (q + sd[p + "self_attn.pos_bias_v"]).transpose(1, 2)
    ac = torch.matmul(qu, k.transpose(-2, -1))
    bd = _rel_shift(torch.matmul(qv, pe.transpose(-2, -1)))
    attn = torch.softmax((ac + bd) / math.sqrt(DK), dim=-1)
    o = torch.matmul(attn, v).transpose(1, 2).reshape(B, T, D)
    x = x + lin("self_attn.linear_out", o)
    h = F.layer_norm(x, (D,), sd[p + "norm_ff.weight"], sd[p + "norm_ff.bias"], 1e-12)
    return x + lin("feed_forward.w_2", F.silu(lin("feed_forward.w_1", h)))


def _embed(sd, p: str, x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """LinearNoSubsampling + EspnetRelPositionalEncoding (subsampling.py:69-113, embedding.py:256-268)."""
    x = F.linear(x, sd[p + "out.0.weight"], sd[p + "out.0.bias"])
    x = F.layer_norm(x, (D,), sd[p + "out.1.weight"], sd[p + "out.1.bias"], 1e-5)
    return x * math.sqrt(D), _rel_pos_emb(x.shape[1], x.device, x.dtype)


def encoder(sd: Dict[str, torch.Tensor], xs: torch.Tensor, pfx: str = "encoder.") -> torch.Tensor:
    """UpsampleConformerEncoder.forward for one un-padded utterance: (1, T, 512) -> (1, 2T, 512)."""
    x, pos = _embed(sd, pfx + "embed.", xs)
    # PreLookaheadLayer (upsample_encoder.py:66-96): right context 3, then causal k3, residual
    h = x.transpose(1, 2)
    h = F.leaky_relu(F.conv1d(F.pad(h, (0, 3)), sd[pfx + "pre_lookahead_layer.conv1.weight"],
                              sd[pfx + "pre_lookahead_layer.conv1.bias"]))
    h = F.conv1d(F.pad(h, (2, 0)), sd[pfx + "pre_lookahead_layer.conv2.weight"], sd[pfx + "pre_lookahead_layer.conv2.bias"])
    x = h.transpose(1, 2) + x
    for i in range(6):
        x = _layer(sd, pfx + f"encoders.{i}.", x, pos)
    # Upsample1D (upsample_encoder.py:37-63): nearest x2, left pad 4, conv k5
    h = F.interpolate(x.transpose(1, 2), scale_factor=2.0, mode="nearest")
    h = F.conv1d(F.pad(h, (4, 0)), sd[pfx + "up_layer.conv.weight"], sd[pfx + "up_layer.conv.bias"])
    x, pos = _embed(sd, pfx + "up_embed.", h.transpose(1, 2))
    for i in range(4):
        x = _layer(sd, pfx + f"up_encoders.{i}.", x, pos)
    return F.layer_norm(x, (D,), sd[pfx + "after_norm.weight"], sd[pfx + "after_norm.bias"], 1e-5)


def flow_inputs(sd: Dict[str, torch.Tensor], token: torch.Tensor, prompt_token: torch.Tensor, prompt_feat: torch.Tensor,
                embedding: torch.Tensor):
    """flow.py:250-270 -> (mu (1,80,L), spks (1,80), cond (1,80,L), mel_len1) for the CFM solve.
    token (1,T) / prompt_token (1,Tp) int64, prompt_feat (1, 2Tp, 80), embedding (1,192)."""
    emb = F.normalize(embedding, dim=1)
    spks = F.linear(emb, sd["spk_embed_affine_layer.weight"], sd["spk_embed_affine_layer.bias"])
    tok = torch.cat([prompt_token, token], dim=1)
    x = sd["input_embedding.weight"][torch.clamp(tok, min=0)]
    h = encoder(sd, x)
    mel_len1 = prompt_feat.shape[1]
    L = h.shape[1]
    mu = F.linear(h, sd["encoder_proj.weight"], sd["encoder_proj.bias"])
    cond = torch.zeros(1, L, 80, device=h.device, dtype=h.dtype)
    cond[:, :mel_len1] = prompt_feat
    return mu.transpose(1, 2).contiguous(), spks, cond.transpose(1, 2).contiguous(), mel_len1
