"""CPU restatement of the reference's S3Gen conditional flow-matching path (TEST INFRASTRUCTURE).

Plain torch fp32 on CPU, no dependence on `/root/reference`, `diffusers` or `einops`, so it travels
to the GPU box.  Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline /
`--impl reference` legs may import it; the product path (`chatterbox_b200/`) never does.

Parity pin: `oracle/make_golden.py cfm` (build container) checks every function here against the
UNMODIFIED reference modules (`decoder.ConditionalDecoder`, `flow_matching.CausalConditionalCFM`)
with identical synthetic weights and commits the reference's outputs under `tests/golden/`;
`tests/test_oracle_cfm.py` re-checks this restatement against those files.  The reference ships no
golden vectors for this path.

What each function follows (paths relative to /root/reference/src/chatterbox/models/s3gen):
  time_embedding     matcha/decoder.py:14-29 (SinusoidalPosEmb, scale 1000) + :73-117 (TimestepEmbedding, SiLU)
  causal_block       decoder.py:48-61 (CausalBlock1D: left-pad-2 conv k3 -> LayerNorm over C -> Mish), :71-97
  resnet             matcha/decoder.py:46-61 via decoder.py:64-68 (CausalResnetBlock1D)
  transformer_block  matcha/transformer.py:243-316 (norm1 -> attn1 -> +res -> norm3 -> ff -> +res) with
                     diffusers 0.29.0 Attention (q/k/v no bias, 8 heads x 64, SDPA, out bias) and GELU (erf)
  estimator          decoder.py:233-317 (ConditionalDecoder.forward; 1 down + 12 mid + 1 up block)
  solve_euler        flow_matching.py:72-123 (batch-2 CFG, rate 0.7) + :194-218 (cosine t schedule)
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

HEADS, HEAD_DIM = 8, 64


def time_embedding(sd: Dict[str, torch.Tensor], t: torch.Tensor, pfx: str = "") -> torch.Tensor:
    """t (B,) -> (B, 1024)."""
    half = 160
    e = math.log(10000) / (half - 1)
    e = torch.exp(torch.arange(half, device=t.device).float() * -e)
    e = 1000 * t.unsqueeze(1) * e.unsqueeze(0)
    e = torch.cat((e.sin(), e.cos()), dim=-1)
    h = F.linear(e, sd[pfx + "time_mlp.linear_1.weight"], sd[pfx + "time_mlp.linear_1.bias"])
    return F.linear(F.silu(h), sd[pfx + "time_mlp.linear_2.weight"], sd[pfx + "time_mlp.linear_2.bias"])


def causal_conv(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    return F.conv1d(F.pad(x, (w.shape[-1] - 1, 0)), w, b)


def causal_block(sd, p: str, x: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    h = causal_conv(x * mask, sd[p + "block.0.weight"], sd[p + "block.0.bias"])
    c = h.shape[1]
    h = F.layer_norm(h.transpose(1, 2), (c,), sd[p + "block.2.weight"], sd[p + "block.2.bias"]).transpose(1, 2)
    return F.mish(h) * mask


def resnet(sd, p: str, x: torch.Tensor, mask: torch.Tensor, temb: torch.Tensor) -> torch.Tensor:
    h = causal_block(sd, p + "block1.", x, mask)
    h = h + F.linear(F.mish(temb), sd[p + "mlp.1.weight"], sd[p + "mlp.1.bias"]).unsqueeze(-1)
    h = causal_block(sd, p + "block2.", h, mask)
    return h + F.conv1d(x * mask, sd[p + "res_conv.weight"], sd[p + "res_conv.bias"])


def transformer_block(sd, p: str, x: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
    """x (B, T, 256); bias additive (B, 1, T) or None."""
    B, T, C = x.shape
    n = F.layer_norm(x, (C,), sd[p + "norm1.weight"], sd[p + "norm1.bias"])
    q = F.linear(n, sd[p + "attn1.to_q.weight"]).view(B, T, HEADS, HEAD_DIM).transpose(1, 2)
    k = F.linear(n, sd[p + "attn1.to_k.weight"]).view(B, T, HEADS, HEAD_DIM).transpose(1, 2)
    v = F.linear(n, sd[p + "attn1.to_v.weight"]).view(B, T, HEADS, HEAD_DIM).transpose(1, 2)
    am = None if bias is None else bias[:, None]           # (B, 1, 1, T) broadcast over heads and queries
    o = F.scaled_dot_product_attention(q, k, v, attn_mask=am)
    o = o.transpose(1, 2).reshape(B, T, HEADS * HEAD_DIM)
    x = x + F.linear(o, sd[p + "attn1.to_out.0.weight"], sd[p + "attn1.to_out.0.bias"])
    n = F.layer_norm(x, (C,), sd[p + "norm3.weight"], sd[p + "norm3.bias"])
    h = F.gelu(F.linear(n, sd[p + "ff.net.0.proj.weight"], sd[p + "ff.net.0.proj.bias"]))
    return x + F.linear(h, sd[p + "ff.net.2.weight"], sd[p + "ff.net.2.bias"])


def estimator(sd, x, mask, mu, t, spks, cond, pfx: str = "", n_mid: int = 12, n_tfm: int = 4) -> torch.Tensor:
    """ConditionalDecoder.forward: x, mu, cond (B,80,T); mask (B,1,T); t (B,); spks (B,80) -> (B,80,T)."""
    temb = time_embedding(sd, t, pfx)
    T = x.shape[-1]
    h = torch.cat([x, mu, spks.unsqueeze(-1).expand(-1, -1, T), cond], dim=1)
    bias = (1.0 - mask) * -1.0e10                           # (B,1,T): key-padding bias (static_chunk_size = 0)

    def tfm(prefix, h):
        h = h.transpose(1, 2).contiguous()
        for j in range(n_tfm):
            h = transformer_block(sd, f"{prefix}{j}.", h, bias)
        return h.transpose(1, 2).contiguous()

    h = resnet(sd, pfx + "down_blocks.0.0.", h, mask, temb)
    h = tfm(pfx + "down_blocks.0.1.", h)
    skip = h
    h = causal_conv(h * mask, sd[pfx + "down_blocks.0.2.weight"], sd[pfx + "down_blocks.0.2.bias"])
    for m in range(n_mid):
        h = resnet(sd, pfx + f"mid_blocks.{m}.0.", h, mask, temb)
        h = tfm(pfx + f"mid_blocks.{m}.1.", h)
    h = torch.cat([h, skip], dim=1)
    h = resnet(sd, pfx + "up_blocks.0.0.", h, mask, temb)
    h = tfm(pfx + "up_blocks.0.1.", h)
    h = causal_conv(h * mask, sd[pfx + "up_blocks.0.2.weight"], sd[pfx + "up_blocks.0.2.bias"])
    h = causal_block(sd, pfx + "final_block.", h, mask)
    out = F.conv1d(h * mask, sd[pfx + "final_proj.weight"], sd[pfx + "final_proj.bias"])
    return out * mask


def t_schedule(n_timesteps: int = 10) -> torch.Tensor:
    """flow_matching.py:215-217: cosine schedule."""
    t = torch.linspace(0, 1, n_timesteps + 1)
    return 1 - torch.cos(t * 0.5 * torch.pi)


def solve_euler(sd, z, mu, spks, cond, n_timesteps: int = 10, cfg_rate: float = 0.7, pfx: str = "",
                n_mid: int = 12, n_tfm: int = 4, trace: Optional[List[torch.Tensor]] = None) -> torch.Tensor:
    """z, mu, cond (1,80,T); spks (1,80).  Returns x after n_timesteps Euler steps (1,80,T)."""
    t_span = t_schedule(n_timesteps).to(z.device)
    T = z.shape[-1]
    mask = torch.ones(1, 1, T, device=z.device)
    x = z
    t, dt = t_span[0].unsqueeze(0), t_span[1] - t_span[0]
    for step in range(1, len(t_span)):
        x_in = torch.cat([x, x], 0)
        mu_in = torch.cat([mu, torch.zeros_like(mu)], 0)
        spks_in = torch.cat([spks, torch.zeros_like(spks)], 0)
        cond_in = torch.cat([cond, torch.zeros_like(cond)], 0)
        d = estimator(sd, x_in, torch.cat([mask, mask], 0), mu_in, t.expand(2), spks_in, cond_in, pfx, n_mid, n_tfm)
        v = (1.0 + cfg_rate) * d[0:1] - cfg_rate * d[1:2]
        if trace is not None:
            trace.append(v)
        x = x + dt * v
        t = t + dt
        if step < len(t_span) - 1:
            dt = t_span[step + 1] - t
    return x.float()
