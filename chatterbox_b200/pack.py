"""One-time weight pre-pack for the T3 decode megakernel (fp32 state dict -> bf16 per-CTA streams).

The kernel's TMA producer walks ONE contiguous byte stream per CTA, in the order the static plan
(`csrc/t3_plan.h`, exported as `cbx_t3_plan_count/unit`) says that CTA consumes 32 KB units.  A
unit is 16 weight rows x 1024 k in bf16, stored in mma.m16n8k16 A-fragment order
[warp k-slice w][k16 step s][lane][a0 a1 a2 a3], so every consumer lane fetches the four A registers of
one MMA with a single conflict-free 128-bit shared-memory read.

State-dict keys are the reference's (`T3.state_dict()`, src/chatterbox/models/t3/t3.py:41-65).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List

import torch

from . import _lib

H, INTER, VOCAB, VOCAB_PAD, UNIT_ROWS, UNIT_ELEMS = 1024, 4096, 8194, 8208, 16, 16384
PH_QKV, PH_O, PH_GU, PH_DOWN, PH_HEAD = range(5)
PHASE_UNITS = (192, 64, 512, 256, 513)
LAYER_UNITS = sum(PHASE_UNITS[:4])


def plan_lists(G: int) -> List[List[List[int]]]:
    """plan[c][ph] = unit ids CTA c consumes in phase ph, in order (from the C plan)."""
    l = _lib.lib()
    return [[[l.cbx_t3_plan_unit(ph, c, G, j) for j in range(l.cbx_t3_plan_count(ph, c, G))]
             for ph in range(5)] for c in range(G)]


def kwarps() -> int:
    """Warps of a megakernel CTA = k-slices of a unit (csrc/t3_plan.h: T3_KW), asked from the library."""
    return _lib.lib().cbx_t3_kwarps()


def _frag_order(units: torch.Tensor) -> torch.Tensor:
    """[U, 16 rows, 1024 k] -> [U, 16384] in A-fragment order: row = rh*8 + g, k = w*(16 KS) + s*16 + kh*8 + tig*2 + e
    -> [w][s][g][tig][kh][rh][e]  (a0 = (kh 0, rh 0), a1 = (0, 1), a2 = (1, 0), a3 = (1, 1)); KS = 64 / KW k16 steps."""
    U = units.shape[0]
    KW = kwarps()
    KS = 1024 // KW // 16
    return units.reshape(U, 2, 8, KW, KS, 2, 4, 2).permute(0, 3, 4, 2, 6, 5, 1, 7).reshape(U, UNIT_ELEMS)


def layer_units(sd: Dict[str, torch.Tensor], i: int, dev) -> torch.Tensor:
    """All 1024 units of decoder layer i in phase order (qkv | o | gate-up | down): [1024, 16384] bf16."""
    p = f"tfmr.layers.{i}."
    g = lambda n: sd[p + n].to(dev, torch.float32)
    qkv = torch.cat([g("self_attn.q_proj.weight"), g("self_attn.k_proj.weight"), g("self_attn.v_proj.weight")], 0)
    a = qkv.reshape(192, 16, H)
    c = g("self_attn.o_proj.weight").reshape(64, 16, H)
    gate = g("mlp.gate_proj.weight").reshape(512, 8, H)
    up = g("mlp.up_proj.weight").reshape(512, 8, H)
    d = torch.cat([gate, up], 1)                                   # unit u: gate rows 8u..8u+7, up rows 8u..8u+7
    wd = g("mlp.down_proj.weight").reshape(64, 16, 4, 1024)        # [r][row][s][k]
    e = wd.permute(0, 2, 1, 3).reshape(256, 16, 1024)              # unit u = r*4 + s
    return _frag_order(torch.cat([a, c, d, e], 0)).to(torch.bfloat16)


def head_units(sd: Dict[str, torch.Tensor], dev) -> torch.Tensor:
    w = sd["speech_head.weight"].to(dev, torch.float32)
    w = torch.cat([w, torch.zeros(VOCAB_PAD - VOCAB, H, device=dev)], 0)
    return _frag_order(w.reshape(513, 16, H)).to(torch.bfloat16)


def plain_weights(sd: Dict[str, torch.Tensor], dev, n_layers: int) -> torch.Tensor:
    """Row-major bf16 copies for the GEMM prefill (csrc/t3_prefill.cu): per layer q|k|v, o, gate/up interleaved in
    128-row groups (tile t = gate rows [128t, 128t+128) then up rows [128t, 128t+128): the SwiGLU epilogue pairs the two
    halves of a 256-column tile), down; then speech_head.  One flat tensor."""
    parts = []
    for i in range(n_layers):
        p = f"tfmr.layers.{i}."
        g = lambda n: sd[p + n].to(dev, torch.float32)
        parts.append(torch.cat([g("self_attn.q_proj.weight"), g("self_attn.k_proj.weight"), g("self_attn.v_proj.weight")], 0))
        parts.append(g("self_attn.o_proj.weight"))
        gate = g("mlp.gate_proj.weight").reshape(32, 128, H)
        up = g("mlp.up_proj.weight").reshape(32, 128, H)
        parts.append(torch.cat([gate, up], 1).reshape(8192, H))
        parts.append(g("mlp.down_proj.weight"))
    parts.append(sd["speech_head.weight"].to(dev, torch.float32))
    return torch.cat([t.reshape(-1).to(torch.bfloat16) for t in parts])


def rope_tables(max_pos: int, dev) -> (torch.Tensor, torch.Tensor):
    """cos/sin [max_pos][32] fp32 of the llama3-scaled RoPE the reference configures
    (llama_configs.py:22-30: theta 5e5, factor 8, low/high 1/4, original ctx 8192), computed in
    fp32 exactly as HF `_compute_llama3_parameters` + `LlamaRotaryEmbedding.forward` do."""
    theta, factor, low_f, high_f, old_ctx, hd = 500000.0, 8.0, 1.0, 4.0, 8192, 64
    inv = 1.0 / (theta ** (torch.arange(0, hd, 2, dtype=torch.int64).to(torch.float) / hd))
    low_wl, high_wl = old_ctx / low_f, old_ctx / high_f
    wl = 2 * math.pi / inv
    inv_l = torch.where(wl > low_wl, inv / factor, inv)
    smooth = (old_ctx / wl - low_f) / (high_f - low_f)
    smoothed = (1 - smooth) * inv_l / factor + smooth * inv_l
    medium = ~(wl < high_wl) * ~(wl > low_wl)
    inv = torch.where(medium, smoothed, inv_l)
    fr = torch.arange(max_pos, dtype=torch.float32)[:, None] * inv[None, :]
    return fr.cos().contiguous().to(dev), fr.sin().contiguous().to(dev)


@dataclass
class T3Packed:
    G: int
    n_layers: int
    stride_units: int
    wstream: torch.Tensor       # [G, stride_units, 16384] bf16
    ln1: torch.Tensor
    ln2: torch.Tensor
    final_norm: torch.Tensor
    speech_emb: torch.Tensor
    speech_pos_emb: torch.Tensor
    rope_cos: torch.Tensor
    rope_sin: torch.Tensor
    max_pos: int
    wplain: torch.Tensor = None   # plain matrices for the GEMM prefill (None: prefill through the decode kernel)

    def weight_bytes_per_step(self) -> int:
        """Algorithmic weight bytes one decode step streams (every unit once + norms)."""
        return (self.n_layers * LAYER_UNITS + PHASE_UNITS[4]) * UNIT_ELEMS * 2 + (2 * self.n_layers + 1) * H * 4


def pack_t3(sd: Dict[str, torch.Tensor], G: int, device, max_pos: int, n_layers: int = 30,
            gemm_prefill: bool = True) -> T3Packed:
    dev = torch.device(device)
    plan = plan_lists(G)
    offs = [0, 192, 192 + 64, 192 + 64 + 512]
    layer_idx = [torch.tensor([offs[ph] + u for ph in range(4) for u in plan[c][ph]], dtype=torch.long, device=dev)
                 for c in range(G)]
    head_idx = [torch.tensor(plan[c][PH_HEAD], dtype=torch.long, device=dev) for c in range(G)]
    lens = [n_layers * len(layer_idx[c]) + len(head_idx[c]) for c in range(G)]
    stride = max(lens)
    stream = torch.zeros(G, stride, UNIT_ELEMS, dtype=torch.bfloat16, device=dev)
    hu = head_units(sd, dev)
    for i in range(n_layers):
        lu = layer_units(sd, i, dev)
        for c in range(G):
            n = len(layer_idx[c])
            stream[c, i * n:(i + 1) * n] = lu[layer_idx[c]]
        del lu
    for c in range(G):
        n = n_layers * len(layer_idx[c])
        stream[c, n:n + len(head_idx[c])] = hu[head_idx[c]]
    f32 = lambda t: t.to(dev, torch.float32).contiguous()
    ln1 = torch.stack([f32(sd[f"tfmr.layers.{i}.input_layernorm.weight"]) for i in range(n_layers)])
    ln2 = torch.stack([f32(sd[f"tfmr.layers.{i}.post_attention_layernorm.weight"]) for i in range(n_layers)])
    cos, sin = rope_tables(max_pos, dev)
    return T3Packed(G=G, n_layers=n_layers, stride_units=stride, wstream=stream, ln1=ln1.contiguous(),
                    ln2=ln2.contiguous(), final_norm=f32(sd["tfmr.norm.weight"]),
                    speech_emb=f32(sd["speech_emb.weight"]), speech_pos_emb=f32(sd["speech_pos_emb.emb.weight"]),
                    rope_cos=cos, rope_sin=sin, max_pos=max_pos,
                    wplain=plain_weights(sd, dev, n_layers) if gemm_prefill else None)
