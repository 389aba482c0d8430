"""Import harness for the UNMODIFIED reference modules (container-only test infrastructure).

TEST INFRASTRUCTURE — never imported by the product path (`chatterbox_b200/`).

`/root/reference` exists only in the build container, not on the GPU box.  This module makes
the reference's own hot-path modules importable there so that (a) the CPU restatements in
`oracle/*.py` can be validated against them and (b) `oracle/make_golden.py` can emit the pinned
vectors under `tests/golden/`.

Two accommodations are needed:

1. `chatterbox/__init__.py:6-10` pulls in `tts.py`/`vc.py` (librosa, perth: absent), so the
   parent packages are pre-registered as empty namespace modules whose `__path__` points into
   `/root/reference/src/chatterbox/...`; sub-modules are then imported directly and unmodified.
2. `matcha/decoder.py:7-8` and `matcha/transformer.py:5-14` import `conformer` and `diffusers`
   (absent).  Minimal import shims restate exactly the symbols the estimator touches
   (diffusers==0.29.0 `Attention` with `AttnProcessor2_0`, `GELU`, `LoRACompatibleLinear`,
   `get_activation`); everything else is an empty placeholder that is never instantiated.
"""
from __future__ import annotations

import os
import sys
import types

import torch
import torch.nn as nn
import torch.nn.functional as F

REF_ROOT = os.environ.get("CBX_REFERENCE_ROOT", "/root/reference")
REF_SRC = os.path.join(REF_ROOT, "src")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_SRC, "chatterbox", "models"))


def _ns(name: str, path: str) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__path__ = [path]
    m.__package__ = name
    sys.modules[name] = m
    return m


# --------------------------------------------------------------------------------------
# diffusers==0.29.0 shim (third-party arithmetic restated; state-dict key names preserved)
# --------------------------------------------------------------------------------------
class _ShimAttention(nn.Module):
    """diffusers.models.attention_processor.Attention (v0.29.0) with the default
    AttnProcessor2_0, self-attention only: to_q/k/v Linear(dim, heads*dim_head, bias=False),
    to_out = [Linear(inner, dim, bias=True), Dropout], SDPA with additive mask, scale 1/sqrt(d)."""

    def __init__(self, query_dim, heads=8, dim_head=64, dropout=0.0, bias=False,
                 cross_attention_dim=None, upcast_attention=False, **kw):
        super().__init__()
        assert cross_attention_dim is None
        inner = heads * dim_head
        self.heads = heads
        self.to_q = nn.Linear(query_dim, inner, bias=bias)
        self.to_k = nn.Linear(query_dim, inner, bias=bias)
        self.to_v = nn.Linear(query_dim, inner, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim, bias=True), nn.Dropout(dropout)])

    def forward(self, hidden_states, encoder_hidden_states=None, attention_mask=None, **kw):
        assert encoder_hidden_states is None
        B, L, _ = hidden_states.shape
        H = self.heads
        if attention_mask is not None:
            # prepare_attention_mask: (B, 1, L) -> repeat_interleave(heads) -> (B, H, 1|L, L)
            attention_mask = attention_mask.repeat_interleave(H, dim=0)
            attention_mask = attention_mask.view(B, H, -1, attention_mask.shape[-1])
        q = self.to_q(hidden_states).view(B, L, H, -1).transpose(1, 2)
        k = self.to_k(hidden_states).view(B, L, H, -1).transpose(1, 2)
        v = self.to_v(hidden_states).view(B, L, H, -1).transpose(1, 2)
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=attention_mask, dropout_p=0.0, is_causal=False)
        o = o.transpose(1, 2).reshape(B, L, -1)
        o = self.to_out[0](o)
        return self.to_out[1](o)


class _ShimGELU(nn.Module):
    """diffusers.models.activations.GELU: gelu(Linear(dim_in, dim_out)), exact erf by default."""

    def __init__(self, dim_in, dim_out, approximate="none", bias=True):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out, bias=bias)
        self.approximate = approximate

    def forward(self, x):
        return F.gelu(self.proj(x), approximate=self.approximate)


class _Unused(nn.Module):
    def __init__(self, *a, **k):
        raise NotImplementedError("placeholder for a diffusers/conformer symbol the hot path never builds")


def _install_shims() -> None:
    if "diffusers" not in sys.modules:
        d = types.ModuleType("diffusers")
        d.__path__ = []
        dm = types.ModuleType("diffusers.models")
        dm.__path__ = []
        att = types.ModuleType("diffusers.models.attention")
        att.GEGLU = _Unused
        att.GELU = _ShimGELU
        att.AdaLayerNorm = _Unused
        att.AdaLayerNormZero = _Unused
        att.ApproximateGELU = _Unused
        ap = types.ModuleType("diffusers.models.attention_processor")
        ap.Attention = _ShimAttention
        lora = types.ModuleType("diffusers.models.lora")
        lora.LoRACompatibleLinear = nn.Linear
        act = types.ModuleType("diffusers.models.activations")

        def get_activation(name):
            return {"silu": nn.SiLU(), "swish": nn.SiLU(), "mish": nn.Mish(), "gelu": nn.GELU()}[name]

        act.get_activation = get_activation
        du = types.ModuleType("diffusers.utils")
        du.__path__ = []
        tu = types.ModuleType("diffusers.utils.torch_utils")
        tu.maybe_allow_in_graph = lambda cls: cls
        for name, mod in [("diffusers", d), ("diffusers.models", dm), ("diffusers.models.attention", att),
                          ("diffusers.models.attention_processor", ap), ("diffusers.models.lora", lora),
                          ("diffusers.models.activations", act), ("diffusers.utils", du),
                          ("diffusers.utils.torch_utils", tu)]:
            sys.modules[name] = mod
    if "conformer" not in sys.modules:
        c = types.ModuleType("conformer")
        c.ConformerBlock = _Unused
        sys.modules["conformer"] = c


_installed = False


def install() -> None:
    """Make `chatterbox.models.*` importable from /root/reference without running its __init__."""
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError(f"reference tree not found under {REF_ROOT} (exists only in the build container)")
    _install_shims()
    base = os.path.join(REF_SRC, "chatterbox")
    _ns("chatterbox", base)
    _ns("chatterbox.models", os.path.join(base, "models"))
    _ns("chatterbox.models.s3gen", os.path.join(base, "models", "s3gen"))
    _ns("chatterbox.models.t3", os.path.join(base, "models", "t3"))
    _installed = True


# --------------------------------------------------------------------------------------
# Constructors with the hyper-parameters the reference uses (s3gen.py:59-98, 222-230)
# --------------------------------------------------------------------------------------
def build_t3():
    install()
    from chatterbox.models.t3.t3 import T3
    return T3().eval()


def build_flow():
    """CausalMaskedDiffWithXvec(encoder, CausalConditionalCFM(ConditionalDecoder)) as s3gen.py:59-98."""
    install()
    from chatterbox.models.s3gen.flow import CausalMaskedDiffWithXvec
    from chatterbox.models.s3gen.transformer.upsample_encoder import UpsampleConformerEncoder
    from chatterbox.models.s3gen.flow_matching import CausalConditionalCFM
    from chatterbox.models.s3gen.decoder import ConditionalDecoder
    from chatterbox.models.s3gen.configs import CFM_PARAMS
    encoder = UpsampleConformerEncoder(
        output_size=512, attention_heads=8, linear_units=2048, num_blocks=6, dropout_rate=0.1,
        positional_dropout_rate=0.1, attention_dropout_rate=0.1, normalize_before=True,
        input_layer='linear', pos_enc_layer_type='rel_pos_espnet', selfattention_layer_type='rel_selfattn',
        input_size=512, use_cnn_module=False, macaron_style=False)
    estimator = ConditionalDecoder(
        in_channels=320, out_channels=80, causal=True, channels=[256], dropout=0.0,
        attention_head_dim=64, n_blocks=4, num_mid_blocks=12, num_heads=8, act_fn='gelu')
    decoder = CausalConditionalCFM(spk_emb_dim=80, cfm_params=CFM_PARAMS, estimator=estimator)
    return CausalMaskedDiffWithXvec(encoder=encoder, decoder=decoder).eval()


def build_hift():
    """HiFTGenerator as s3gen.py:222-230."""
    install()
    from chatterbox.models.s3gen.f0_predictor import ConvRNNF0Predictor
    from chatterbox.models.s3gen.hifigan import HiFTGenerator
    return HiFTGenerator(
        sampling_rate=24000, upsample_rates=[8, 5, 3], upsample_kernel_sizes=[16, 11, 7],
        source_resblock_kernel_sizes=[7, 7, 11],
        source_resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]],
        f0_predictor=ConvRNNF0Predictor()).eval()


def trim_fade() -> torch.Tensor:
    """s3gen.py:233-236."""
    n_trim = 24000 // 50
    tf = torch.zeros(2 * n_trim)
    tf[n_trim:] = (torch.cos(torch.linspace(torch.pi, 0, n_trim)) + 1) / 2
    return tf
