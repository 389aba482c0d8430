"""Deterministic synthetic weights and inputs of the named architecture (no checkpoints offline).

Everything here is derived from a key name and a seed only, so the build container (where the
reference can be imported and golden vectors are made) and the GPU box (where it cannot)
materialise bit-identical tensors: `torch.randn` on a CPU `torch.Generator` is platform-stable
for a given torch version.

State-dict key names and shapes are the reference's own (so a real checkpoint loads through the
same code):
  * T3      — `src/chatterbox/models/t3/t3.py:41-65`, `modules/cond_enc.py:40-62`,
              `modules/perceiver.py:173-198`, HF `LlamaModel` configured by `llama_configs.py:1-33`
  * flow    — `s3gen/s3gen.py:59-98` (`CausalMaskedDiffWithXvec`, encoder, `ConditionalDecoder`)
  * HiFT    — `s3gen/s3gen.py:222-230`, `s3gen/hifigan.py:291-380`, `s3gen/f0_predictor.py:19-50`
`tests/test_synthetic.py` checks the generated key/shape set against the manifest of the
reference's `state_dict()` committed under `tests/golden/ref_manifest.json`.
"""
from __future__ import annotations

import math
import zlib
from collections import OrderedDict
from typing import Dict, List, Tuple

import torch

Spec = "OrderedDict[str, Tuple[int, ...]]"

# ---------------------------------------------------------------------------------------------
# architecture constants (reference file:line in the module docstring)
# ---------------------------------------------------------------------------------------------
T3_HIDDEN = 1024
T3_LAYERS = 30
T3_HEADS = 16
T3_HEAD_DIM = 64
T3_INTER = 4096
T3_TEXT_VOCAB = 704
T3_SPEECH_VOCAB = 8194
T3_MAX_TEXT_POS = 2048 + 2
T3_MAX_SPEECH_POS = 4096 + 2 + 2
T3_START_TEXT, T3_STOP_TEXT = 255, 0
T3_START_SPEECH, T3_STOP_SPEECH = 6561, 6562
S3_VOCAB = 6561


def t3_spec(n_layers: int = T3_LAYERS) -> Spec:
    s: Spec = OrderedDict()
    H, I = T3_HIDDEN, T3_INTER
    s["tfmr.embed_tokens.weight"] = (8, H)
    for i in range(n_layers):
        p = f"tfmr.layers.{i}."
        for n in ("q_proj", "k_proj", "v_proj", "o_proj"):
            s[p + f"self_attn.{n}.weight"] = (H, H)
        s[p + "mlp.gate_proj.weight"] = (I, H)
        s[p + "mlp.up_proj.weight"] = (I, H)
        s[p + "mlp.down_proj.weight"] = (H, I)
        s[p + "input_layernorm.weight"] = (H,)
        s[p + "post_attention_layernorm.weight"] = (H,)
    s["tfmr.norm.weight"] = (H,)
    s["cond_enc.spkr_enc.weight"] = (H, 256)
    s["cond_enc.spkr_enc.bias"] = (H,)
    s["cond_enc.emotion_adv_fc.weight"] = (H, 1)
    s["cond_enc.perceiver.pre_attention_query"] = (1, 32, H)
    s["cond_enc.perceiver.attn.norm.weight"] = (H,)
    s["cond_enc.perceiver.attn.norm.bias"] = (H,)
    for n in ("to_q", "to_k", "to_v", "proj_out"):
        s[f"cond_enc.perceiver.attn.{n}.weight"] = (H, H)
        s[f"cond_enc.perceiver.attn.{n}.bias"] = (H,)
    s["text_emb.weight"] = (T3_TEXT_VOCAB, H)
    s["speech_emb.weight"] = (T3_SPEECH_VOCAB, H)
    s["text_pos_emb.emb.weight"] = (T3_MAX_TEXT_POS, H)
    s["speech_pos_emb.emb.weight"] = (T3_MAX_SPEECH_POS, H)
    s["text_head.weight"] = (T3_TEXT_VOCAB, H)
    s["speech_head.weight"] = (T3_SPEECH_VOCAB, H)
    return s


def _conformer_layer(s: Spec, p: str) -> None:
    s[p + "self_attn.pos_bias_u"] = (8, 64)
    s[p + "self_attn.pos_bias_v"] = (8, 64)
    for n in ("linear_q", "linear_k", "linear_v", "linear_out"):
        s[p + f"self_attn.{n}.weight"] = (512, 512)
        s[p + f"self_attn.{n}.bias"] = (512,)
    s[p + "self_attn.linear_pos.weight"] = (512, 512)
    s[p + "feed_forward.w_1.weight"] = (2048, 512)
    s[p + "feed_forward.w_1.bias"] = (2048,)
    s[p + "feed_forward.w_2.weight"] = (512, 2048)
    s[p + "feed_forward.w_2.bias"] = (512,)
    for n in ("norm_ff", "norm_mha"):
        s[p + n + ".weight"] = (512,)
        s[p + n + ".bias"] = (512,)


def _resnet(s: Spec, p: str, cin: int, cout: int) -> None:
    s[p + "mlp.1.weight"] = (cout, 1024)
    s[p + "mlp.1.bias"] = (cout,)
    for blk, ci in (("block1", cin), ("block2", cout)):
        s[p + f"{blk}.block.0.weight"] = (cout, ci, 3)
        s[p + f"{blk}.block.0.bias"] = (cout,)
        s[p + f"{blk}.block.2.weight"] = (cout,)
        s[p + f"{blk}.block.2.bias"] = (cout,)
    s[p + "res_conv.weight"] = (cout, cin, 1)
    s[p + "res_conv.bias"] = (cout,)


def _tfm_block(s: Spec, p: str) -> None:
    s[p + "norm1.weight"] = (256,)
    s[p + "norm1.bias"] = (256,)
    for n in ("to_q", "to_k", "to_v"):
        s[p + f"attn1.{n}.weight"] = (512, 256)
    s[p + "attn1.to_out.0.weight"] = (256, 512)
    s[p + "attn1.to_out.0.bias"] = (256,)
    s[p + "norm3.weight"] = (256,)
    s[p + "norm3.bias"] = (256,)
    s[p + "ff.net.0.proj.weight"] = (1024, 256)
    s[p + "ff.net.0.proj.bias"] = (1024,)
    s[p + "ff.net.2.weight"] = (256, 1024)
    s[p + "ff.net.2.bias"] = (256,)


def estimator_spec(prefix: str = "") -> Spec:
    """ConditionalDecoder(in=320, out=80, channels=[256], n_blocks=4, num_mid_blocks=12)."""
    s: Spec = OrderedDict()
    e = prefix
    s[e + "time_mlp.linear_1.weight"] = (1024, 320)
    s[e + "time_mlp.linear_1.bias"] = (1024,)
    s[e + "time_mlp.linear_2.weight"] = (1024, 1024)
    s[e + "time_mlp.linear_2.bias"] = (1024,)
    _resnet(s, e + "down_blocks.0.0.", 320, 256)
    for j in range(4):
        _tfm_block(s, e + f"down_blocks.0.1.{j}.")
    s[e + "down_blocks.0.2.weight"] = (256, 256, 3)
    s[e + "down_blocks.0.2.bias"] = (256,)
    for m in range(12):
        _resnet(s, e + f"mid_blocks.{m}.0.", 256, 256)
        for j in range(4):
            _tfm_block(s, e + f"mid_blocks.{m}.1.{j}.")
    _resnet(s, e + "up_blocks.0.0.", 512, 256)
    for j in range(4):
        _tfm_block(s, e + f"up_blocks.0.1.{j}.")
    s[e + "up_blocks.0.2.weight"] = (256, 256, 3)
    s[e + "up_blocks.0.2.bias"] = (256,)
    s[e + "final_block.block.0.weight"] = (256, 256, 3)
    s[e + "final_block.block.0.bias"] = (256,)
    s[e + "final_block.block.2.weight"] = (256,)
    s[e + "final_block.block.2.bias"] = (256,)
    s[e + "final_proj.weight"] = (80, 256, 1)
    s[e + "final_proj.bias"] = (80,)
    return s


def flow_spec() -> Spec:
    s: Spec = OrderedDict()
    s["input_embedding.weight"] = (S3_VOCAB, 512)
    s["spk_embed_affine_layer.weight"] = (80, 192)
    s["spk_embed_affine_layer.bias"] = (80,)
    for emb in ("encoder.embed.",):
        s[emb + "out.0.weight"] = (512, 512)
        s[emb + "out.0.bias"] = (512,)
        s[emb + "out.1.weight"] = (512,)
        s[emb + "out.1.bias"] = (512,)
    s["encoder.after_norm.weight"] = (512,)
    s["encoder.after_norm.bias"] = (512,)
    s["encoder.pre_lookahead_layer.conv1.weight"] = (512, 512, 4)
    s["encoder.pre_lookahead_layer.conv1.bias"] = (512,)
    s["encoder.pre_lookahead_layer.conv2.weight"] = (512, 512, 3)
    s["encoder.pre_lookahead_layer.conv2.bias"] = (512,)
    for i in range(6):
        _conformer_layer(s, f"encoder.encoders.{i}.")
    s["encoder.up_layer.conv.weight"] = (512, 512, 5)
    s["encoder.up_layer.conv.bias"] = (512,)
    s["encoder.up_embed.out.0.weight"] = (512, 512)
    s["encoder.up_embed.out.0.bias"] = (512,)
    s["encoder.up_embed.out.1.weight"] = (512,)
    s["encoder.up_embed.out.1.bias"] = (512,)
    for i in range(4):
        _conformer_layer(s, f"encoder.up_encoders.{i}.")
    s["encoder_proj.weight"] = (80, 512)
    s["encoder_proj.bias"] = (80,)
    s.update(estimator_spec("decoder.estimator."))
    return s


def _wn(s: Spec, p: str, cout: int, cin: int, k: int, transposed: bool = False) -> None:
    """weight_norm parametrisation: original0 = g, original1 = v (hifigan.py:28,120-142)."""
    s[p + "bias"] = (cin if transposed else cout,)
    s[p + "parametrizations.weight.original0"] = (cout, 1, 1)
    s[p + "parametrizations.weight.original1"] = (cout, cin, k)


def _resblock(s: Spec, p: str, ch: int, k: int) -> None:
    for c in ("convs1", "convs2"):
        for j in range(3):
            _wn(s, p + f"{c}.{j}.", ch, ch, k)
    for a in ("activations1", "activations2"):
        for j in range(3):
            s[p + f"{a}.{j}.alpha"] = (ch,)


def hift_spec() -> Spec:
    s: Spec = OrderedDict()
    s["m_source.l_linear.weight"] = (1, 9)
    s["m_source.l_linear.bias"] = (1,)
    _wn(s, "conv_pre.", 512, 80, 7)
    # ConvTranspose1d weight is (in, out, k); weight_norm dim=0 -> g per *input* channel
    _wn(s, "ups.0.", 512, 256, 16, transposed=True)
    _wn(s, "ups.1.", 256, 128, 11, transposed=True)
    _wn(s, "ups.2.", 128, 64, 7, transposed=True)
    for i, (c, k) in enumerate(((256, 30), (128, 6), (64, 1))):
        s[f"source_downs.{i}.weight"] = (c, 18, k)
        s[f"source_downs.{i}.bias"] = (c,)
    for i, (c, k) in enumerate(((256, 7), (128, 7), (64, 11))):
        _resblock(s, f"source_resblocks.{i}.", c, k)
    for i, c in enumerate((256, 128, 64)):
        for j, k in enumerate((3, 7, 11)):
            _resblock(s, f"resblocks.{i * 3 + j}.", c, k)
    _wn(s, "conv_post.", 18, 64, 7)
    _wn(s, "f0_predictor.condnet.0.", 512, 80, 3)
    for j in (2, 4, 6, 8):
        _wn(s, f"f0_predictor.condnet.{j}.", 512, 512, 3)
    s["f0_predictor.classifier.weight"] = (1, 512)
    s["f0_predictor.classifier.bias"] = (1,)
    return s


# ---------------------------------------------------------------------------------------------
# deterministic initialisation (a function of key name + seed only)
# ---------------------------------------------------------------------------------------------
def _gen(key: str, seed: int) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed((zlib.crc32(key.encode()) + 1000003 * int(seed)) & 0x7FFFFFFF)
    return g


def _randn(key: str, seed: int, shape) -> torch.Tensor:
    return torch.randn(tuple(shape), generator=_gen(key, seed), dtype=torch.float32)


def _fan_in(shape) -> int:
    n = 1
    for d in shape[1:]:
        n *= d
    return max(n, 1)


def init_tensor(family: str, key: str, shape, seed: int, bf16_exact: bool = True) -> torch.Tensor:
    """Initialisation rule for one tensor.  Scales follow the reference constructors
    (HF `initializer_range=0.02`, `nn.Embedding` N(0,1), `LearnedPositionEmbeddings` 0.02,
    weight-norm g≈|v|) but norm gains, biases and Snake alphas are made non-trivial so that a
    kernel that drops one of them cannot pass parity."""
    r = _randn(key, seed, shape)
    last = key.rsplit(".", 1)[-1]
    if family == "t3":
        if key in ("text_emb.weight", "speech_emb.weight"):
            return r
        if key.endswith("pre_attention_query"):
            a = math.sqrt(3.0) * math.sqrt(2.0 / 64.0)
            return (torch.rand(tuple(shape), generator=_gen(key + "#u", seed)) * 2 - 1) * a
        if key.endswith("emotion_adv_fc.weight"):
            return r * 0.5
        if "norm" in key and last == "weight":
            return 1.0 + 0.1 * r
        if last == "bias":
            return 0.02 * r
        w = 0.02 * r
        if key == "speech_head.weight":
            # a trained T3 does not emit the ids S3Gen cannot consume (>= 6561: BOS/EOS/unused, tts.py:257-262);
            # zero rows give those ids logit 0, below the winning logit of any step, so synthetic utterances keep
            # every generated token and the S3Gen workload is exactly 2 x 256 mel frames
            w[S3_VOCAB:] = 0.0
        if bf16_exact and (key == "speech_head.weight" or (key.startswith("tfmr.layers.") and key.endswith("_proj.weight"))):
            # the streamed matrices are made bf16-representable, so the reference (fp32 math) and the
            # build (bf16 storage, fp32 accumulate) run on IDENTICAL weights.  `bf16_exact=False` keeps the
            # fp32 values (what a real t3_cfg.safetensors holds): the build then pays its bf16 storage
            # rounding against the fp32 reference (tests/test_t3_fp32w_gpu.py)
            w = w.to(torch.bfloat16).to(torch.float32)
        return w
    if family == "flow":
        if key == "input_embedding.weight":
            return r
        if key.endswith("pos_bias_u") or key.endswith("pos_bias_v"):
            return 0.15 * r
        is_norm = (len(shape) == 1 and last == "weight")  # LayerNorm gains are the only 1-D weights
        if is_norm:
            return 1.0 + 0.1 * r
        if last == "bias":
            return 0.02 * r
        return (0.7 / math.sqrt(_fan_in(shape))) * r
    if family == "hift":
        if last == "alpha":
            return (1.0 + 0.2 * r).clamp_min(0.2)
        if last == "original0":
            return (0.5 + 0.05 * r).abs()
        if last == "original1":
            return 0.02 * r
        if last == "bias":
            return 0.02 * r
        if key.startswith("m_source.l_linear"):
            return 0.3 * r
        if key == "f0_predictor.classifier.bias":
            return torch.full(tuple(shape), 140.0)       # random-init f0 lands in the voiced range (> 10 Hz)
        if key == "f0_predictor.classifier.weight":
            return 40.0 * r
        return (1.0 / math.sqrt(_fan_in(shape))) * r
    raise ValueError(family)


def make_state_dict(family: str, seed: int = 1234, n_layers: int | None = None,
                    bf16_exact: bool = True) -> "OrderedDict[str, torch.Tensor]":
    """family in {"t3", "flow", "hift"}; `n_layers` truncates the T3 stack (tests only); `bf16_exact=False`
    leaves the streamed T3 matrices un-rounded fp32 (see init_tensor)."""
    if family == "t3":
        spec = t3_spec(T3_LAYERS if n_layers is None else n_layers)
    elif family == "flow":
        spec = flow_spec()
    elif family == "hift":
        spec = hift_spec()
    else:
        raise ValueError(family)
    return OrderedDict((k, init_tensor(family, k, shp, seed, bf16_exact)) for k, shp in spec.items())


def cfm_noise(seed: int = 1234, length: int = 50 * 300) -> torch.Tensor:
    """Stand-in for `CausalConditionalCFM.rand_noise` (flow_matching.py:191): fixed (1,80,15000)."""
    return _randn("decoder.rand_noise", seed, (1, 80, length))


# ---------------------------------------------------------------------------------------------
# synthetic inputs
# ---------------------------------------------------------------------------------------------
def text_tokens(n: int = 64, seed: int = 1) -> torch.Tensor:
    """(1, n+2) int64: SOT 255 + n ids uniform in [1,703] excluding 255 + EOT 0 (tts.py:234-243)."""
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(1, T3_TEXT_VOCAB - 1, (n,), generator=g)
    ids = torch.where(ids >= T3_START_TEXT, ids + 1, ids)  # skip 255
    return torch.cat([torch.tensor([T3_START_TEXT]), ids, torch.tensor([T3_STOP_TEXT])])[None]


def t3_cond(seed: int = 2, prompt_len: int = 150, exaggeration: float = 0.5) -> Dict[str, torch.Tensor]:
    """Fields of `T3Cond` (cond_enc.py:11-24) for a 6 s synthetic reference voice."""
    g = torch.Generator().manual_seed(seed)
    tokens = torch.randint(0, S3_VOCAB, (1, prompt_len), generator=g)
    g3 = torch.Generator().manual_seed(seed + 1)
    spk = torch.randn(1, 256, generator=g3).abs()
    spk = spk / spk.norm(dim=1, keepdim=True)
    return dict(speaker_emb=spk, cond_prompt_speech_tokens=tokens,
                emotion_adv=exaggeration * torch.ones(1, 1, 1))


def s3gen_ref(seed: int = 4, prompt_len: int = 150) -> Dict[str, torch.Tensor]:
    """`ref_dict` of `S3Token2Mel.embed_ref` (s3gen.py:151-157) for a synthetic voice."""
    g = torch.Generator().manual_seed(seed)
    prompt_token = torch.randint(0, S3_VOCAB, (1, prompt_len), generator=g)
    prompt_feat = (torch.randn(1, 2 * prompt_len, 80, generator=g) * 2.0 - 5.0).clamp(-11.5, 2.0)
    embedding = torch.randn(1, 192, generator=g)
    return dict(prompt_token=prompt_token, prompt_token_len=torch.tensor([prompt_len]),
                prompt_feat=prompt_feat, prompt_feat_len=None, embedding=embedding)


def speech_tokens(n: int, seed: int = 5) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, S3_VOCAB, (1, n), generator=g)


def cfm_inputs(L: int, seed: int = 6, prompt_frames: int | None = None) -> Dict[str, torch.Tensor]:
    """Inputs of `CausalConditionalCFM.forward` (flow_matching.py:194-218) for one utterance of L mel
    frames: mu (1,80,L) encoder output, spks (1,80) projected x-vector, cond (1,80,L) = prompt mel in
    the first `prompt_frames` frames and zeros after (flow.py:268-270), z = the fixed noise."""
    g = torch.Generator().manual_seed(seed)
    mu = torch.randn(1, 80, L, generator=g)
    spks = torch.randn(1, 80, generator=g) * 0.5
    pf = L // 3 if prompt_frames is None else prompt_frames
    cond = torch.zeros(1, 80, L)
    cond[:, :, :pf] = (torch.randn(1, 80, pf, generator=g) * 2.0 - 5.0).clamp(-11.5, 2.0)
    return dict(mu=mu, spks=spks, cond=cond, z=cfm_noise()[:, :, :L].clone())


def mel_input(L: int, seed: int = 7) -> torch.Tensor:
    """(1, 80, L) log-mel-like features (utils/mel.py range) for the vocoder tests."""
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(1, 80, L, generator=g) * 2.0 - 5.0).clamp(-11.5, 2.0)
