"""Seeded synthetic weights and codes for the Tokenizer-12Hz decoder.

No checkpoint is mounted on the build or GPU boxes, so throughput and parity are measured on
random-init weights of the named architecture.  A freshly constructed reference decoder outputs
exact silence (embedding_sum = 0, Snake alpha = beta = 0, LayerScale 0.01, gamma 1e-6;
reference modeling :673, :596-597, :224), so every parameter is overwritten here with a seeded
draw whose gains keep the pre-clamp waveform inside (-1, 1).

Keys are the reference `Qwen3TTSTokenizerV2Decoder.state_dict()` keys (no `decoder.` prefix);
shapes follow reference modeling :823-866.  Each tensor is drawn from its own generator seeded
by (seed, crc32(key)) so the result does not depend on iteration order.
"""
from __future__ import annotations

import math
import zlib
from collections import OrderedDict
from typing import Dict, List, Tuple

import torch

from .configuration import Qwen3TTSTokenizerV2DecoderConfig


def decoder_param_shapes(cfg: Qwen3TTSTokenizerV2DecoderConfig) -> "OrderedDict[str, Tuple[int, ...]]":
    """All persistent tensors of the reference decoder, in module order."""
    s: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    H, Lat, I = cfg.hidden_size, cfg.latent_dim, cfg.intermediate_size
    nh, nkv, hd = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
    D, dq = cfg.codebook_dim, cfg.codebook_dim // 2
    # pre_transformer (modeling :481-493)
    for l in range(cfg.num_hidden_layers):
        p = f"pre_transformer.layers.{l}."
        s[p + "self_attn.q_proj.weight"] = (nh * hd, H)
        s[p + "self_attn.k_proj.weight"] = (nkv * hd, H)
        s[p + "self_attn.v_proj.weight"] = (nkv * hd, H)
        s[p + "self_attn.o_proj.weight"] = (H, nh * hd)
        s[p + "mlp.gate_proj.weight"] = (I, H)
        s[p + "mlp.up_proj.weight"] = (I, H)
        s[p + "mlp.down_proj.weight"] = (H, I)
        s[p + "input_layernorm.weight"] = (H,)
        s[p + "post_attention_layernorm.weight"] = (H,)
        s[p + "self_attn_layer_scale.scale"] = (H,)
        s[p + "mlp_layer_scale.scale"] = (H,)
    s["pre_transformer.norm.weight"] = (H,)
    s["pre_transformer.input_proj.weight"] = (H, Lat)
    s["pre_transformer.input_proj.bias"] = (H,)
    s["pre_transformer.output_proj.weight"] = (Lat, H)
    s["pre_transformer.output_proj.bias"] = (Lat,)
    # quantizer (modeling :829-836, :757-770)
    for name, nq in (("rvq_first", 1), ("rvq_rest", cfg.num_quantizers - 1)):
        s[f"quantizer.{name}.input_proj.weight"] = (dq, D, 1)
        s[f"quantizer.{name}.output_proj.weight"] = (D, dq, 1)
        for k in range(nq):
            s[f"quantizer.{name}.vq.layers.{k}._codebook.cluster_usage"] = (cfg.codebook_size,)
            s[f"quantizer.{name}.vq.layers.{k}._codebook.embedding_sum"] = (cfg.codebook_size, dq)
    # pre_conv (modeling :838-842)
    s["pre_conv.conv.weight"] = (Lat, D, 3)
    s["pre_conv.conv.bias"] = (Lat,)
    # upsample (modeling :844-854, :210-224)
    for i, f in enumerate(cfg.upsampling_ratios):
        p = f"upsample.{i}."
        s[p + "0.conv.weight"] = (Lat, Lat, f)
        s[p + "0.conv.bias"] = (Lat,)
        s[p + "1.dwconv.conv.weight"] = (Lat, 1, 7)
        s[p + "1.dwconv.conv.bias"] = (Lat,)
        s[p + "1.norm.weight"] = (Lat,)
        s[p + "1.norm.bias"] = (Lat,)
        s[p + "1.pwconv1.weight"] = (4 * Lat, Lat)
        s[p + "1.pwconv1.bias"] = (4 * Lat,)
        s[p + "1.pwconv2.weight"] = (Lat, 4 * Lat)
        s[p + "1.pwconv2.bias"] = (Lat,)
        s[p + "1.gamma"] = (Lat,)
    # decoder (modeling :856-864, :637-657, :618-634)
    Dd = cfg.decoder_dim
    s["decoder.0.conv.weight"] = (Dd, Lat, 7)
    s["decoder.0.conv.bias"] = (Dd,)
    for i, r in enumerate(cfg.upsample_rates):
        cin, cout = Dd // 2 ** i, Dd // 2 ** (i + 1)
        p = f"decoder.{i + 1}.block."
        s[p + "0.alpha"] = (cin,)
        s[p + "0.beta"] = (cin,)
        s[p + "1.conv.weight"] = (cin, cout, 2 * r)
        s[p + "1.conv.bias"] = (cout,)
        for j in (2, 3, 4):
            s[p + f"{j}.act1.alpha"] = (cout,)
            s[p + f"{j}.act1.beta"] = (cout,)
            s[p + f"{j}.conv1.conv.weight"] = (cout, cout, 7)
            s[p + f"{j}.conv1.conv.bias"] = (cout,)
            s[p + f"{j}.act2.alpha"] = (cout,)
            s[p + f"{j}.act2.beta"] = (cout,)
            s[p + f"{j}.conv2.conv.weight"] = (cout, cout, 1)
            s[p + f"{j}.conv2.conv.bias"] = (cout,)
    n = len(cfg.upsample_rates)
    cl = Dd // 2 ** n
    s[f"decoder.{n + 1}.alpha"] = (cl,)
    s[f"decoder.{n + 1}.beta"] = (cl,)
    s[f"decoder.{n + 2}.conv.weight"] = (1, cl, 7)
    s[f"decoder.{n + 2}.conv.bias"] = (1,)
    return s


def _gen(seed: int, key: str) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed((int(seed) * 1000003 + zlib.crc32(key.encode())) % (2 ** 63 - 1))
    return g


def _fan_in(key: str, shape: Tuple[int, ...]) -> int:
    if key.endswith("1.conv.weight") and ".block." in key:  # ConvTranspose1d [Cin, Cout, k]: 2 taps hit each output
        return shape[0] * 2
    if key.startswith("upsample.") and key.endswith("0.conv.weight"):  # k = stride: 1 tap per output
        return shape[0]
    if len(shape) == 3:
        return shape[1] * shape[2]
    return shape[1]


def make_decoder_state_dict(cfg: Qwen3TTSTokenizerV2DecoderConfig, seed: int = 0, stress: bool = False) -> Dict[str, torch.Tensor]:
    """Deterministic fp32 CPU state dict with reference key names.

    stress=True draws Snake beta from N(0, 0.3) exactly as SURVEY 8(c) words it (instead of 1 + N(0, 0.3)): the periodic term
    of every Snake is then ~2.7x larger and the 29-Snake stack amplifies any perturbation ~10x more (fp32 vs fp64 99 dB instead
    of 113-121 dB).  It is the ill-conditioned end of what a checkpoint could be; parity numbers are reported for both
    (profiles/r2_parity.json)."""
    out: Dict[str, torch.Tensor] = OrderedDict()
    for key, shape in decoder_param_shapes(cfg).items():
        g = _gen(seed, key)
        if key.endswith("embedding_sum"):
            t = torch.randn(shape, generator=g)
        elif key.endswith("cluster_usage"):
            t = torch.rand(shape, generator=g) * 2.0 + 0.5
        elif key.endswith(".alpha"):
            t = torch.randn(shape, generator=g) * 0.3
        elif key.endswith(".beta"):
            # log-scale magnitude parameter; mean +1 keeps the periodic term's slope a*ib*sin(2ax) well
            # below 1 so the random-init stack amplifies perturbations about as little as the survey's
            # reference run did (fp32 vs fp64 ~ 113-121 dB) -- with mean 0 it is chaotic
            # (fp32 vs fp64 only 99 dB) and no 16-bit operand scheme can be compared meaningfully.
            t = (0.0 if stress else 1.0) + torch.randn(shape, generator=g) * 0.3
        elif key.endswith("gamma") or key.endswith("layer_scale.scale"):
            t = torch.rand(shape, generator=g) * 0.5 + 0.25
        elif key.endswith("layernorm.weight") or key.endswith("norm.weight"):
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif key.endswith("norm.bias") or key.endswith(".bias"):
            t = 0.02 * torch.randn(shape, generator=g)
        elif key.endswith(".weight"):
            gain = _GAIN(key)
            t = torch.randn(shape, generator=g) * (gain / math.sqrt(_fan_in(key, shape)))
        else:  # pragma: no cover
            raise KeyError(key)
        out[key] = t.contiguous()
    return out


def _GAIN(key: str) -> float:
    # Gains chosen so activations stay O(1) through the stack and the pre-clamp waveform stays
    # inside (-1, 1) (a naive gain-1 fan-in init clamps ~90 % of samples).
    if key.startswith("quantizer.") and "output_proj" in key:
        return 0.35  # sum of up to 15 unit-variance codebook rows feeds this
    if key.startswith("decoder.") and ".block." in key and "conv2.conv" in key:
        return 0.35
    if key.startswith("decoder.") and ".block." in key and "conv1.conv" in key:
        return 0.5
    if key.startswith("decoder.") and key.endswith("conv.weight") and ".block." not in key and key != "decoder.0.conv.weight":
        return 0.09  # final 96 -> 1 projection
    return 1.0


ENCODER_RATIOS = (4, 5, 6, 8)     # reversed(MimiConfig.upsampling_ratios): strides of the encoder's four down-sampling convs


def encoder_param_shapes(num_filters: int = 64, hidden: int = 512, layers: int = 8, intermediate: int = 2048, codebook_size: int = 2048,
                         codebook_dim: int = 512, num_quantizers: int = 16) -> "OrderedDict[str, Tuple[int, ...]]":
    """Key names and shapes of `Qwen3TTSTokenizerV2Model.encoder` (reference modeling :898-907: HF MimiModel with the decoder parts
    nulled; MimiConfig defaults), prefixed `encoder.` as in the model's state dict.  Only the first `num_quantizers` codebooks are
    listed: the reference keeps `encoder_valid_num_quantizers` = 16 of Mimi's 32 (modeling :982)."""
    s: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    e = "encoder.encoder.layers."
    s[e + "0.conv.weight"] = (num_filters, 1, 7)
    s[e + "0.conv.bias"] = (num_filters,)
    c, i = num_filters, 1
    for r in ENCODER_RATIOS:
        s[e + f"{i}.block.1.conv.weight"] = (c // 2, c, 3)
        s[e + f"{i}.block.1.conv.bias"] = (c // 2,)
        s[e + f"{i}.block.3.conv.weight"] = (c, c // 2, 1)
        s[e + f"{i}.block.3.conv.bias"] = (c,)
        s[e + f"{i + 2}.conv.weight"] = (2 * c, c, 2 * r)
        s[e + f"{i + 2}.conv.bias"] = (2 * c,)
        c, i = 2 * c, i + 3
    s[e + f"{i + 1}.conv.weight"] = (hidden, c, 3)
    s[e + f"{i + 1}.conv.bias"] = (hidden,)
    for l in range(layers):
        p = f"encoder.encoder_transformer.layers.{l}."
        for n in ("q_proj", "k_proj", "v_proj", "o_proj"):
            s[p + f"self_attn.{n}.weight"] = (hidden, hidden)
        s[p + "mlp.fc1.weight"] = (intermediate, hidden)
        s[p + "mlp.fc2.weight"] = (hidden, intermediate)
        for n in ("input_layernorm", "post_attention_layernorm"):
            s[p + n + ".weight"] = (hidden,)
            s[p + n + ".bias"] = (hidden,)
        s[p + "self_attn_layer_scale.scale"] = (hidden,)
        s[p + "mlp_layer_scale.scale"] = (hidden,)
    s["encoder.downsample.conv.weight"] = (hidden, hidden, 4)
    dq = codebook_dim // 2
    for grp, n in (("semantic", 1), ("acoustic", num_quantizers - 1)):
        q = f"encoder.quantizer.{grp}_residual_vector_quantizer."
        s[q + "input_proj.weight"] = (dq, codebook_dim, 1)
        for l in range(n):
            s[q + f"layers.{l}.codebook.embed_sum"] = (codebook_size, dq)
            s[q + f"layers.{l}.codebook.cluster_usage"] = (codebook_size,)
    return s


def make_encoder_state_dict(seed: int = 0, **shape_kwargs) -> Dict[str, torch.Tensor]:
    """Deterministic fp32 CPU state dict of the 12 Hz ENCODER with the reference's key names (no checkpoint is mounted).  Gains keep
    the activations O(1) from a waveform of amplitude ~0.1 to the 512-d latents, so the RVQ search sees latents of the scale its
    (synthetic) codebooks were drawn at."""
    out: Dict[str, torch.Tensor] = OrderedDict()
    for key, shape in encoder_param_shapes(**shape_kwargs).items():
        g = _gen(seed + 7919, key)
        if key.endswith("embed_sum"):
            layer = int(key.split("layers.")[1].split(".")[0])
            t = torch.randn(shape, generator=g) * (0.7 ** layer)            # later layers quantise ever smaller residuals
        elif key.endswith("cluster_usage"):
            t = torch.rand(shape, generator=g) * 2.0 + 0.5
        elif key.endswith("layer_scale.scale"):
            t = torch.rand(shape, generator=g) * 0.3 + 0.1
        elif key.endswith("layernorm.weight"):
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif key.endswith(".bias"):
            t = 0.02 * torch.randn(shape, generator=g)
        elif key.endswith(".weight"):
            fan_in = shape[1] * (shape[2] if len(shape) == 3 else 1)
            # measured on make_audio(): every stage's std stays between 1.1 and 1.3, the transformer changes the signal by ~0.7 std
            if key.endswith("layers.0.conv.weight"):
                gain = 6.0
            elif "_proj" in key or "mlp." in key or "downsample" in key:
                gain = 1.0
            else:
                gain = 0.6 if "block.3" in key else 1.2
            t = torch.randn(shape, generator=g) * (gain / math.sqrt(fan_in))
        else:  # pragma: no cover
            raise KeyError(key)
        out[key] = t.contiguous()
    return out


def make_audio(batch: int, samples: int, seed: int = 0, lengths=None) -> torch.Tensor:
    """Seeded synthetic 24 kHz audio [batch, samples] in (-1, 1): a few drifting partials plus noise, zero beyond every item's length
    (the reference's feature extractor pads with zeros)."""
    g = torch.Generator().manual_seed(99991 + seed)
    t = torch.arange(samples, dtype=torch.float32) / 24000.0
    wav = torch.zeros((batch, samples))
    for b in range(batch):
        f0 = 90.0 + 160.0 * torch.rand((), generator=g)
        for h in range(1, 6):
            ph = 2 * math.pi * torch.rand((), generator=g)
            wav[b] += (0.25 / h) * torch.sin(2 * math.pi * f0 * h * t * (1.0 + 0.02 * torch.sin(2 * math.pi * 0.7 * t)) + ph)
        wav[b] += 0.03 * torch.randn((samples,), generator=g)
        wav[b] *= 0.5 + 0.5 * torch.sin(2 * math.pi * 1.3 * t + b) ** 2
        if lengths is not None:
            wav[b, int(lengths[b]):] = 0.0
    return wav.clamp_(-1.0, 1.0)


def make_codes(batch: int, frames: int, cfg: Qwen3TTSTokenizerV2DecoderConfig, seed: int = 0,
               lengths: List[int] | None = None) -> torch.Tensor:
    """[B, T, Q] int64.  Codebook 0 is drawn from [1, size) because the reference counts a 0 in
    codebook 0 as padding (modeling :1014); padding frames (t >= lengths[b]) are all-zero like
    `pad_sequence(padding_value=0)` in the wrapper (qwen3_tts_tokenizer.py:329)."""
    g = _gen(seed, f"codes/{batch}/{frames}")
    q = cfg.num_quantizers
    codes = torch.randint(0, cfg.codebook_size, (batch, frames, q), generator=g, dtype=torch.int64)
    codes[..., 0] = torch.randint(1, cfg.codebook_size, (batch, frames), generator=g, dtype=torch.int64)
    if lengths is not None:
        for b, n in enumerate(lengths):
            codes[b, int(n):] = 0
    return codes
