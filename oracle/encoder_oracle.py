"""ORACLE - test infrastructure, not product code.

CPU restatement (float64 by default) of the reference's 12 Hz ENCODE path, SURVEY 8(f)-2:

    Qwen3TTSTokenizerV2Model.encode   /root/reference/qwen_tts/core/tokenizer_12hz/modeling_qwen3_tts_tokenizer_v2.py:960-990
        encoder.encode(input_values.unsqueeze(1))      -> Qwen3TTSTokenizerV2Encoder(MimiModel) (:898-907): MimiModel.encode
        audio_codes[:, :encoder_valid_num_quantizers]  (:982)
        per item: codes[..., :ceil(mask.sum() / encode_downsample_rate)].T   (:983)

MimiModel lives in the third-party `transformers` package (pinned ==4.57.3 by the reference's pyproject.toml:23, 5.5.0 in this image,
NOT vendored under /root/reference); line numbers below are transformers/models/mimi/modeling_mimi.py of this image:

    MimiModel._encode_frame  :1455-1489   encoder -> encoder_transformer -> downsample -> quantizer.encode
    MimiEncoder              :454-496     conv k7 (1 -> 64); per ratio r in (4, 5, 6, 8): ResnetBlock, ELU, conv k = 2r stride r (C -> 2C);
                                          ELU, conv k3 (1024 -> 512)
    MimiResnetBlock          :412-451     x + conv_k1(ELU(conv_k3(ELU(x))))   (hidden = dim / 2, dilation 1, identity shortcut)
    MimiConv1d.forward       :331-352     causal: left pad (k - 1) * dil + 1 - stride, right pad up to a whole last frame (:273-285);
                                          pad_mode "constant" (zeros) everywhere except the downsample conv ("replicate", :1421-1431)
    MimiTransformerLayer     :926-993     x + ls1 * attn(LN1(x));  x + ls2 * fc2(gelu(fc1(LN2(x))))   (LayerNorm with bias, eps 1e-5)
    MimiAttention            :645-738     q/k/v/o without bias, RoPE (rotate-half, theta 1e4, :515-612), scale 1/sqrt(64), causal
                                          sliding-window mask (create_sliding_window_causal_mask, :1106-1112): key j visible to
                                          query i iff 0 <= i - j < sliding_window
    quantizer.encode         :1296-1350   -> oracle/encode_oracle.py

Pinning: tests/golden/encode_*.npz hold latents and codes written by the reference's own Qwen3TTSTokenizerV2Encoder class (imported
unmodified through oracle/ref_shim.py) on seeded weights and audio (oracle/make_golden_encoder.py); tests/test_oracle_encoder.py
checks this file against them and, when /root/reference is present, against the live reference."""
from __future__ import annotations

import math
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from . import encode_oracle as EO

RATIOS = (4, 5, 6, 8)          # reversed(MimiConfig.upsampling_ratios)
P = "encoder."                 # model-level key prefix (Qwen3TTSTokenizerV2Model.encoder)


def causal_conv1d(x: torch.Tensor, w: torch.Tensor, b, stride: int = 1, dilation: int = 1, pad_mode: str = "constant") -> torch.Tensor:
    """MimiConv1d.forward, causal branch (modeling_mimi.py:331-352, padding arithmetic :273-285).  x [B, Cin, L]."""
    k_eff = (w.shape[-1] - 1) * dilation + 1
    pad_total = k_eff - stride
    length = x.shape[-1]
    n_frames = math.ceil((length - k_eff + pad_total) / stride + 1) - 1
    extra = n_frames * stride + k_eff - pad_total - length
    x = F.pad(x, (pad_total, extra), mode=pad_mode)
    return F.conv1d(x, w, b, stride=stride, dilation=dilation)


def seanet_encoder(sd: Dict[str, torch.Tensor], x: torch.Tensor, dt=torch.float64) -> torch.Tensor:
    """MimiEncoder.forward (:454-496).  x [B, 1, T] -> [B, 512, ceil-chain(T / 960)]."""
    def g(k):
        return sd[P + "encoder." + k].to(dt)
    x = causal_conv1d(x, g("layers.0.conv.weight"), g("layers.0.conv.bias"))
    i = 1
    for r in RATIOS:
        h = causal_conv1d(F.elu(x), g(f"layers.{i}.block.1.conv.weight"), g(f"layers.{i}.block.1.conv.bias"))
        h = causal_conv1d(F.elu(h), g(f"layers.{i}.block.3.conv.weight"), g(f"layers.{i}.block.3.conv.bias"))
        x = x + h
        x = causal_conv1d(F.elu(x), g(f"layers.{i + 2}.conv.weight"), g(f"layers.{i + 2}.conv.bias"), stride=r)
        i += 3
    return causal_conv1d(F.elu(x), g(f"layers.{i + 1}.conv.weight"), g(f"layers.{i + 1}.conv.bias"))


def rope_tables(n: int, head_dim: int, theta: float, dt) -> Tuple[torch.Tensor, torch.Tensor]:
    """MimiRotaryEmbedding (:515-576): inv_freq and the position products are evaluated in float32 by the reference."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.int64).float() / head_dim))
    freqs = torch.arange(n).float()[:, None] * inv[None, :]
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos().to(dt), emb.sin().to(dt)


def rotate_half(x: torch.Tensor) -> torch.Tensor:
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


def encoder_transformer(sd: Dict[str, torch.Tensor], x: torch.Tensor, num_heads: int = 8, sliding_window: int = 250, theta: float = 10000.0,
                        eps: float = 1e-5, dt=torch.float64) -> torch.Tensor:
    """MimiTransformerModel.forward (:996-1140) without cache.  x [B, T, 512]."""
    B, T, H = x.shape
    hd = H // num_heads
    cos, sin = rope_tables(T, hd, theta, dt)
    i = torch.arange(T)[:, None]
    j = torch.arange(T)[None, :]
    visible = (j <= i) & (i - j < sliding_window)
    layer = 0
    while P + f"encoder_transformer.layers.{layer}.self_attn.q_proj.weight" in sd:
        def g(k, layer=layer):
            return sd[P + f"encoder_transformer.layers.{layer}." + k].to(dt)
        h = F.layer_norm(x, (H,), g("input_layernorm.weight"), g("input_layernorm.bias"), eps)
        q = (h @ g("self_attn.q_proj.weight").T).view(B, T, num_heads, hd).transpose(1, 2)
        k = (h @ g("self_attn.k_proj.weight").T).view(B, T, num_heads, hd).transpose(1, 2)
        v = (h @ g("self_attn.v_proj.weight").T).view(B, T, num_heads, hd).transpose(1, 2)
        q = q * cos + rotate_half(q) * sin
        k = k * cos + rotate_half(k) * sin
        s = (q @ k.transpose(2, 3)) / math.sqrt(hd)
        s = s.masked_fill(~visible, float("-inf"))
        a = torch.softmax(s, dim=-1) @ v
        a = a.transpose(1, 2).reshape(B, T, H) @ g("self_attn.o_proj.weight").T
        x = x + g("self_attn_layer_scale.scale") * a
        h = F.layer_norm(x, (H,), g("post_attention_layernorm.weight"), g("post_attention_layernorm.bias"), eps)
        h = F.gelu(h @ g("mlp.fc1.weight").T) @ g("mlp.fc2.weight").T
        x = x + g("mlp_layer_scale.scale") * h
        layer += 1
    return x


def encode_latents(sd: Dict[str, torch.Tensor], wav: torch.Tensor, dt=torch.float64, **kw) -> torch.Tensor:
    """MimiModel._encode_frame up to the quantiser (:1455-1486).  wav [B, T] -> latents [B, 512, frames]."""
    x = seanet_encoder(sd, wav.to(dt).unsqueeze(1), dt)
    x = encoder_transformer(sd, x.transpose(1, 2), dt=dt, **kw).transpose(1, 2)
    return causal_conv1d(x, sd[P + "downsample.conv.weight"].to(dt), None, stride=2, pad_mode="replicate")


def num_frames(samples: int) -> int:
    """Frames MimiModel.encode emits for `samples` input samples: every strided conv rounds up (:273-285)."""
    n = samples
    for r in RATIOS + (2,):
        n = -(-n // r)
    return n


def model_encode(sd: Dict[str, torch.Tensor], input_values: torch.Tensor, padding_mask: torch.Tensor, num_quantizers: int = 16,
                 encode_downsample_rate: int = 1920, dt=torch.float64):
    """Qwen3TTSTokenizerV2Model.encode (reference modeling :960-990) -> (list of [T_i, K] int64 arrays, latents [B, 512, T] float64,
    margin [B, K, T]).  The padded batch is encoded as it is (the reference passes no mask into the encoder); every item is then cut
    to ceil(mask.sum() / encode_downsample_rate) frames."""
    with torch.inference_mode():
        lat = encode_latents(sd, input_values, dt)
    codes, margin = EO.rvq_encode(sd, lat, num_quantizers=num_quantizers)
    out: List[np.ndarray] = []
    for c, m in zip(codes, padding_mask):
        n = -(-int(m.sum()) // encode_downsample_rate)
        out.append(np.ascontiguousarray(c[:, :n].T))
    return out, lat.numpy(), margin


def decidable(codes_a: np.ndarray, codes_b: np.ndarray, margin: np.ndarray, tol: float, num_semantic: int = 1) -> Tuple[int, int]:
    """Compare two [K, T] code arrays of one item given the oracle's relative margins [K, T].  Inside a quantiser group (semantic:
    codebook 0; acoustic: the rest - both start from the same latents) a frame's codes must agree up to its first codebook whose
    margin is below `tol` AND was resolved differently: from there on the residuals differ and the group's later codebooks are
    not comparable.  Returns (codes compared, codes that differ although their margin is >= tol)."""
    K, T = codes_a.shape
    compared = wrong = 0
    for t in range(T):
        for lo, hi in ((0, num_semantic), (num_semantic, K)):
            for k in range(lo, hi):
                same = codes_a[k, t] == codes_b[k, t]
                if margin[k, t] >= tol:
                    compared += 1
                    wrong += 0 if same else 1
                if not same:
                    break
    return compared, wrong
