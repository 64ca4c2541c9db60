"""ORACLE — test infrastructure, not product code.

CPU restatement (plain torch functional ops, fp32 or fp64, no `transformers`) of the reference's
Tokenizer-12Hz decode path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product path (qwen3_tts_b200) never does.

Parity pinning: the reference's own tests hold NO golden vectors for this path, so
this restatement is pinned against the unmodified reference module executed through
oracle/ref_shim.py in the build container (tests/test_oracle_vs_reference.py, fixtures in
tests/golden/ written by oracle/make_golden.py).

Every function cites the reference lines it follows; "modeling" is
/root/reference/qwen_tts/core/tokenizer_12hz/modeling_qwen3_tts_tokenizer_v2.py.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]


def _w(sd: SD, key: str, dtype) -> torch.Tensor:
    return sd[key].to(dtype)


# ----------------------------------------------------------------------------- RVQ (modeling :660-820)
def codebook_table(sd: SD, prefix: str, dtype) -> torch.Tensor:
    """EuclideanCodebook.decode's table: embedding_sum / clamp(cluster_usage, 1e-5) (modeling :675-676)."""
    return _w(sd, prefix + "embedding_sum", dtype) / _w(sd, prefix + "cluster_usage", dtype).clamp(min=1e-5)[:, None]


def rvq_decode(sd: SD, cfg, codes: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
    """SplitResidualVectorQuantizer.decode (modeling :814-820): codes [B, K, T] -> [B, D, T].

    rvq_first handles codebook 0, rvq_rest codebooks 1..K-1; each is a left-to-right sum of
    embedding rows (modeling :720-726) followed by its own 1x1 output_proj (modeling :772-776)."""
    def one(name: str, c: torch.Tensor) -> torch.Tensor:
        q = None
        for k in range(c.shape[1]):
            tab = codebook_table(sd, f"quantizer.{name}.vq.layers.{k}._codebook.", dtype)
            e = F.embedding(c[:, k], tab).transpose(1, 2)  # VectorQuantization.decode (modeling :706-710)
            q = e if q is None else q + e
        return F.conv1d(q, _w(sd, f"quantizer.{name}.output_proj.weight", dtype))
    out = one("rvq_first", codes[:, :1])
    if codes.shape[1] > 1:
        out = out + one("rvq_rest", codes[:, 1:])
    return out


# ----------------------------------------------------------------------------- convs (modeling :159-207)
def causal_conv(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], dilation: int = 1, groups: int = 1) -> torch.Tensor:
    """Qwen3TTSTokenizerV2CausalConvNet.forward, stride 1: left-pad (k-1)*dilation zeros (modeling :179-192)."""
    k = w.shape[-1]
    x = F.pad(x, ((k - 1) * dilation, 0))
    return F.conv1d(x, w, b, dilation=dilation, groups=groups)


def causal_trans_conv(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, stride: int) -> torch.Tensor:
    """Qwen3TTSTokenizerV2CausalTransConvNet.forward: ConvTranspose1d then trim (k - stride) from
    BOTH ends (modeling :200-207)."""
    k = w.shape[-1]
    y = F.conv_transpose1d(x, w, b, stride=stride)
    pad = k - stride
    return y[..., pad: y.shape[-1] - pad]


def snake_beta(x: torch.Tensor, alpha: torch.Tensor, beta: torch.Tensor) -> torch.Tensor:
    """SnakeBeta.forward (modeling :601-615): x + 1/(exp(beta)+1e-9) * sin(x*exp(alpha))^2."""
    a = torch.exp(alpha)[None, :, None]
    bt = torch.exp(beta)[None, :, None]
    return x + (1.0 / (bt + 1e-9)) * torch.pow(torch.sin(x * a), 2)


def convnext_block(sd: SD, p: str, x: torch.Tensor, dtype) -> torch.Tensor:
    """Qwen3TTSTokenizerV2ConvNeXtBlock.forward (modeling :226-242)."""
    C = x.shape[1]
    h = causal_conv(x, _w(sd, p + "dwconv.conv.weight", dtype), _w(sd, p + "dwconv.conv.bias", dtype), groups=C)
    h = h.permute(0, 2, 1)
    h = F.layer_norm(h, (C,), _w(sd, p + "norm.weight", dtype), _w(sd, p + "norm.bias", dtype), eps=1e-6)
    h = F.linear(h, _w(sd, p + "pwconv1.weight", dtype), _w(sd, p + "pwconv1.bias", dtype))
    h = F.gelu(h)  # nn.GELU() default = exact erf
    h = F.linear(h, _w(sd, p + "pwconv2.weight", dtype), _w(sd, p + "pwconv2.bias", dtype))
    h = _w(sd, p + "gamma", dtype) * h
    return x + h.permute(0, 2, 1)


# ----------------------------------------------------------------------------- transformer (modeling :75-144, :245-574)
def rms_norm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    """Qwen3TTSTokenizerV2DecoderRMSNorm.forward (modeling :382-387)."""
    var = x.pow(2).mean(-1, keepdim=True)
    return w * (x * torch.rsqrt(var + eps))


def rope_cos_sin(cfg, T: int, dtype) -> Tuple[torch.Tensor, torch.Tensor]:
    """DecoderRotatoryEmbedding.forward with rope_type 'default' (modeling :259-278): positions
    restart at 0 for every forward call because no cache is passed (modeling :524-531, :875)."""
    d = cfg.head_dim
    inv = 1.0 / (cfg.rope_theta ** (torch.arange(0, d, 2, dtype=torch.int64).to(torch.float32) / d))
    pos = torch.arange(T, dtype=torch.float32)
    freqs = torch.outer(pos, inv)  # computed in fp32 like the reference (autocast disabled, :272-276)
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos().to(dtype), emb.sin().to(dtype)


def _rotate_half(x: torch.Tensor) -> torch.Tensor:
    x1, x2 = x[..., : x.shape[-1] // 2], x[..., x.shape[-1] // 2:]
    return torch.cat((-x2, x1), dim=-1)  # modeling :75-79


def attention(sd: SD, cfg, p: str, x: torch.Tensor, cos, sin, dtype) -> torch.Tensor:
    """DecoderAttention.forward (modeling :311-353) with the eager softmax path (:121-144) and the
    mask transformers builds for sliding attention: kv <= q and kv > q - sliding_window."""
    B, T, _ = x.shape
    nh, nkv, hd = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
    q = F.linear(x, _w(sd, p + "q_proj.weight", dtype)).view(B, T, nh, hd).transpose(1, 2)
    k = F.linear(x, _w(sd, p + "k_proj.weight", dtype)).view(B, T, nkv, hd).transpose(1, 2)
    v = F.linear(x, _w(sd, p + "v_proj.weight", dtype)).view(B, T, nkv, hd).transpose(1, 2)
    c, s = cos[None, None], sin[None, None]
    q = q * c + _rotate_half(q) * s  # modeling :82-106
    k = k * c + _rotate_half(k) * s
    if nkv != nh:
        rep = nh // nkv
        k = k.repeat_interleave(rep, dim=1)
        v = v.repeat_interleave(rep, dim=1)
    att = torch.matmul(q, k.transpose(2, 3)) * (hd ** -0.5)
    qi = torch.arange(T)[:, None]
    ki = torch.arange(T)[None, :]
    allowed = (ki <= qi) & (ki > qi - cfg.sliding_window)
    att = att.masked_fill(~allowed[None, None], float("-inf"))
    att = torch.softmax(att.to(torch.float32) if dtype != torch.float64 else att, dim=-1).to(dtype)
    o = torch.matmul(att, v).transpose(1, 2).reshape(B, T, nh * hd)
    return F.linear(o, _w(sd, p + "o_proj.weight", dtype))


def transformer(sd: SD, cfg, x: torch.Tensor, dtype) -> torch.Tensor:
    """DecoderTransformerModel.forward (modeling :500-574), inputs_embeds only, no cache."""
    P = "pre_transformer."
    x = F.linear(x, _w(sd, P + "input_proj.weight", dtype), _w(sd, P + "input_proj.bias", dtype))
    cos, sin = rope_cos_sin(cfg, x.shape[1], dtype)
    for l in range(cfg.num_hidden_layers):
        p = f"{P}layers.{l}."
        h = rms_norm(x, _w(sd, p + "input_layernorm.weight", dtype), cfg.rms_norm_eps)
        h = attention(sd, cfg, p + "self_attn.", h, cos, sin, dtype)
        x = x + _w(sd, p + "self_attn_layer_scale.scale", dtype) * h  # modeling :463
        h = rms_norm(x, _w(sd, p + "post_attention_layernorm.weight", dtype), cfg.rms_norm_eps)
        g = F.silu(F.linear(h, _w(sd, p + "mlp.gate_proj.weight", dtype)))
        u = F.linear(h, _w(sd, p + "mlp.up_proj.weight", dtype))
        h = F.linear(g * u, _w(sd, p + "mlp.down_proj.weight", dtype))  # modeling :367-369
        x = x + _w(sd, p + "mlp_layer_scale.scale", dtype) * h  # modeling :469
    x = rms_norm(x, _w(sd, P + "norm.weight", dtype), cfg.rms_norm_eps)
    return F.linear(x, _w(sd, P + "output_proj.weight", dtype), _w(sd, P + "output_proj.bias", dtype))


# ----------------------------------------------------------------------------- vocoder (modeling :618-657)
def residual_unit(sd: SD, p: str, x: torch.Tensor, dilation: int, dtype) -> torch.Tensor:
    """DecoderDecoderResidualUnit.forward (modeling :627-634)."""
    h = snake_beta(x, _w(sd, p + "act1.alpha", dtype), _w(sd, p + "act1.beta", dtype))
    h = causal_conv(h, _w(sd, p + "conv1.conv.weight", dtype), _w(sd, p + "conv1.conv.bias", dtype), dilation=dilation)
    h = snake_beta(h, _w(sd, p + "act2.alpha", dtype), _w(sd, p + "act2.beta", dtype))
    h = causal_conv(h, _w(sd, p + "conv2.conv.weight", dtype), _w(sd, p + "conv2.conv.bias", dtype))
    return h + x


def decoder_block(sd: SD, p: str, x: torch.Tensor, stride: int, dtype, taps: Optional[dict] = None) -> torch.Tensor:
    """DecoderDecoderBlock.forward (modeling :644-657): Snake, TransConv(k=2s, s), 3 residual units."""
    x = snake_beta(x, _w(sd, p + "0.alpha", dtype), _w(sd, p + "0.beta", dtype))
    x = causal_trans_conv(x, _w(sd, p + "1.conv.weight", dtype), _w(sd, p + "1.conv.bias", dtype), stride)
    if taps is not None:
        taps[p + "1"] = x
    for j, d in ((2, 1), (3, 3), (4, 9)):
        x = residual_unit(sd, f"{p}{j}.", x, d, dtype)
        if taps is not None:
            taps[f"{p}{j}"] = x
    return x


# ----------------------------------------------------------------------------- decoder forward / chunking / length rule
def decoder_forward(sd: SD, cfg, codes: torch.Tensor, dtype=torch.float32, taps: Optional[dict] = None,
                    clamp: bool = True) -> torch.Tensor:
    """Qwen3TTSTokenizerV2Decoder.forward (modeling :868-883): codes [B, Q, T] int64 -> [B, 1, T*1920-555]."""
    if codes.shape[1] != cfg.num_quantizers:
        raise ValueError(f"Expected {cfg.num_quantizers} layer of codes, got {codes.shape[1]}")
    def tap(name, v):
        if taps is not None:
            taps[name] = v
    h = rvq_decode(sd, cfg, codes, dtype); tap("quantizer", h)
    h = causal_conv(h, _w(sd, "pre_conv.conv.weight", dtype), _w(sd, "pre_conv.conv.bias", dtype)); tap("pre_conv", h)
    h = transformer(sd, cfg, h.transpose(1, 2), dtype); tap("pre_transformer", h)
    h = h.permute(0, 2, 1)
    for i, f in enumerate(cfg.upsampling_ratios):
        h = causal_trans_conv(h, _w(sd, f"upsample.{i}.0.conv.weight", dtype), _w(sd, f"upsample.{i}.0.conv.bias", dtype), f)
        tap(f"upsample.{i}.0", h)
        h = convnext_block(sd, f"upsample.{i}.1.", h, dtype); tap(f"upsample.{i}.1", h)
    h = causal_conv(h, _w(sd, "decoder.0.conv.weight", dtype), _w(sd, "decoder.0.conv.bias", dtype)); tap("decoder.0", h)
    n = len(cfg.upsample_rates)
    for i, r in enumerate(cfg.upsample_rates):
        h = decoder_block(sd, f"decoder.{i + 1}.block.", h, r, dtype, taps); tap(f"decoder.{i + 1}", h)
    h = snake_beta(h, _w(sd, f"decoder.{n + 1}.alpha", dtype), _w(sd, f"decoder.{n + 1}.beta", dtype))
    h = causal_conv(h, _w(sd, f"decoder.{n + 2}.conv.weight", dtype), _w(sd, f"decoder.{n + 2}.conv.bias", dtype))
    tap("preclamp", h)
    return h.clamp(min=-1, max=1) if clamp else h


def chunk_plan(T: int, chunk_size: int = 300, left_context_size: int = 25) -> List[Tuple[int, int, int]]:
    """(start - ctx, end, ctx) per chunk, exactly as the loop at modeling :887-894 computes them
    (note `start - left_context_size > 0`, so a chunk starting AT frame 25 gets ctx = 25 only
    because start - 25 == 0 fails the test and falls to ctx = start = 25 — same value)."""
    plan, start = [], 0
    while start < T:
        end = min(start + chunk_size, T)
        ctx = left_context_size if start - left_context_size > 0 else start
        plan.append((start - ctx, end, ctx))
        start = end
    return plan


def chunked_decode(sd: SD, cfg, codes: torch.Tensor, chunk_size: int = 300, left_context_size: int = 25,
                   dtype=torch.float32) -> torch.Tensor:
    """Qwen3TTSTokenizerV2Decoder.chunked_decode (modeling :885-895)."""
    up = cfg.total_upsample
    wavs = []
    for lo, hi, ctx in chunk_plan(codes.shape[-1], chunk_size, left_context_size):
        w = decoder_forward(sd, cfg, codes[..., lo:hi], dtype)
        wavs.append(w[..., ctx * up:])
    return torch.cat(wavs, dim=-1)


def model_decode(sd: SD, cfg, audio_codes: torch.Tensor, dtype=torch.float32, decode_upsample_rate: int = 1920) -> List[torch.Tensor]:
    """Qwen3TTSTokenizerV2Model.decode (modeling :1012-1015): audio_codes [B, T, Q] ->
    list of 1-D waveforms cut to (codes[..., 0] > 0).sum(1) * 1920 samples."""
    wav = chunked_decode(sd, cfg, audio_codes.transpose(1, 2), dtype=dtype).squeeze(1)
    lengths = (audio_codes[..., 0] > 0).sum(1) * decode_upsample_rate
    return [a[:l] for a, l in zip(wav, lengths)]


def expected_num_samples(T: int, chunk_size: int = 300, up: int = 1920, lookahead: int = 555) -> int:
    """Length of chunked_decode's output: each chunk loses the 555-sample transposed-conv trim."""
    return T * up - lookahead * len(chunk_plan(T, chunk_size))


def snr_db(ref: torch.Tensor, out: torch.Tensor) -> float:
    ref = ref.double().flatten(); out = out.double().flatten()
    num = (ref * ref).sum().item(); den = ((ref - out) ** 2).sum().item()
    return float("inf") if den == 0 else 10.0 * math.log10(max(num, 1e-300) / den)
