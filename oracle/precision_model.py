"""ORACLE-side numerics model (test infrastructure): the fp32 restatement with 16-bit rounding injected
where the CUDA path rounds (GEMM operands, stored activations).  Used on CPU to choose the precision
scheme and to predict the SNR the GPU path should reach; never imported by the product path."""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import decoder_oracle as O


class Scheme:
    """Which tensors are rounded to `dt` (torch.float16 / torch.bfloat16 / None = keep fp32)."""

    def __init__(self, dt=torch.float16, early_act=True, early_w=True, tconv_act=True, tconv_w=True, unit_act=True, unit_w=True,
                 stream=True, rvq=True):
        self.dt = dt
        self.early_act, self.early_w = early_act, early_w
        self.tconv_act, self.tconv_w = tconv_act, tconv_w
        self.unit_act, self.unit_w = unit_act, unit_w
        self.stream = stream  # vocoder residual stream stored in 16 bit
        self.rvq = rvq

    def r(self, x, on=True):
        if not on or self.dt is None:
            return x
        return x.to(self.dt).to(torch.float32)


def forward(sd, cfg, codes, s: Scheme):
    f32 = torch.float32
    W = lambda k, on: s.r(sd[k].to(f32), on)
    V = lambda k: sd[k].to(f32)
    h = s.r(O.rvq_decode(sd, cfg, codes, f32), s.rvq)
    h = O.causal_conv(h, W("pre_conv.conv.weight", s.early_w), V("pre_conv.conv.bias"))
    x = s.r(h, s.early_act).transpose(1, 2)
    P = "pre_transformer."
    x = F.linear(x, W(P + "input_proj.weight", s.early_w), V(P + "input_proj.bias"))
    cos, sin = O.rope_cos_sin(cfg, x.shape[1], f32)
    B, T, _ = x.shape
    nh, hd = cfg.num_attention_heads, cfg.head_dim
    for l in range(cfg.num_hidden_layers):
        p = f"{P}layers.{l}."
        hn = s.r(O.rms_norm(x, V(p + "input_layernorm.weight"), cfg.rms_norm_eps), s.early_act)
        qkv = [s.r(F.linear(hn, W(p + f"self_attn.{n}_proj.weight", s.early_w)), s.early_act).view(B, T, nh, hd).transpose(1, 2) for n in "qkv"]
        q, k, v = qkv
        c, sn = cos[None, None], sin[None, None]
        q = q * c + O._rotate_half(q) * sn
        k = s.r(k * c + O._rotate_half(k) * sn, s.early_act)
        att = torch.matmul(q, k.transpose(2, 3)) * (hd ** -0.5)
        qi = torch.arange(T)[:, None]; ki = torch.arange(T)[None, :]
        att = att.masked_fill(~((ki <= qi) & (ki > qi - cfg.sliding_window))[None, None], float("-inf"))
        o = s.r(torch.matmul(torch.softmax(att, -1), v).transpose(1, 2).reshape(B, T, nh * hd), s.early_act)
        x = x + V(p + "self_attn_layer_scale.scale") * F.linear(o, W(p + "self_attn.o_proj.weight", s.early_w))
        hn = s.r(O.rms_norm(x, V(p + "post_attention_layernorm.weight"), cfg.rms_norm_eps), s.early_act)
        g = F.silu(F.linear(hn, W(p + "mlp.gate_proj.weight", s.early_w))) * F.linear(hn, W(p + "mlp.up_proj.weight", s.early_w))
        x = x + V(p + "mlp_layer_scale.scale") * F.linear(s.r(g, s.early_act), W(p + "mlp.down_proj.weight", s.early_w))
    hn = s.r(O.rms_norm(x, V(P + "norm.weight"), cfg.rms_norm_eps), s.early_act)
    h = s.r(F.linear(hn, W(P + "output_proj.weight", s.early_w), V(P + "output_proj.bias")), s.early_act).permute(0, 2, 1)
    for i, f in enumerate(cfg.upsampling_ratios):
        p = f"upsample.{i}."
        h = s.r(O.causal_trans_conv(h, W(p + "0.conv.weight", s.early_w), V(p + "0.conv.bias"), f), s.early_act)
        C = h.shape[1]
        t = O.causal_conv(h, V(p + "1.dwconv.conv.weight"), V(p + "1.dwconv.conv.bias"), groups=C).permute(0, 2, 1)
        t = s.r(F.layer_norm(t, (C,), V(p + "1.norm.weight"), V(p + "1.norm.bias"), eps=1e-6), s.early_act)
        t = s.r(F.gelu(F.linear(t, W(p + "1.pwconv1.weight", s.early_w), V(p + "1.pwconv1.bias"))), s.early_act)
        t = F.linear(t, W(p + "1.pwconv2.weight", s.early_w), V(p + "1.pwconv2.bias"))
        h = s.r(h + (V(p + "1.gamma") * t).permute(0, 2, 1), s.early_act)
    h = O.causal_conv(h, W("decoder.0.conv.weight", s.early_w), V("decoder.0.conv.bias"))
    n = len(cfg.upsample_rates)
    for i, r in enumerate(cfg.upsample_rates):
        p = f"decoder.{i + 1}.block."
        a = s.r(O.snake_beta(h, V(p + "0.alpha"), V(p + "0.beta")), s.tconv_act)
        y = O.causal_trans_conv(a, W(p + "1.conv.weight", s.tconv_w), V(p + "1.conv.bias"), r)
        for j, d in ((2, 1), (3, 3), (4, 9)):
            u = f"{p}{j}."
            a1 = s.r(O.snake_beta(y, V(u + "act1.alpha"), V(u + "act1.beta")), s.unit_act)
            y = s.r(y, s.stream)
            t = O.causal_conv(a1, W(u + "conv1.conv.weight", s.unit_w), V(u + "conv1.conv.bias"), dilation=d)
            a2 = s.r(O.snake_beta(t, V(u + "act2.alpha"), V(u + "act2.beta")), s.unit_act)
            y = y + O.causal_conv(a2, W(u + "conv2.conv.weight", s.unit_w), V(u + "conv2.conv.bias"))
        h = y
    a = s.r(O.snake_beta(h, V(f"decoder.{n + 1}.alpha"), V(f"decoder.{n + 1}.beta")), s.unit_act)
    w = O.causal_conv(a, V(f"decoder.{n + 2}.conv.weight"), V(f"decoder.{n + 2}.conv.bias"))
    return w.clamp(-1, 1)
