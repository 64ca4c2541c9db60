"""Configuration mirror for the Tokenizer-12Hz decode path.

Mirrors the fields the reference decoder reads
(reference: qwen_tts/core/tokenizer_12hz/configuration_qwen3_tts_tokenizer_v2.py:72-114 for the
decoder config, :137-170 for the outer tokenizer config).  Plain Python on purpose: the hot path
needs the numbers, not HuggingFace's PretrainedConfig machinery.

`codebook_dim` has NO default in the reference (it reaches the model through **kwargs from the
checkpoint's config.json, modeling_qwen3_tts_tokenizer_v2.py:830-839).  We default it to 512
(the Mimi latent width) and read the real value from config.json when a checkpoint is given.
"""
from __future__ import annotations

import json
import os
from typing import Any, Dict, Optional, Tuple


class Qwen3TTSTokenizerV2DecoderConfig:
    model_type = "qwen3_tts_tokenizer_12hz_decoder"

    def __init__(
        self,
        codebook_size: int = 2048,
        hidden_size: int = 1024,
        latent_dim: int = 1024,
        max_position_embeddings: int = 8000,
        rope_theta: float = 10000,
        num_attention_heads: int = 16,
        num_key_value_heads: int = 16,
        attention_bias: bool = False,
        sliding_window: int = 72,
        intermediate_size: int = 3072,
        hidden_act: str = "silu",
        layer_scale_initial_scale: float = 0.01,
        rms_norm_eps: float = 1e-5,
        num_hidden_layers: int = 8,
        num_quantizers: int = 16,
        upsample_rates: Tuple[int, ...] = (8, 5, 4, 3),
        upsampling_ratios: Tuple[int, ...] = (2, 2),
        decoder_dim: int = 1536,
        attention_dropout: float = 0.0,
        codebook_dim: int = 512,
        head_dim: Optional[int] = None,
        **kwargs: Any,
    ):
        self.codebook_size = int(codebook_size)
        self.hidden_size = int(hidden_size)
        self.latent_dim = int(latent_dim)
        self.max_position_embeddings = int(max_position_embeddings)
        self.rope_theta = float(rope_theta)
        self.num_attention_heads = int(num_attention_heads)
        self.num_key_value_heads = int(num_key_value_heads)
        self.attention_bias = bool(attention_bias)
        self.sliding_window = int(sliding_window)
        self.intermediate_size = int(intermediate_size)
        self.hidden_act = hidden_act
        self.layer_scale_initial_scale = float(layer_scale_initial_scale)
        self.rms_norm_eps = float(rms_norm_eps)
        self.num_hidden_layers = int(num_hidden_layers)
        self.num_quantizers = int(num_quantizers)
        self.upsample_rates = tuple(int(x) for x in upsample_rates)
        self.upsampling_ratios = tuple(int(x) for x in upsampling_ratios)
        self.decoder_dim = int(decoder_dim)
        self.attention_dropout = float(attention_dropout)
        self.codebook_dim = int(codebook_dim)
        # reference: getattr(config, "head_dim", hidden_size // num_attention_heads) (modeling :288)
        self.head_dim = int(head_dim) if head_dim is not None else self.hidden_size // self.num_attention_heads
        self.extra = dict(kwargs)

    @property
    def layer_types(self):
        # reference configuration :116-121 — every layer is sliding attention
        return ["sliding_attention"] * self.num_hidden_layers

    @property
    def total_upsample(self) -> int:
        # reference modeling :826
        t = 1
        for r in self.upsample_rates + self.upsampling_ratios:
            t *= r
        return t

    def to_dict(self) -> Dict[str, Any]:
        d = {k: v for k, v in self.__dict__.items() if k != "extra"}
        d["upsample_rates"] = list(self.upsample_rates)
        d["upsampling_ratios"] = list(self.upsampling_ratios)
        return d


class Qwen3TTSTokenizerV2Config:
    model_type = "qwen3_tts_tokenizer_12hz"

    def __init__(
        self,
        encoder_config: Optional[dict] = None,
        decoder_config: Optional[dict] = None,
        encoder_valid_num_quantizers: int = 16,
        input_sample_rate: int = 24000,
        output_sample_rate: int = 24000,
        decode_upsample_rate: int = 1920,
        encode_downsample_rate: int = 1920,
        **kwargs: Any,
    ):
        if isinstance(decoder_config, Qwen3TTSTokenizerV2DecoderConfig):
            self.decoder_config = decoder_config
        else:
            self.decoder_config = Qwen3TTSTokenizerV2DecoderConfig(**(decoder_config or {}))
        # MimiConfig fields as a plain dict (the reference builds transformers' MimiConfig from it, configuration :155-158); the encode
        # path reads what tensor shapes cannot tell from here: num_attention_heads, sliding_window, max_position_embeddings, rope_theta, norm_eps
        self.encoder_config = encoder_config or {}
        self.encoder_valid_num_quantizers = int(encoder_valid_num_quantizers)
        self.input_sample_rate = int(input_sample_rate)
        self.output_sample_rate = int(output_sample_rate)
        self.decode_upsample_rate = int(decode_upsample_rate)
        self.encode_downsample_rate = int(encode_downsample_rate)
        self.return_dict = bool(kwargs.pop("return_dict", True))
        self.extra = dict(kwargs)

    @classmethod
    def from_pretrained(cls, path: str) -> "Qwen3TTSTokenizerV2Config":
        with open(os.path.join(path, "config.json"), "r") as f:
            d = json.load(f)
        d.pop("model_type", None)
        return cls(**d)
