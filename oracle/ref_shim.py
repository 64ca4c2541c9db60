"""ORACLE — test infrastructure.  Imports the UNMODIFIED reference decoder from /root/reference.

Reads /root/reference in the build container; on the GPU box (no /root/reference) the same two files staged byte for byte under
the git-ignored oracle/_ref/ by oracle/stage_ref.py.  Used by oracle/make_golden.py to write tests/golden/, by
tests/test_oracle_vs_reference.py and by bench.py's baseline legs (--impl reference, cpu_baseline, torch_gpu context).  Two shims are needed because the reference pins
transformers==4.57.3 (pyproject.toml:23) and this image has 5.5.0:
  1. `@check_model_inputs()` (modeling :498) — 5.x decorator takes the function directly;
  2. `ROPE_INIT_FUNCTIONS["default"]` (modeling :259) — dropped in 5.x.
`qwen_tts/__init__` is bypassed (it imports librosa/soundfile, absent here): the two
tokenizer_12hz files are loaded directly under a synthetic parent package.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

import torch

REF_ROOT = os.environ.get("Q3TTS_REFERENCE_ROOT", "/root/reference")
_T12 = os.path.join(REF_ROOT, "qwen_tts", "core", "tokenizer_12hz")
if not os.path.isfile(os.path.join(_T12, "modeling_qwen3_tts_tokenizer_v2.py")):
    # GPU box: no /root/reference; the same two files, staged byte for byte by oracle/stage_ref.py (git-ignored oracle/_ref/)
    _T12 = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "qwen_tts", "core", "tokenizer_12hz")
_cache = {}


def available() -> bool:
    return os.path.isfile(os.path.join(_T12, "modeling_qwen3_tts_tokenizer_v2.py"))


def load_reference_modules():
    if "mod" in _cache:
        return _cache["cfg"], _cache["mod"]
    import transformers.modeling_rope_utils as R
    import transformers.utils.generic as G

    _orig = G.check_model_inputs

    def _compat(*a, **k):
        if len(a) == 1 and callable(a[0]) and not k:
            return _orig(a[0])
        return lambda f: _orig(f)

    G.check_model_inputs = _compat

    def _default_rope(config, device=None, seq_len=None, **kw):
        d = getattr(config, "head_dim", None) or config.hidden_size // config.num_attention_heads
        inv = 1.0 / (config.rope_theta ** (torch.arange(0, d, 2, dtype=torch.int64).to(device=device, dtype=torch.float) / d))
        return inv, 1.0

    R.ROPE_INIT_FUNCTIONS.setdefault("default", _default_rope)
    pkg = types.ModuleType("q3ref_t12")
    pkg.__path__ = [_T12]
    sys.modules["q3ref_t12"] = pkg

    def _load(name):
        spec = importlib.util.spec_from_file_location("q3ref_t12." + name, os.path.join(_T12, name + ".py"))
        m = importlib.util.module_from_spec(spec)
        sys.modules["q3ref_t12." + name] = m
        spec.loader.exec_module(m)
        return m

    cfg = _load("configuration_qwen3_tts_tokenizer_v2")
    mod = _load("modeling_qwen3_tts_tokenizer_v2")
    _cache["cfg"], _cache["mod"] = cfg, mod
    return cfg, mod


def build_reference_decoder(state_dict, codebook_dim: int = 512, attn_implementation: str = "eager", **cfg_kwargs):
    """Reference `Qwen3TTSTokenizerV2Decoder` (modeling :823) loaded with `state_dict` (decoder-level keys)."""
    cfgm, mod = load_reference_modules()
    cfg = cfgm.Qwen3TTSTokenizerV2DecoderConfig(codebook_dim=codebook_dim, **cfg_kwargs)
    cfg._attn_implementation = attn_implementation
    dec = mod.Qwen3TTSTokenizerV2Decoder._from_config(cfg).eval()
    missing, unexpected = dec.load_state_dict(state_dict, strict=False)
    missing = [m for m in missing if "inv_freq" not in m]
    if missing or unexpected:
        raise RuntimeError(f"state_dict mismatch: missing={missing[:5]} unexpected={unexpected[:5]}")
    return dec


def reference_model_decode(dec, audio_codes: torch.Tensor, decode_upsample_rate: int = 1920):
    """`Qwen3TTSTokenizerV2Model.decode` body (modeling :1012-1015) over a reference decoder; the
    outer model class is not constructed because it also builds the (out-of-scope) Mimi encoder."""
    with torch.inference_mode():
        wav = dec.chunked_decode(audio_codes.transpose(1, 2)).squeeze(1)
    lengths = (audio_codes[..., 0] > 0).sum(1) * decode_upsample_rate
    return [a[:l] for a, l in zip(wav, lengths)]


def build_reference_encoder(state_dict, attn_implementation: str = "eager"):
    """The reference's own `Qwen3TTSTokenizerV2Encoder` (modeling :898-907: HF MimiModel, decoder parts nulled) with MimiConfig
    defaults (what `Qwen3TTSTokenizerV2Config(encoder_config=None)` builds, configuration :155-158), loaded with the `encoder.*` keys
    of `state_dict`.  Codebooks beyond the ones in `state_dict` keep their (unused) initial values: the reference reads only the
    first encoder_valid_num_quantizers = 16 (modeling :982)."""
    from transformers import MimiConfig
    _, mod = load_reference_modules()
    cfg = MimiConfig()
    cfg._attn_implementation = attn_implementation
    enc = mod.Qwen3TTSTokenizerV2Encoder._from_config(cfg).eval()
    own = enc.state_dict()
    for k, v in state_dict.items():
        if not k.startswith("encoder."):
            continue
        kk = k[len("encoder."):]
        if kk not in own or own[kk].shape != v.shape:
            raise RuntimeError(f"state_dict mismatch at {kk}: {tuple(v.shape)}")
        own[kk] = v.clone()
    enc.load_state_dict(own)
    return enc


def reference_model_encode(enc, input_values: torch.Tensor, padding_mask: torch.Tensor, valid_num_quantizers: int = 16,
                           encode_downsample_rate: int = 1920):
    """`Qwen3TTSTokenizerV2Model.encode` body (modeling :979-985) over a reference encoder -> list of [T_i, 16] LongTensors."""
    with torch.inference_mode():
        frames = enc.encode(input_values=input_values.unsqueeze(1), return_dict=True)
    codes = frames.audio_codes[:, :valid_num_quantizers]
    return [c[..., :-(-m.sum() // encode_downsample_rate)].transpose(0, 1) for c, m in zip(codes, padding_mask)]
