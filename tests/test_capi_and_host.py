"""CPU-side checks: the C-ABI library loads and exports every symbol include/q3tts_decode.h declares,
refuses to run without a CUDA device (no CPU fallback), and the host-side mirror of the reference
interface behaves like the reference on bad input."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from qwen3_tts_b200 import _capi
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2Config, Qwen3TTSTokenizerV2DecoderConfig
from qwen3_tts_b200.synthetic import decoder_param_shapes, make_codes, make_decoder_state_dict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "q3tts_decode.h")).read()
    declared = set(re.findall(r"\b(q3tts_[a-z_0-9]+)\s*\(", hdr))
    assert len(declared) >= 16
    lib = _capi.load_library()
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == set(_capi.EXPORTS)      # the ctypes table binds exactly the header
    # the debug / design-evidence exports live in their own header and are not part of the boundary
    dbg = open(os.path.join(ROOT, "include", "q3tts_debug.h")).read()
    declared_dbg = set(re.findall(r"\b(q3tts_[a-z_0-9]+)\s*\(", dbg))
    assert declared_dbg == set(_capi.DEBUG_EXPORTS) and not (declared_dbg & declared)
    for name in declared_dbg:
        assert hasattr(lib, name), name
    assert not any(w in hdr for w in ("selftest", "bench_mma", "bench_tma", "unit_trace"))


def test_config_struct_matches_header_field_order():
    hdr = open(os.path.join(ROOT, "include", "q3tts_decode.h")).read()
    body = hdr[hdr.index("typedef struct {"):hdr.index("} q3tts_config;")]
    fields = re.findall(r"(?:int32_t|float)\s+([a-z_]+)(?:\[\d+\])?;", body)
    assert fields == [f[0] for f in _capi.q3tts_config._fields_]


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    lib = _capi.load_library()
    cs = _capi.make_config_struct(Qwen3TTSTokenizerV2DecoderConfig())
    h = C.c_void_p()
    assert lib.q3tts_create(C.byref(cs), 0, C.byref(h)) == _capi.CUDA_ERROR
    assert not h.value
    from qwen3_tts_b200 import Qwen3TTSTokenizerV2Decoder
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Qwen3TTSTokenizerV2Decoder(Qwen3TTSTokenizerV2DecoderConfig(), device="cuda:0")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        Qwen3TTSTokenizerV2Decoder(Qwen3TTSTokenizerV2DecoderConfig(), device="cpu")


def test_config_defaults_mirror_reference():
    c = Qwen3TTSTokenizerV2DecoderConfig()
    assert (c.codebook_size, c.hidden_size, c.latent_dim, c.num_hidden_layers, c.sliding_window) == (2048, 1024, 1024, 8, 72)
    assert c.upsample_rates == (8, 5, 4, 3) and c.upsampling_ratios == (2, 2) and c.total_upsample == 1920
    assert c.head_dim == 64 and c.layer_types == ["sliding_attention"] * 8
    o = Qwen3TTSTokenizerV2Config(decoder_config={"codebook_dim": 256})
    assert o.decoder_config.codebook_dim == 256 and o.model_type == "qwen3_tts_tokenizer_12hz"
    assert (o.input_sample_rate, o.output_sample_rate, o.decode_upsample_rate, o.encode_downsample_rate) == (24000, 24000, 1920, 1920)


def test_param_inventory_matches_reference_count(sd, cfg):
    shapes = decoder_param_shapes(cfg)
    assert set(shapes) == set(sd)
    n = sum(int(np.prod(s)) for k, s in shapes.items())
    # 195 080 897 parameters in the decoder
    assert n == 195_080_897


def test_synthetic_inputs_are_deterministic_and_dodge_the_padding_rule(cfg):
    a = make_codes(3, 20, cfg, seed=5, lengths=[20, 7, 0])
    b = make_codes(3, 20, cfg, seed=5, lengths=[20, 7, 0])
    assert torch.equal(a, b) and a.dtype == torch.int64 and a.shape == (3, 20, 16)
    assert (a[0, :, 0] > 0).all() and (a[1, :7, 0] > 0).all() and (a[1, 7:] == 0).all() and (a[2] == 0).all()
    assert int(a.max()) < cfg.codebook_size


def test_tokenizer_rejects_bad_payload_like_reference():
    from qwen3_tts_b200.tokenizer import Qwen3TTSTokenizer

    class _Stub:
        def get_model_type(self):
            return "qwen3_tts_tokenizer_12hz"

    t = Qwen3TTSTokenizer()
    t.model = _Stub()
    with pytest.raises(TypeError, match="must be an encode output, a dict, or a list of dicts"):
        t.decode(12345)
    t.model.get_model_type = lambda: "something_else"
    with pytest.raises(ValueError, match="Unknown model type"):
        t.decode({"audio_codes": torch.zeros(4, 16, dtype=torch.long)})


def test_compute_mode_table_matches_header():
    """_capi.MODES (what compute_mode= accepts) must be exactly the Q3TTS_MODE_* enum of the header."""
    hdr = open(os.path.join(ROOT, "include", "q3tts_decode.h")).read()
    enum = {m.group(1).lower().replace("_", "-"): int(m.group(2)) for m in re.finditer(r"Q3TTS_MODE_([A-Z0-9_]+)\s*=\s*(\d+)", hdr)}
    assert enum == _capi.MODES and len(enum) == 6
    assert all((v & 1) == (1 if k.startswith("bf16") else 0) for k, v in enum.items())     # bit 0 = operand type


def test_encoder_key_filter_and_audio_input_rules():
    """Host logic of the encode side that needs no GPU: which `encoder.*` checkpoint keys are uploaded (the reference reads 16 of
    Mimi's 32 codebooks, modeling :982; decoder-side parts of MimiModel are nulled, :898-907) and the wrapper's input rules
    (wrapper :162-206)."""
    import types

    import numpy as np

    from qwen3_tts_b200 import Qwen3TTSTokenizer, Qwen3TTSTokenizerV2Model
    from qwen3_tts_b200.synthetic import encoder_param_shapes, make_encoder_state_dict
    m = object.__new__(Qwen3TTSTokenizerV2Model)
    m.config = Qwen3TTSTokenizerV2Config()
    sd = make_encoder_state_dict(seed=0)
    assert set(sd) == set(encoder_param_shapes()) and all(m._wants_encoder_key(k) for k in sd)
    q = "encoder.quantizer.acoustic_residual_vector_quantizer."
    assert m._wants_encoder_key(q + "layers.14.codebook.embed_sum") and not m._wants_encoder_key(q + "layers.15.codebook.embed_sum")
    assert not m._wants_encoder_key(q + "output_proj.weight") and not m._wants_encoder_key(q + "layers.0.codebook.initialized")
    assert not m._wants_encoder_key("encoder.quantizer.semantic_residual_vector_quantizer.layers.1.codebook.embed_sum")
    assert not m._wants_encoder_key("encoder.decoder_transformer.layers.0.mlp.fc1.weight") and not m._wants_encoder_key("decoder.pre_conv.conv.weight")
    tok = object.__new__(Qwen3TTSTokenizer)
    tok.model = types.SimpleNamespace(get_input_sample_rate=lambda: 24000)
    a = np.zeros((100, 2), dtype=np.float64)
    out = tok._normalize_audio_inputs(a, sr=24000)                   # a single array is a batch of one; channels are averaged; float32
    assert len(out) == 1 and out[0].shape == (100,) and out[0].dtype == np.float32
    with pytest.raises(ValueError):
        tok._normalize_audio_inputs([np.zeros(10, dtype=np.float32)], sr=None)
    with pytest.raises(TypeError):
        tok._normalize_audio_inputs([np.zeros(10, dtype=np.float32), "a.wav"], sr=24000)
    with pytest.raises(NotImplementedError):
        tok._normalize_audio_inputs([np.zeros(10, dtype=np.float32)], sr=16000)   # resampling needs librosa: not shipped
    assert tok._normalize_audio_inputs([], sr=24000) == []
