"""B200-native Qwen3-TTS-Tokenizer-12Hz decode path (codes -> 24 kHz waveform).

Public surface mirrors the reference package for this path (qwen_tts/__init__.py:21-22,
qwen_tts/core/tokenizer_12hz): Qwen3TTSTokenizer, Qwen3TTSTokenizerV2Model,
Qwen3TTSTokenizerV2Decoder and their configs.  All arithmetic runs in the in-tree CUDA
library lib/libq3tts_decode.so (hand-written sm_100a kernels behind a C-ABI)."""
from .configuration import Qwen3TTSTokenizerV2Config, Qwen3TTSTokenizerV2DecoderConfig
from .modeling import (Qwen3TTSTokenizerV2Decoder, Qwen3TTSTokenizerV2DecoderOutput, Qwen3TTSTokenizerV2EncoderOutput,
                       Qwen3TTSTokenizerV2Model)
from .inference import Qwen3TTSModel, TalkerStreamSink
from .sharding import ShardedDecoder
from .tokenizer import Qwen3TTSTokenizer


def synthetic_model(seed: int = 0, device="cuda:0", compute_mode: str = "fp16", state_dict=None, stress: bool = False, with_encoder: bool = False):
    """Model with the seeded random-init weights of the named architecture (no checkpoint is mounted).  with_encoder: also load
    seeded `encoder.*` weights (Mimi shapes) so that encode() works."""
    from .synthetic import make_decoder_state_dict, make_encoder_state_dict
    cfg = Qwen3TTSTokenizerV2Config()
    model = Qwen3TTSTokenizerV2Model(cfg, device=device, compute_mode=compute_mode)
    sd = state_dict if state_dict is not None else make_decoder_state_dict(cfg.decoder_config, seed, stress=stress)
    if with_encoder:
        sd = dict(sd)
        sd.update(make_encoder_state_dict(seed))
    model.decoder.load_state_dict(sd)
    return model


__all__ = ["ShardedDecoder", "Qwen3TTSModel", "TalkerStreamSink", "Qwen3TTSTokenizer", "Qwen3TTSTokenizerV2Model", "Qwen3TTSTokenizerV2Decoder", "Qwen3TTSTokenizerV2DecoderOutput", "Qwen3TTSTokenizerV2EncoderOutput",
           "Qwen3TTSTokenizerV2Config", "Qwen3TTSTokenizerV2DecoderConfig", "synthetic_model"]
