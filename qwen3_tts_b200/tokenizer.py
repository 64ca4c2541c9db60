"""`Qwen3TTSTokenizer` — drop-in for qwen_tts/inference/qwen3_tts_tokenizer.py ("wrapper"), decode side.

Same constructor-less class, `from_pretrained`, `decode(encoded) -> (List[np.float32 1-D], sr)` accepting
the reference's payload forms (wrapper :298-329), `encode(audios, sr)` for numpy waveforms (wrapper :208-257) and the five getters
(wrapper :367-411).  Audio file I/O and resampling (librosa / soundfile, wrapper :122-206) are not part of this package: encode takes
waveforms that are already at the model's input rate.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np
import torch
from torch.nn.utils.rnn import pad_sequence

from .configuration import Qwen3TTSTokenizerV2Config
from .modeling import Qwen3TTSTokenizerV2Model


class Qwen3TTSTokenizer:
    def __init__(self):
        self.model = None
        self.feature_extractor = None
        self.config = None
        self.device = None

    @classmethod
    def from_pretrained(cls, pretrained_model_name_or_path: str, **kwargs) -> "Qwen3TTSTokenizer":
        """kwargs as the reference forwards to AutoModel.from_pretrained: device_map, dtype (wrapper :64-97)."""
        inst = cls()
        inst.model = Qwen3TTSTokenizerV2Model.from_pretrained(
            pretrained_model_name_or_path, device_map=kwargs.get("device_map"), dtype=kwargs.get("dtype", kwargs.get("torch_dtype")))
        inst.config = inst.model.config
        inst.device = inst.model.device
        return inst

    @classmethod
    def from_model(cls, model: Qwen3TTSTokenizerV2Model) -> "Qwen3TTSTokenizer":
        """Wrap an already-built model (synthetic weights, tests, benches)."""
        inst = cls()
        inst.model = model
        inst.config = model.config
        inst.device = model.device
        return inst

    def _normalize_audio_inputs(self, audios, sr) -> List[np.ndarray]:
        """wrapper :162-206 for numpy input: list of 1-D float32 waveforms at the model's input rate.  Paths / base64 strings need
        soundfile + librosa (file decoding, resampling), which this package does not ship: load them with the reference's
        `load_audio` and pass the arrays."""
        target_sr = int(self.model.get_input_sample_rate())
        if isinstance(audios, (str, np.ndarray)):
            audios = [audios]
        if len(audios) == 0:
            return []
        if isinstance(audios[0], str):
            raise NotImplementedError("audio paths / base64 strings need soundfile + librosa; pass numpy waveforms and `sr`")
        if sr is None:
            raise ValueError("For numpy waveform input, you must provide `sr` (original sampling rate).")
        out = []
        for a in audios:
            if not isinstance(a, np.ndarray):
                raise TypeError("Mixed input types are not supported. Use all paths/base64 or all numpy arrays.")
            if a.ndim > 1:
                a = np.mean(a, axis=-1)
            if int(sr) != target_sr:
                raise NotImplementedError(f"resampling {sr} -> {target_sr} Hz needs librosa; resample before calling encode")
            out.append(a.astype(np.float32))
        return out

    def encode(self, audios, sr=None, return_dict: bool = True):
        """Batch-encode audio into discrete codes (wrapper :208-257): numpy waveform(s) + `sr` -> Qwen3TTSTokenizerV2EncoderOutput with
        audio_codes = List[LongTensor (codes_len_i, 16)].  Padding follows the reference's feature extractor (EncodecFeatureExtractor,
        padding=True): zeros up to the longest item, padding_mask = 1 on real samples."""
        wavs = self._normalize_audio_inputs(audios, sr=sr)
        if not wavs:
            raise ValueError("encode: no audio given")
        T = max(int(w.shape[0]) for w in wavs)
        input_values = torch.zeros((len(wavs), T), dtype=torch.float32)
        padding_mask = torch.zeros((len(wavs), T), dtype=torch.long)
        for b, w in enumerate(wavs):
            input_values[b, :w.shape[0]] = torch.from_numpy(w)
            padding_mask[b, :w.shape[0]] = 1
        return self.model.encode(input_values.to(self.device), padding_mask, return_dict=return_dict)

    def num_decoded_samples(self, frames: int) -> int:
        """Samples decode() returns for an item that fills a batch padded to `frames` code frames:
        frames * 1920 - 555 per 300-frame chunk (chunked_decode, modeling :885-895)."""
        d = self.model.decoder
        return int(d._lib.q3tts_chunked_num_samples(d._h, int(frames), 300, 25))

    def decode(self, encoded, skip_samples=None) -> Tuple[List[np.ndarray], int]:
        """Decode back to waveform (wrapper :259-365).  `skip_samples` (additive, one int per item): leading samples of each
        waveform that are cut inside the decode call (voice-clone prefix, model wrapper :627-630)."""
        model_type = self.model.get_model_type()

        def _to_tensor(x, dtype=None):
            if isinstance(x, torch.Tensor):
                return x
            t = torch.from_numpy(np.asarray(x))
            return t.to(dtype) if dtype is not None else t

        if hasattr(encoded, "audio_codes"):
            audio_codes_list = encoded.audio_codes
        elif isinstance(encoded, dict):
            audio_codes_list = encoded["audio_codes"]
        elif isinstance(encoded, list):
            audio_codes_list = [e["audio_codes"] for e in encoded]
        else:
            raise TypeError("`encoded` must be an encode output, a dict, or a list of dicts.")

        if isinstance(audio_codes_list, torch.Tensor):
            t = audio_codes_list
            if t.dim() == 2:  # single sample (T, Q) -> (1, T, Q)
                t = t.unsqueeze(0)
            audio_codes_padded = t
        else:
            audio_codes_list = [_to_tensor(c, dtype=torch.long) for c in audio_codes_list]
            audio_codes_padded = pad_sequence(audio_codes_list, batch_first=True, padding_value=0)

        if model_type != "qwen3_tts_tokenizer_12hz":
            raise ValueError(f"Unknown model type: {model_type}")

        with torch.inference_mode():
            if audio_codes_padded.is_cuda and skip_samples is None:
                dec = self.model.decode(audio_codes_padded, return_dict=True)
                wavs = [w.to(torch.float32).detach().cpu().numpy() for w in dec.audio_values]
            else:
                # host payload: one C call does H2D, chunked decode, the length rule, the prefix cut and D2H
                wavs = self.model.decode_host(audio_codes_padded.cpu(), skip_samples=skip_samples)
        return wavs, int(self.model.get_output_sample_rate())

    def decode_pcm16(self, encoded, skip_samples=None) -> Tuple[List[np.ndarray], int]:
        """Additive: same payload forms as decode(), result as PCM16 (np.int16) ready for a WAV writer."""
        if hasattr(encoded, "audio_codes"):
            lst = encoded.audio_codes
        elif isinstance(encoded, dict):
            lst = encoded["audio_codes"]
        elif isinstance(encoded, list):
            lst = [e["audio_codes"] for e in encoded]
        else:
            raise TypeError("`encoded` must be an encode output, a dict, or a list of dicts.")
        if isinstance(lst, torch.Tensor):
            padded = lst.unsqueeze(0) if lst.dim() == 2 else lst
        else:
            padded = pad_sequence([c if isinstance(c, torch.Tensor) else torch.from_numpy(np.asarray(c)) for c in lst], batch_first=True,
                                  padding_value=0)
        with torch.inference_mode():
            pcm = self.model.decode_host_pcm16(padded.cpu(), skip_samples=skip_samples)
        return pcm, int(self.model.get_output_sample_rate())

    def get_model_type(self) -> str:
        return self.model.get_model_type()

    def get_input_sample_rate(self) -> int:
        return int(self.model.get_input_sample_rate())

    def get_output_sample_rate(self) -> int:
        return int(self.model.get_output_sample_rate())

    def get_encode_downsample_rate(self) -> int:
        return int(self.model.get_encode_downsample_rate())

    def get_decode_upsample_rate(self) -> int:
        return int(self.model.get_decode_upsample_rate())
