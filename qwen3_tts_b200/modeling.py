"""Host-side mirror of the reference's Tokenizer-12Hz decoder classes over the CUDA C-ABI.

Same class names, argument meaning and error behaviour as
qwen_tts/core/tokenizer_12hz/modeling_qwen3_tts_tokenizer_v2.py ("modeling" below):
  Qwen3TTSTokenizerV2Decoder.forward / chunked_decode   (modeling :868, :885)
  Qwen3TTSTokenizerV2Model.decode + getters             (modeling :945-958, :992-1022)
PyTorch is used for device memory and streams only; all arithmetic runs in libq3tts_decode.so.
"""
from __future__ import annotations

import ctypes as C
import glob
import json
import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Union

import numpy as np
import torch

from . import _capi
from .configuration import Qwen3TTSTokenizerV2Config, Qwen3TTSTokenizerV2DecoderConfig


@dataclass
class Qwen3TTSTokenizerV2DecoderOutput:
    """audio_values: List[FloatTensor (segment_length_i,)] — reference modeling :63-72."""
    audio_values: List[torch.Tensor] = None

    def __getitem__(self, i):
        return (self.audio_values,)[i]

    def to_tuple(self):
        return (self.audio_values,)


@dataclass
class Qwen3TTSTokenizerV2EncoderOutput:
    """audio_codes: List[LongTensor (codes_length_i, num_quantizers)] - reference modeling :50-60."""
    audio_codes: List[torch.Tensor] = None

    def __getitem__(self, i):
        return (self.audio_codes,)[i]

    def to_tuple(self):
        return (self.audio_codes,)


def _stream_ptr(device: torch.device) -> int:
    return int(torch.cuda.current_stream(device).cuda_stream)


class _Quantizer:
    """`decoder.quantizer.decode(codes)` — SplitResidualVectorQuantizer.decode (modeling :814-820)."""

    def __init__(self, dec: "Qwen3TTSTokenizerV2Decoder"):
        self._dec = dec

    def encode(self, latents: torch.Tensor) -> torch.Tensor:
        """The encoder's quantiser stage (MimiSplitResidualVectorQuantizer.encode, the tail of modeling :960-990): latents
        [B, codebook_dim, T] fp32 -> codes [B, Q, T] int64.  Needs the model's `encoder.quantizer.*` tensors (load_state_dict)."""
        d = self._dec
        if latents.dim() != 3 or latents.shape[1] != d.config.codebook_dim:
            raise ValueError(f"Expected latents of shape (B, {d.config.codebook_dim}, T), got {tuple(latents.shape)}")
        x = latents.to(device=d.device, dtype=torch.float32).contiguous()
        B, _, T = x.shape
        codes = torch.empty((B, d.config.num_quantizers, T), dtype=torch.int64, device=d.device)
        rc = d._lib.q3tts_rvq_encode(d._h, x.data_ptr(), B, T, codes.data_ptr(), _stream_ptr(d.device))
        _capi.check(d._h, rc, "rvq_encode")
        return codes

    def decode(self, codes: torch.Tensor) -> torch.Tensor:
        d = self._dec
        codes = d._check_codes(codes)
        B, _, T = codes.shape
        out = torch.empty((B, d.config.codebook_dim, T), dtype=torch.float32, device=d.device)
        rc = d._lib.q3tts_rvq_decode(d._h, codes.data_ptr(), B, T, out.data_ptr(), _stream_ptr(d.device))
        _capi.check(d._h, rc, "rvq_decode")
        return out


class Qwen3TTSTokenizerV2Decoder:
    """Drop-in for the reference decoder module (modeling :823-895), inference only."""

    def __init__(self, config: Qwen3TTSTokenizerV2DecoderConfig, device: Union[str, torch.device] = "cuda:0",
                 compute_mode: str = "fp16"):
        self.config = config
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("Qwen3TTSTokenizerV2Decoder (B200-native) needs a CUDA device; there is no CPU fallback")
        if not torch.cuda.is_available():
            raise RuntimeError("CUDA is not available; the B200-native decoder has no CPU fallback")
        self.total_upsample = config.total_upsample
        self.compute_mode = _capi.MODES[compute_mode]
        self._lib = _capi.load_library()
        self._h = C.c_void_p()
        dev_index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", dev_index)
        cs = _capi.make_config_struct(config)
        rc = self._lib.q3tts_create(C.byref(cs), dev_index, C.byref(self._h))
        if rc != _capi.OK:
            raise RuntimeError(f"q3tts_create failed ({rc}): needs an sm_100 (B200) device and a valid config")
        self.quantizer = _Quantizer(self)
        self._ready = False

    # ------------------------------------------------------------------ weights
    def load_state_dict(self, state_dict: Dict[str, torch.Tensor], strict: bool = True):
        """Consumes reference `Qwen3TTSTokenizerV2Decoder.state_dict()` keys."""
        any_dev = False
        for key, t in state_dict.items():
            if key.endswith("inv_freq") or (key.startswith("quantizer.") and ".input_proj." in key):
                continue  # non-persistent rope buffer; RVQ input_proj is unused on decode (modeling :757-759)
            t = t.detach()
            if t.dtype != torch.float32:
                t = t.to(torch.float32)
            t = t.contiguous()
            on_dev = 1 if t.is_cuda else 0
            any_dev = any_dev or t.is_cuda
            if t.is_cuda:
                if t.device != self.device:
                    t = t.to(self.device)
                torch.cuda.current_stream(self.device).synchronize()
            shape = (C.c_int64 * t.dim())(*t.shape)
            rc = self._lib.q3tts_load_tensor(self._h, key.encode(), t.data_ptr(), t.dim(), shape, on_dev)
            _capi.check(self._h, rc, f"load_tensor({key})")
        if any_dev:
            torch.cuda.synchronize(self.device)
        rc = self._lib.q3tts_finalize_weights(self._h, self.compute_mode)
        _capi.check(self._h, rc, "finalize_weights")
        self._ready = True
        return self

    def reserve(self, max_frames_per_launch: int):
        _capi.check(self._h, self._lib.q3tts_reserve(self._h, int(max_frames_per_launch)), "reserve")

    def set_engine(self, engine: str):
        _capi.check(self._h, self._lib.q3tts_set_engine(self._h, {"tc": _capi.ENGINE_TC, "simt": _capi.ENGINE_SIMT}[engine]), "set_engine")

    def set_check_codes(self, on: bool):
        """Opt-in range check for device codes (forward / chunked_decode / decode_step / quantizer.decode): raises IndexError where
        the reference's F.embedding asserts (modeling :675-678).  Costs a stream synchronise per call, so it is off by default."""
        _capi.check(self._h, self._lib.q3tts_set_check_codes(self._h, 1 if on else 0), "set_check_codes")

    def activation_range(self, frames: int = 24, seed: int = 0) -> Dict[str, float]:
        """max |activation| at every oracle tap for one seeded random code sequence.  16-bit operands overflow at 65504 (fp16): a
        checkpoint whose Snake / ConvNeXt activations get near that must run in bf16 mode (from_pretrained checks this)."""
        g = torch.Generator().manual_seed(seed)
        codes = torch.randint(1, self.config.codebook_size, (1, self.config.num_quantizers, frames), generator=g, dtype=torch.int64)
        stages = ["quantizer", "pre_conv", "pre_transformer"]
        stages += [f"upsample.{i}.{j}" for i in range(len(self.config.upsampling_ratios)) for j in (0, 1)] + ["decoder.0.snake"]
        stages += [f"decoder.{i + 1}.block.{j}" for i in range(len(self.config.upsample_rates)) for j in (1, 2, 3, 4)]
        out = {}
        for st in stages:
            v = self.probe(codes, st)
            out[st] = float(v.abs().max().item()) if bool(torch.isfinite(v).all()) else float("inf")
        return out

    def launch_count(self) -> int:
        return int(self._lib.q3tts_launch_count(self._h))

    def set_profile(self, on: bool):
        _capi.check(self._h, self._lib.q3tts_set_profile(self._h, 1 if on else 0), "set_profile")

    def get_profile(self):
        """[(kernel code, ms, flops, bytes, (rows, n, cin, taps))] for every op launched since the last call (profiling must be on);
        reading clears the profile.  Kernel codes: include/q3tts_decode.h q3tts_get_profile."""
        n = 4096
        ty, ms = (C.c_int32 * n)(), (C.c_float * n)()
        fl, by = (C.c_double * n)(), (C.c_double * n)()
        dm = (C.c_int32 * (4 * n))()
        k = self._lib.q3tts_get_profile(self._h, n, ty, ms, fl, by, dm)
        if k < 0:
            _capi.check(self._h, k, "get_profile")
        return [(int(ty[i]), float(ms[i]), float(fl[i]), float(by[i]), tuple(dm[4 * i: 4 * i + 4])) for i in range(min(k, n))]

    # ------------------------------------------------------------------ forward paths
    def _check_codes(self, codes: torch.Tensor) -> torch.Tensor:
        if codes.dim() != 3 or codes.shape[1] != self.config.num_quantizers:
            got = codes.shape[1] if codes.dim() >= 2 else "?"
            raise ValueError(f"Expected {self.config.num_quantizers} layer of codes, got {got}")  # modeling :869-870
        if not self._ready:
            raise RuntimeError("decoder weights are not loaded")
        return codes.to(device=self.device, dtype=torch.int64).contiguous()

    def forward(self, codes: torch.Tensor) -> torch.Tensor:
        """codes [B, Q, T] int64 -> wav [B, 1, T*1920-555] fp32, clamped (modeling :868-883)."""
        codes = self._check_codes(codes)
        B, _, T = codes.shape
        n = int(self._lib.q3tts_forward_num_samples(self._h, T))
        wav = torch.empty((B, 1, n), dtype=torch.float32, device=self.device)
        rc = self._lib.q3tts_forward(self._h, codes.data_ptr(), B, T, wav.data_ptr(), _stream_ptr(self.device))
        _capi.check(self._h, rc, "forward")
        return wav

    __call__ = forward

    def chunked_decode(self, codes: torch.Tensor, chunk_size: int = 300, left_context_size: int = 25) -> torch.Tensor:
        """modeling :885-895."""
        codes = self._check_codes(codes)
        B, _, T = codes.shape
        n = int(self._lib.q3tts_chunked_num_samples(self._h, T, chunk_size, left_context_size))
        wav = torch.empty((B, 1, n), dtype=torch.float32, device=self.device)
        rc = self._lib.q3tts_chunked_decode(self._h, codes.data_ptr(), B, T, int(chunk_size), int(left_context_size), wav.data_ptr(),
                                            _stream_ptr(self.device))
        _capi.check(self._h, rc, "chunked_decode")
        return wav

    # ------------------------------------------------------------------ streaming (additive API, see include/q3tts_decode.h)
    def reserve_streams(self, max_streams: int, max_frames_per_step: int = 4):
        """Allocates per-stream state (conv history rows, K/V rings) and the step workspace."""
        _capi.check(self._h, self._lib.q3tts_stream_reserve(self._h, int(max_streams), int(max_frames_per_step)), "stream_reserve")

    def open_stream(self) -> int:
        sid = C.c_int(-1)
        _capi.check(self._h, self._lib.q3tts_stream_open(self._h, C.byref(sid), _stream_ptr(self.device)), "stream_open")
        return int(sid.value)

    def reset_stream(self, stream_id: int):
        _capi.check(self._h, self._lib.q3tts_stream_reset(self._h, int(stream_id), _stream_ptr(self.device)), "stream_reset")

    def close_stream(self, stream_id: int):
        _capi.check(self._h, self._lib.q3tts_stream_close(self._h, int(stream_id)), "stream_close")

    def stream_frames(self, stream_id: int) -> int:
        return int(self._lib.q3tts_stream_frames(self._h, int(stream_id)))

    def stream_state_bytes(self) -> int:
        return int(self._lib.q3tts_stream_state_bytes(self._h))

    def decode_step(self, stream_ids, codes: torch.Tensor, out: Optional[torch.Tensor] = None):
        """One packet for each of S streams: codes [S, Q, P] int64 -> (wav [S, P*1920] fp32 on the device, n_valid list).

        Concatenating wav[i, :n_valid[i]] over the steps of a stream equals forward(all its codes) (the first step of a
        stream emits P*1920-555 samples, later steps P*1920)."""
        codes = self._check_codes(codes)
        S, _, P = codes.shape
        ids = (C.c_int32 * S)(*[int(i) for i in stream_ids])
        if len(stream_ids) != S:
            raise ValueError(f"{len(stream_ids)} stream ids for {S} code rows")
        stride = P * self.total_upsample
        if out is None:
            out = torch.empty((S, stride), dtype=torch.float32, device=self.device)
        elif out.shape != (S, stride) or out.dtype != torch.float32 or not out.is_contiguous():
            raise ValueError("out must be a contiguous float32 [S, P*upsample] tensor")
        nv = (C.c_int32 * S)()
        rc = self._lib.q3tts_stream_step(self._h, ids, S, codes.data_ptr(), P, out.data_ptr(), stride, nv, _stream_ptr(self.device))
        _capi.check(self._h, rc, "stream_step")
        return out, [int(v) for v in nv]

    def probe(self, codes: torch.Tensor, stage: str) -> torch.Tensor:
        """Parity probe: activation after `stage` as fp32 [B, C, L] (oracle tap names)."""
        codes = self._check_codes(codes)
        B, _, T = codes.shape
        cap = B * T * self.total_upsample * 96 + 1024
        cap = max(cap, B * T * 4 * 4096)
        buf = torch.empty(cap, dtype=torch.float32, device=self.device)
        c, l = C.c_int(0), C.c_int(0)
        rc = self._lib.q3tts_forward_probe(self._h, codes.data_ptr(), B, T, stage.encode(), buf.data_ptr(), cap, C.byref(c), C.byref(l),
                                           _stream_ptr(self.device))
        _capi.check(self._h, rc, f"forward_probe({stage})")
        return buf[: B * c.value * l.value].view(B, c.value, l.value)

    def __del__(self):
        try:
            if getattr(self, "_h", None) and self._h.value:
                self._lib.q3tts_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass


class Qwen3TTSTokenizerV2Model:
    """Drop-in for the reference model shell (modeling :927-1022): decode(), encode() and the five getters.

    encode() (wav -> codes; the reference delegates to HF MimiModel, modeling :898-907, :960-990) runs on the same handle when the
    state dict carries the `encoder.*` tensors; SURVEY 8(f)-2."""

    def __init__(self, config: Qwen3TTSTokenizerV2Config, device: Union[str, torch.device] = "cuda:0", compute_mode: str = "fp16"):
        self.config = config
        self.encoder_valid_num_quantizers = config.encoder_valid_num_quantizers
        self.input_sample_rate = config.input_sample_rate
        self.output_sample_rate = config.output_sample_rate
        self.decode_upsample_rate = config.decode_upsample_rate
        self.encode_downsample_rate = config.encode_downsample_rate
        self.decoder = Qwen3TTSTokenizerV2Decoder(config.decoder_config, device=device, compute_mode=compute_mode)
        self._apply_encoder_config(config.encoder_config)
        self.device = self.decoder.device
        self.dtype = torch.float32

    def _apply_encoder_config(self, enc: Optional[dict]):
        """Hands the encoder's non-shape hyper-parameters (MimiConfig fields; defaults = MimiConfig()) to the handle before the weights
        are finalised.  Shapes (filters, ratios, layer count, widths) are read from the tensors themselves."""
        enc = dict(enc or {})
        rope = enc.get("rope_parameters") or {}
        s = _capi.q3tts_encoder_config(
            num_attention_heads=int(enc.get("num_attention_heads", 8)), sliding_window=int(enc.get("sliding_window", 250)),
            max_position_embeddings=int(enc.get("max_position_embeddings", 8000)),
            rope_theta=float(rope.get("rope_theta", enc.get("rope_theta", 10000.0))), norm_eps=float(enc.get("norm_eps", 1e-5)))
        d = self.decoder
        _capi.check(d._h, d._lib.q3tts_set_encoder_config(d._h, C.byref(s)), "set_encoder_config")

    # ---- getters (modeling :945-958)
    def get_model_type(self):
        return self.config.model_type

    def get_input_sample_rate(self):
        return self.input_sample_rate

    def get_output_sample_rate(self):
        return self.output_sample_rate

    def get_encode_downsample_rate(self):
        return self.encode_downsample_rate

    def get_decode_upsample_rate(self):
        return self.decode_upsample_rate

    def load_state_dict(self, state_dict: Dict[str, torch.Tensor], strict: bool = True):
        """Accepts the reference model's state_dict: `decoder.*` keys feed the decode path, `encoder.*` keys the encode path
        (conv stack, transformer, down-sampling conv, and the first `num_quantizers` codebooks of the quantiser: the reference keeps
        encoder_valid_num_quantizers = 16 of Mimi's 32, modeling :982).  A state dict without `encoder.*` gives a decode-only model."""
        dec = {k[len("decoder."):]: v for k, v in state_dict.items() if k.startswith("decoder.")}
        if not dec:  # already decoder-level keys
            dec = {k: v for k, v in state_dict.items() if not k.startswith("encoder.")}
        dec.update({k: v for k, v in state_dict.items() if self._wants_encoder_key(k)})
        self.decoder.load_state_dict(dec, strict=strict)
        self._has_encoder = any(k.startswith("encoder.encoder.") for k in state_dict)
        return self

    def _wants_encoder_key(self, k: str) -> bool:
        if not k.startswith("encoder."):
            return False
        if k.startswith("encoder.quantizer."):
            if not (k.endswith("input_proj.weight") or k.endswith("embed_sum") or k.endswith("cluster_usage")):
                return False
            if ".layers." in k:   # acoustic codebooks beyond the ones the reference reads are not uploaded
                layer = int(k.split(".layers.")[1].split(".")[0])
                limit = 1 if "semantic_" in k else self.config.decoder_config.num_quantizers - 1
                return layer < limit
            return True
        return k.startswith(("encoder.encoder.", "encoder.encoder_transformer.", "encoder.downsample.")) and not k.endswith("inv_freq")

    @classmethod
    def from_pretrained(cls, path: str, device_map: Union[str, torch.device, None] = None, dtype=None, **kwargs):
        """Reads `<path>/config.json` and `*.safetensors` like AutoModel.from_pretrained would (wrapper :88).
        dtype=torch.bfloat16 selects bf16 tensor-core operands; anything else (float32, float16, None) asks for the fp16 parity mode
        (fp16 operands, fp32 accumulation).  fp16 operands overflow at 65504 and parity has only been measured on synthetic weights
        (no checkpoint is mounted on the build boxes), so after loading the activation range of every stage is probed once on random
        codes: if any stage comes within 2x of the fp16 limit the model is rebuilt in bf16 mode and a warning says so."""
        import warnings

        from safetensors import safe_open

        config = Qwen3TTSTokenizerV2Config.from_pretrained(path)
        device = device_map if device_map not in (None, "auto") else "cuda:0"
        mode = "bf16" if dtype is torch.bfloat16 else "fp16"
        model = cls(config, device=device, compute_mode=mode)
        sd: Dict[str, torch.Tensor] = {}
        files = sorted(glob.glob(os.path.join(path, "*.safetensors")))
        if not files:
            raise FileNotFoundError(f"no .safetensors under {path}")
        for f in files:
            with safe_open(f, framework="pt", device="cpu") as sf:
                for k in sf.keys():
                    if k.startswith("decoder.") or model._wants_encoder_key(k):
                        sd[k] = sf.get_tensor(k)
        model.load_state_dict(sd)
        if mode == "fp16":
            rng = model.decoder.activation_range()
            worst = max(rng, key=rng.get)
            if not rng[worst] < 0.5 * 65504.0:
                warnings.warn(f"fp16 operands would overflow on this checkpoint (max |activation| {rng[worst]:.3g} at {worst}); "
                              "loading in bf16 mode instead (wider range, 8-bit mantissa)")
                del model
                model = cls(config, device=device, compute_mode="bf16")
                model.load_state_dict(sd)
        return model

    def encode_num_frames(self, samples: int) -> int:
        """Code frames the encoder emits for `samples` input samples (every strided conv rounds up, modeling_mimi.py:273-285)."""
        d = self.decoder
        n = int(d._lib.q3tts_encode_num_frames(d._h, int(samples)))
        _capi.check(d._h, min(n, 0), "encode_num_frames")
        return n

    def encode(self, input_values: torch.Tensor, padding_mask: Optional[torch.Tensor] = None, return_dict: Optional[bool] = None,
               return_latents: bool = False):
        """input_values [B, T] float waveform (zero-padded), padding_mask [B, T] (1 = real sample) ->
        Qwen3TTSTokenizerV2EncoderOutput(audio_codes=List[LongTensor (ceil(mask_i.sum() / 1920), 16)]) (modeling :960-990).
        return_latents (additive): also return the quantiser's input [B, 512, F] for parity checks."""
        return_dict = return_dict if return_dict is not None else self.config.return_dict
        if input_values.dim() != 2:
            raise ValueError(f"Expected input_values of shape (B, T), got {tuple(input_values.shape)}")
        d = self.decoder
        wav = input_values.to(device=d.device, dtype=torch.float32).contiguous()
        B, T = wav.shape
        F = self.encode_num_frames(T)
        Q = d.config.num_quantizers
        codes = torch.empty((B, Q, F), dtype=torch.int64, device=d.device)
        lat = torch.empty((B, d.config.codebook_dim, F), dtype=torch.float32, device=d.device) if return_latents else None
        rc = d._lib.q3tts_encode(d._h, wav.data_ptr(), B, T, T, codes.data_ptr(), lat.data_ptr() if return_latents else None, _stream_ptr(d.device))
        _capi.check(d._h, rc, "encode")
        n = [T] * B if padding_mask is None else padding_mask.sum(dim=-1).tolist()   # no mask: every sample is real
        rate = self.encode_downsample_rate
        audio_codes = [codes[b, :self.encoder_valid_num_quantizers, :-(-int(n[b]) // rate)].transpose(0, 1) for b in range(B)]
        if return_latents:
            return audio_codes, lat
        if not return_dict:
            return (audio_codes,)
        return Qwen3TTSTokenizerV2EncoderOutput(audio_codes)

    def encode_host(self, wavs: List[np.ndarray]) -> List[np.ndarray]:
        """Host form (additive): list of 1-D float32 waveforms at 24 kHz -> list of int64 arrays [frames_i, 16] in host memory, through
        q3tts_encode_host (one pinned staging buffer each way; the layout q3tts_decode_host reads)."""
        d = self.decoder
        B = len(wavs)
        T = max(int(w.shape[0]) for w in wavs)
        F = self.encode_num_frames(T)
        Q = d.config.num_quantizers
        stage = getattr(self, "_enc_stage", None)
        if stage is None or stage[0].numel() < B * T or stage[1].numel() < B * F * Q:
            stage = (torch.empty(B * T, dtype=torch.float32).pin_memory(), torch.empty(B * F * Q, dtype=torch.int64).pin_memory())
            self._enc_stage = stage
        wav = stage[0][:B * T].view(B, T)
        valid = np.empty(B, dtype=np.int64)
        for b, w in enumerate(wavs):
            n = int(w.shape[0])
            wav[b, :n] = torch.from_numpy(np.ascontiguousarray(w, dtype=np.float32))
            wav[b, n:] = 0.0
            valid[b] = n
        codes = stage[1][:B * F * Q].view(B, F, Q)
        lengths = np.empty(B, dtype=np.int64)
        rc = d._lib.q3tts_encode_host(d._h, wav.data_ptr(), B, T, T, valid.ctypes.data, codes.data_ptr(), lengths.ctypes.data, _stream_ptr(d.device))
        _capi.check(d._h, rc, "encode_host")
        return [codes[b, :int(lengths[b]), :self.encoder_valid_num_quantizers].numpy().copy() for b in range(B)]

    def decode(self, audio_codes: torch.Tensor, return_dict: Optional[bool] = None):
        """audio_codes [B, T, Q] -> Qwen3TTSTokenizerV2DecoderOutput(audio_values=List[Tensor]) (modeling :992-1022)."""
        return_dict = return_dict if return_dict is not None else self.config.return_dict
        audio_values = self.decoder.chunked_decode(audio_codes.transpose(1, 2)).squeeze(1)
        audio_lengths = (audio_codes[..., 0] > 0).sum(1) * self.decode_upsample_rate
        audio_values = [a[:l] for a, l in zip(audio_values, audio_lengths.tolist())]
        if not return_dict:
            return (audio_values,)
        return Qwen3TTSTokenizerV2DecoderOutput(audio_values)

    def decode_host(self, audio_codes_cpu: torch.Tensor, skip_samples=None):
        """End-to-end path used by Qwen3TTSTokenizer.decode: HOST codes [B,T,Q] -> list of HOST fp32
        waveforms, with the H2D / D2H copies and the length rule inside one C call.  skip_samples (one int per item,
        optional): leading samples that are never stored (voice-clone prefix cut, model wrapper :627-630)."""
        import numpy as np
        codes = audio_codes_cpu.to(torch.int64).contiguous()
        if codes.dim() != 3 or codes.shape[2] != self.config.decoder_config.num_quantizers:
            raise ValueError(f"Expected {self.config.decoder_config.num_quantizers} layer of codes, got {codes.shape[-1]}")
        B, T, _ = codes.shape
        d = self.decoder
        total = int(d._lib.q3tts_chunked_num_samples(d._h, T, 300, 25))
        wav = torch.empty((B, max(total, 1)), dtype=torch.float32, pin_memory=True)
        lengths = (C.c_int64 * B)()
        skip = self._skip_array(skip_samples, B)
        rc = d._lib.q3tts_decode_host(d._h, codes.data_ptr(), B, T, wav.data_ptr(), wav.shape[1], lengths, skip, _stream_ptr(d.device))
        _capi.check(d._h, rc, "decode_host")
        arr = wav.numpy()  # fresh pinned buffer per call: the returned rows are views that keep it alive, no extra host copy
        return [arr[b, : lengths[b]] for b in range(B)]

    @staticmethod
    def _skip_array(skip_samples, B: int):
        if skip_samples is None:
            return None
        if len(skip_samples) != B:
            raise ValueError(f"{len(skip_samples)} skip_samples for {B} items")
        return (C.c_int64 * B)(*[int(v) for v in skip_samples])

    def decode_host_pcm16(self, audio_codes_cpu: torch.Tensor, skip_samples=None):
        """decode_host with the wire formats of SURVEY 8f-3: codes travel as int16, the result is PCM16
        (np.int16 = rint(clamp(wav) * 32767), what sf.write stores for a float waveform), converted in the output conv."""
        if audio_codes_cpu.dim() != 3 or audio_codes_cpu.shape[2] != self.config.decoder_config.num_quantizers:
            raise ValueError(f"Expected {self.config.decoder_config.num_quantizers} layer of codes, got {audio_codes_cpu.shape[-1]}")
        if audio_codes_cpu.numel() and (int(audio_codes_cpu.min()) < 0 or int(audio_codes_cpu.max()) >= self.config.decoder_config.codebook_size):
            # checked BEFORE narrowing: .to(int16) would wrap a leaked special / EOS id into a valid-looking index
            raise IndexError(f"code index outside [0, {self.config.decoder_config.codebook_size}) (reference: F.embedding raises, modeling :675-678)")
        codes = audio_codes_cpu.to(torch.int16).contiguous()
        B, T, _ = codes.shape
        d = self.decoder
        total = int(d._lib.q3tts_chunked_num_samples(d._h, T, 300, 25))
        pcm = torch.empty((B, max(total, 1)), dtype=torch.int16, pin_memory=True)
        lengths = (C.c_int64 * B)()
        skip = self._skip_array(skip_samples, B)
        rc = d._lib.q3tts_decode_host_pcm16(d._h, codes.data_ptr(), B, T, pcm.data_ptr(), pcm.shape[1], lengths, skip, _stream_ptr(d.device))
        _capi.check(d._h, rc, "decode_host_pcm16")
        arr = pcm.numpy()
        return [arr[b, : lengths[b]] for b in range(B)]
